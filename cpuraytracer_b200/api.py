"""Host-side mirror of the reference's `src/common-lib` + `src/spheres` interface over the C ABI.

Class and argument names follow the reference (camera.h:9, ray-tracing.h:59, material.h:24,40,57,73,
texture.h:15,25, light.h:15); the objects only *record* into an rt_scene -- all rendering happens in
the CUDA library.  `Renderer.render(spp, max_depth)` == `spp` x SpheresApp::DrawBitmap
(spheres-app.cpp:163-222) with AppSettings::k_recursionDepth = max_depth.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional, Sequence

import numpy as np

from . import _lib as L


def _f(values: Sequence[float], n: int):
    if len(values) != n:
        raise ValueError(f"expected {n} floats, got {len(values)}")
    return (C.c_float * n)(*[float(v) for v in values])


# ---------------------------------------------------------------- textures / materials / hitables
@dataclass
class XMCOLOR:
    """DirectX::PackedVector::XMCOLOR(float r, g, b, a): quantised to 8 bit by the library."""
    r: float
    g: float
    b: float
    a: float = 1.0

    def tuple(self):
        return (self.r, self.g, self.b, self.a)


class Texture:
    pass


@dataclass
class ConstTexture(Texture):        # texture.h:12-20
    color: XMCOLOR


@dataclass
class CheckerTexture(Texture):      # texture.h:22-30
    color0: XMCOLOR
    color1: XMCOLOR
    tiling: float


class Material:
    pass


@dataclass
class Metal(Material):              # material.h:21-35
    reflectance: Texture
    smoothness: float = 0.0


@dataclass
class DielectricOpaque(Material):   # material.h:37-52
    albedo: Texture
    smoothness: float = 16.0


@dataclass
class DielectricTransparent(Material):   # material.h:54-68
    smoothness: float = 16.0
    ior: float = 1.5


@dataclass
class Emissive(Material):           # material.h:70-84 (sky only)
    luminance: float
    color: Texture


@dataclass
class Sphere:                       # ray-tracing.h:53-65
    center: Sequence[float]
    radius: float
    material: Material


@dataclass
class DirectionalLight:             # light.h:12-22 (the occlusion callback is the scene itself)
    direction: Sequence[float]
    color: XMCOLOR
    luminance: float


@dataclass
class Camera:                       # camera.h:6-20
    origin: Sequence[float]
    look_at: Sequence[float]
    vertical_fov: float
    aspect_ratio: float
    focal_length: float
    aperture: float

    def params(self) -> L.CameraParams:
        return L.CameraParams(_f(self.origin, 3), _f(self.look_at, 3), self.vertical_fov, self.aspect_ratio,
                              self.focal_length, self.aperture)

    def basis(self) -> L.CameraBasis:
        out = L.CameraBasis()
        p = self.params()
        L.check(L.load().rt_camera_basis_of(C.byref(p), C.byref(out)))
        return out

    def GetOrigin(self):
        return tuple(self.origin)


# ---------------------------------------------------------------- scene recorder
class Scene:
    """m_scene / m_textures / m_lights / m_skyMaterial of SpheresApp (spheres-app.h:35-41)."""

    def __init__(self):
        self._lib = L.load()
        h = C.c_void_p()
        L.check(self._lib.rt_scene_create(C.byref(h)))
        self._h = h
        self._tex_ids = {}
        self._mat_ids = {}

    def close(self):
        if getattr(self, "_h", None):
            self._lib.rt_scene_destroy(self._h)
            self._h = None

    __del__ = close

    @property
    def handle(self):
        return self._h

    def accel_image(self, min_group_slots: int = 0):
        """(info, image bytes) of the box hierarchy RT_FLAG_ACCEL stages for this scene (rt_scene_accel_image_slots: host only,
        no device needed); layout in include/rt_b200.h.  min_group_slots: finest group size to try (8, 16, 32; 0 = default).
        Raises RtError(RT_ERR_UNSUPPORTED) when the scene has no hierarchy."""
        info = L.AccelImageInfo()
        L.check(self._lib.rt_scene_accel_image_slots(self._h, min_group_slots, None, 0, C.byref(info)))
        buf = np.zeros(info.bytes, dtype=np.uint8)
        L.check(self._lib.rt_scene_accel_image_slots(self._h, min_group_slots, buf.ctypes.data_as(C.c_void_p), buf.nbytes, C.byref(info)))
        return info, buf

    def cull_image(self, light_index: int = 0):
        """(info, A (np, 4), B (np, 4), G (2 np, 4)) float32: the arrays the brute-force query stages for light `light_index`
        (-1: no light) -- pair arrays in the light-aligned frame + world-space (c, r); rt_scene_cull_image, host only."""
        info = L.CullImageInfo()
        L.check(self._lib.rt_scene_cull_image(self._h, light_index, None, 0, C.byref(info)))
        buf = np.zeros(info.bytes, dtype=np.uint8)
        L.check(self._lib.rt_scene_cull_image(self._h, light_index, buf.ctypes.data_as(C.c_void_p), buf.nbytes, C.byref(info)))
        f = buf.view(np.float32)
        return (info, f[: info.off_B // 4].reshape(-1, 4), f[info.off_B // 4: info.off_G // 4].reshape(-1, 4), f[info.off_G // 4:].reshape(-1, 4))

    # -- low-level recorder (ids) --
    def add_texture(self, tex: Texture) -> int:
        key = id(tex)
        if key in self._tex_ids:
            return self._tex_ids[key][0]
        if isinstance(tex, ConstTexture):
            i = L.check_id(self._lib.rt_scene_add_texture_const(self._h, _f(tex.color.tuple(), 4)))
        elif isinstance(tex, CheckerTexture):
            i = L.check_id(self._lib.rt_scene_add_texture_checker(self._h, _f(tex.color0.tuple(), 4),
                                                                  _f(tex.color1.tuple(), 4), tex.tiling))
        else:
            raise TypeError(f"unknown texture {tex!r}")
        self._tex_ids[key] = (i, tex)
        return i

    def add_material(self, mat: Material) -> int:
        """Each call creates a new material instance, as every reference Sphere owns its own (ray-tracing.h:57)."""
        if isinstance(mat, DielectricOpaque):
            return L.check_id(self._lib.rt_scene_add_material_opaque(self._h, self.add_texture(mat.albedo), mat.smoothness))
        if isinstance(mat, Metal):
            return L.check_id(self._lib.rt_scene_add_material_metal(self._h, self.add_texture(mat.reflectance), mat.smoothness))
        if isinstance(mat, DielectricTransparent):
            return L.check_id(self._lib.rt_scene_add_material_transparent(self._h, mat.smoothness, mat.ior))
        raise TypeError(f"{type(mat).__name__} cannot be attached to a Sphere")

    def add(self, sphere: Sphere) -> int:
        return L.check_id(self._lib.rt_scene_add_sphere(self._h, _f(sphere.center, 3), sphere.radius,
                                                        self.add_material(sphere.material)))

    def set_sky(self, sky: Emissive) -> None:
        L.check(self._lib.rt_scene_set_sky(self._h, sky.luminance, self.add_texture(sky.color)))

    def add_light(self, light: DirectionalLight) -> None:
        L.check(self._lib.rt_scene_add_directional_light(self._h, _f(light.direction, 3), _f(light.color.tuple(), 4),
                                                         light.luminance))

    # -- flat view --
    def counts(self):
        n = [C.c_int32() for _ in range(4)]
        L.check(self._lib.rt_scene_counts(self._h, *[C.byref(x) for x in n]))
        return tuple(x.value for x in n)

    def export(self) -> "FlatScene":
        ns, nm, nt, nl = self.counts()
        sp = np.zeros(ns, L.SPHERE_DTYPE)
        ma = np.zeros(nm, L.MATERIAL_DTYPE)
        te = np.zeros(nt, L.TEXTURE_DTYPE)
        li = np.zeros(nl, L.LIGHT_DTYPE)
        lum, tex = C.c_float(), C.c_int32()
        L.check(self._lib.rt_scene_export(self._h, sp.ctypes.data, ma.ctypes.data, te.ctypes.data, li.ctypes.data,
                                          C.byref(lum), C.byref(tex)))
        return FlatScene(sp, ma, te, li, lum.value, tex.value)

    @staticmethod
    def from_flat(flat: "FlatScene") -> "Scene":
        s = Scene()
        sp = np.ascontiguousarray(flat.spheres, L.SPHERE_DTYPE)
        ma = np.ascontiguousarray(flat.materials, L.MATERIAL_DTYPE)
        te = np.ascontiguousarray(flat.textures, L.TEXTURE_DTYPE)
        li = np.ascontiguousarray(flat.lights, L.LIGHT_DTYPE)
        L.check(s._lib.rt_scene_import(s._h, len(sp), sp.ctypes.data, len(ma), ma.ctypes.data, len(te), te.ctypes.data,
                                       len(li), li.ctypes.data, flat.sky_luminance, flat.sky_texture))
        return s


@dataclass
class FlatScene:
    """Flat scene description v1: the exact bytes both the CUDA library and an external checker consume."""
    spheres: np.ndarray
    materials: np.ndarray
    textures: np.ndarray
    lights: np.ndarray
    sky_luminance: float
    sky_texture: int

    def save(self, path: str) -> None:
        np.savez_compressed(path, spheres=self.spheres, materials=self.materials, textures=self.textures,
                            lights=self.lights, sky=np.array([self.sky_luminance, self.sky_texture], np.float64))

    @staticmethod
    def load(path: str) -> "FlatScene":
        z = np.load(path)
        return FlatScene(z["spheres"], z["materials"], z["textures"], z["lights"], float(z["sky"][0]), int(z["sky"][1]))


# ---------------------------------------------------------------- renderer (one GPU)
class Renderer:
    """One rt_context: the B200 replacement of SpheresApp's render loop.

    m_backbufferHdr / m_backbufferLdr (app.h:26-27) live in HBM; `render` adds samples, `read_hdr_sum`,
    `resolve_ldr` and `primary_hits` copy results to host numpy arrays.
    """

    def __init__(self, device: int = 0, width: int = 1280, height: int = 720, stream: Optional[int] = None):
        self._lib = L.load()
        h = C.c_void_p()
        L.check(self._lib.rt_context_create(device, C.byref(h)))
        self._h = h
        self.device = device
        self.width = self.height = 0
        self.sample_count = 0          # m_sampleCount, spheres-app.h:42
        if stream is not None:
            self.set_stream(stream)
        self.resize(width, height)

    def close(self):
        if getattr(self, "_h", None):
            self._lib.rt_context_destroy(self._h)
            self._h = None

    __del__ = close

    def set_stream(self, cuda_stream: int) -> None:
        L.check(self._lib.rt_set_stream(self._h, C.c_void_p(cuda_stream)))

    def resize(self, width: int, height: int) -> None:
        L.check(self._lib.rt_framebuffer_resize(self._h, width, height))
        self.width, self.height, self.sample_count = width, height, 0

    def clear(self) -> None:
        L.check(self._lib.rt_framebuffer_clear(self._h))
        self.sample_count = 0

    def set_scene(self, scene: Scene) -> None:
        L.check(self._lib.rt_set_scene(self._h, scene.handle))

    def set_camera(self, camera: Camera) -> None:
        p = camera.params()
        L.check(self._lib.rt_set_camera(self._h, C.byref(p)))

    def set_exposure(self, stops: float) -> None:
        L.check(self._lib.rt_set_exposure(self._h, stops))

    def set_seed(self, seed: int) -> None:
        L.check(self._lib.rt_set_seed(self._h, seed))

    def reset_sampler(self) -> None:
        """RT_FLAG_HALTON: every material's Halton indices back to 0 (as after InitScene)."""
        L.check(self._lib.rt_sampler_reset(self._h))

    def set_chunk_spp(self, chunk: int) -> None:
        L.check(self._lib.rt_set_chunk_spp(self._h, chunk))

    def set_accel_group_slots(self, slots: int) -> None:
        """RT_FLAG_ACCEL tuning: finest group size (8, 16 or 32 slots; 0 = default) the next set_scene may choose."""
        L.check(self._lib.rt_set_accel_group_slots(self._h, slots))

    def render(self, spp: int, max_depth: int = 50, flags: int = 0, first_sample: Optional[int] = None) -> None:
        """`spp` x DrawBitmap.  first_sample defaults to continuing after the samples already accumulated."""
        first = self.sample_count + 1 if first_sample is None else first_sample
        L.check(self._lib.rt_render(self._h, first, spp, max_depth, flags))
        self.sample_count += spp

    def render_static(self, spp: int, max_depth: int = 50, flags: int = 0, first_sample: Optional[int] = None) -> None:
        """rt_render_static: the same render with its scratch in the library's static device array -- allocates nothing
        (include/rt_b200.h has the limit and the sharing rule; no counterpart in the reference)."""
        first = self.sample_count + 1 if first_sample is None else first_sample
        L.check(self._lib.rt_render_static(self._h, first, spp, max_depth, flags))
        self.sample_count += spp

    def synchronize(self) -> None:
        L.check(self._lib.rt_synchronize(self._h))

    def read_hdr_sum(self, out: Optional[np.ndarray] = None) -> np.ndarray:
        n = self.width * self.height
        if out is None:
            out = np.empty((self.height, self.width, 4), np.float32)
        assert out.dtype == np.float32 and out.size == 4 * n and out.flags.c_contiguous
        L.check(self._lib.rt_read_hdr_sum(self._h, out.ctypes.data, n))
        return out

    def write_hdr_sum(self, hdr: np.ndarray, sample_count: int) -> None:
        hdr = np.ascontiguousarray(hdr, np.float32)
        L.check(self._lib.rt_write_hdr_sum(self._h, hdr.ctypes.data, self.width * self.height))
        self.sample_count = sample_count

    def resolve_ldr(self, sample_count: Optional[int] = None, out: Optional[np.ndarray] = None, copy: bool = True):
        n = self.width * self.height
        sc = self.sample_count if sample_count is None else sample_count
        if not copy:
            L.check(self._lib.rt_resolve_ldr(self._h, sc, None, 0))
            return None
        if out is None:
            out = np.empty((self.height, self.width), np.uint32)
        L.check(self._lib.rt_resolve_ldr(self._h, sc, out.ctypes.data, n))
        return out

    def primary_hits(self, sample: int = 1, flags: int = L.RT_FLAG_JITTER_OFF | L.RT_FLAG_LENS_OFF):
        n = self.width * self.height
        ids = np.empty((self.height, self.width), np.int32)
        mids = np.empty((self.height, self.width), np.int32)
        L.check(self._lib.rt_primary_hits(self._h, sample, flags, ids.ctypes.data, mids.ctypes.data, n))
        return ids, mids

    def hdr_device_ptr(self) -> int:
        return int(self._lib.rt_hdr_device_ptr(self._h) or 0)

    def bind_external_hdr(self, device_ptr: int) -> None:
        L.check(self._lib.rt_hdr_bind_external(self._h, C.c_void_p(device_ptr)))

    def stats(self) -> L.Stats:
        s = L.Stats()
        L.check(self._lib.rt_get_stats(self._h, C.byref(s)))
        return s

    def reset_stats(self) -> None:
        L.check(self._lib.rt_stats_reset(self._h))
