"""ctypes binding of oracle/liborc.so (rt_oracle.h) -- CPU ORACLE, TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import
this module.  The product package (cpuraytracer_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liborc.so")

SPHERE_DTYPE = np.dtype([("cx", "<f4"), ("cy", "<f4"), ("cz", "<f4"), ("radius", "<f4"), ("material", "<i4")])
MATERIAL_DTYPE = np.dtype([("kind", "<i4"), ("texture", "<i4"), ("smoothness", "<f4"), ("ior", "<f4")])
TEXTURE_DTYPE = np.dtype([("kind", "<i4"), ("tiling", "<f4"), ("c0", "<f4", (4,)), ("c1", "<f4", (4,))])
LIGHT_DTYPE = np.dtype([("dir", "<f4", (3,)), ("color", "<f4", (4,)), ("luminance", "<f4")])

FLAG_JITTER_OFF, FLAG_LENS_OFF = 1, 2
FLAG_REF_DRAWBITMAP = 4      # tier A only: serial GenerateRays + parallel trace + parallel resolve (the reference's frame structure)
RNG_COUNTER, RNG_HALTON = 0, 1
ACCEL_BRUTE, ACCEL_BVH = 0, 1
INTEGRATOR_ITERATIVE, INTEGRATOR_RECURSIVE = 0, 1
MATH_PINNED, MATH_REFERENCE = 0, 1
MAT_OPAQUE, MAT_METAL, MAT_TRANSPARENT, MAT_SKY = 0, 1, 2, 3


class OrcScene(C.Structure):
    _fields_ = [("n_spheres", C.c_int32), ("spheres", C.c_void_p),
                ("n_materials", C.c_int32), ("materials", C.c_void_p),
                ("n_textures", C.c_int32), ("textures", C.c_void_p),
                ("n_lights", C.c_int32), ("lights", C.c_void_p),
                ("sky_luminance", C.c_float), ("sky_texture", C.c_int32)]


class OrcCameraParams(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("look_at", C.c_float * 3), ("vfov_deg", C.c_float),
                ("aspect", C.c_float), ("focus", C.c_float), ("aperture", C.c_float)]


class OrcCamera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("x", C.c_float * 3), ("y", C.c_float * 3),
                ("image_origin", C.c_float * 3), ("aperture", C.c_float), ("focus", C.c_float)]


class OrcConfig(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("first_sample", C.c_uint64), ("spp", C.c_int32),
                ("max_depth", C.c_int32), ("flags", C.c_uint32), ("rng_mode", C.c_int32), ("accel", C.c_int32),
                ("integrator", C.c_int32), ("threads", C.c_int32), ("exposure_stops", C.c_float), ("seed", C.c_uint32), ("math_mode", C.c_int32)]


class OrcStats(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("closest_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("seconds", C.c_double)]


_lib = None


def build(force: bool = False) -> str:
    """Compiles liborc.so (and oracle/_ref when /root/reference is present) with the committed Makefile."""
    src = [os.path.join(HERE, f) for f in ("rt_oracle.cpp", "rt_oracle.h", "Makefile")]
    stale = (not os.path.exists(LIB_PATH)) or any(os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in src)
    if force or stale:
        r = subprocess.run(["make", "-C", HERE, "liborc.so"] + (["-B"] if force else []), capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)
    return LIB_PATH


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        build()
    lib = C.CDLL(LIB_PATH)
    F, P = C.c_float, C.c_void_p
    FP = C.POINTER(C.c_float)
    lib.orc_render.restype = C.c_int
    lib.orc_render.argtypes = [C.POINTER(OrcScene), C.POINTER(OrcCameraParams), C.POINTER(OrcConfig), P, C.POINTER(OrcStats)]
    lib.orc_primary_hits.restype = C.c_int
    lib.orc_primary_hits.argtypes = [C.POINTER(OrcScene), C.POINTER(OrcCameraParams), C.POINTER(OrcConfig), P, P]
    lib.orc_resolve_ldr.restype = None
    lib.orc_resolve_ldr.argtypes = [P, C.c_int64, C.c_uint64, P]
    lib.orc_set_math_mode.restype = None
    lib.orc_set_math_mode.argtypes = [C.c_int]
    lib.orc_halton.restype = F
    lib.orc_halton.argtypes = [C.c_uint64, C.c_uint32]
    lib.orc_halton_disk.restype = None
    lib.orc_halton_disk.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, FP]
    lib.orc_halton_hemisphere.restype = None
    lib.orc_halton_hemisphere.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, FP]
    lib.orc_hemisphere_from_u.restype = None
    lib.orc_hemisphere_from_u.argtypes = [F, F, FP]
    lib.orc_sincos.restype = None
    lib.orc_sincos.argtypes = [F, FP, FP]
    lib.orc_camera_make.restype = None
    lib.orc_camera_make.argtypes = [C.POINTER(OrcCameraParams), C.POINTER(OrcCamera)]
    lib.orc_camera_get_ray.restype = None
    lib.orc_camera_get_ray.argtypes = [C.POINTER(OrcCamera), F, F, F, F, FP, FP]
    lib.orc_sphere_intersect.restype = C.c_int
    lib.orc_sphere_intersect.argtypes = [P, FP, FP, FP, FP, FP, FP]
    lib.orc_closest_hit.restype = C.c_int
    lib.orc_closest_hit.argtypes = [C.POINTER(OrcScene), FP, FP, C.c_int, FP]
    lib.orc_texture_eval.restype = None
    lib.orc_texture_eval.argtypes = [P, F, F, FP]
    lib.orc_schlick.restype = F
    lib.orc_schlick.argtypes = [F, F]
    lib.orc_fresnel_term.restype = F
    lib.orc_fresnel_term.argtypes = [F, F]
    lib.orc_reflect.restype = None
    lib.orc_reflect.argtypes = [FP, FP, FP]
    lib.orc_refract.restype = None
    lib.orc_refract.argtypes = [FP, FP, F, FP]
    lib.orc_xmcolor_quantize.restype = C.c_uint8
    lib.orc_xmcolor_quantize.argtypes = [F]
    lib.orc_xmcolor_load.restype = F
    lib.orc_xmcolor_load.argtypes = [C.c_uint8]
    lib.orc_rng_counter.restype = None
    lib.orc_rng_counter.argtypes = [C.c_uint32, C.c_uint64, C.c_uint32, C.c_uint32, FP]
    lib.orc_scatter.restype = C.c_int
    lib.orc_scatter.argtypes = [C.POINTER(OrcScene), C.c_int, FP, FP, F, F, F, FP, FP, FP]
    lib.orc_shade_hit.restype = C.c_int
    lib.orc_shade_hit.argtypes = [C.POINTER(OrcScene), FP, FP, FP, C.c_int, FP]
    lib.orc_build_info.restype = C.c_char_p
    lib.orc_build_info.argtypes = []
    _lib = lib
    return lib


def f3(*v):
    return (C.c_float * len(v))(*[float(x) for x in v])


class Scene:
    """Holds the flat arrays alive and exposes the orc_scene struct."""

    def __init__(self, spheres, materials, textures, lights, sky_luminance, sky_texture):
        self.spheres = np.ascontiguousarray(spheres, SPHERE_DTYPE)
        self.materials = np.ascontiguousarray(materials, MATERIAL_DTYPE)
        self.textures = np.ascontiguousarray(textures, TEXTURE_DTYPE)
        self.lights = np.ascontiguousarray(lights, LIGHT_DTYPE)
        self.c = OrcScene(len(self.spheres), self.spheres.ctypes.data, len(self.materials), self.materials.ctypes.data,
                          len(self.textures), self.textures.ctypes.data, len(self.lights), self.lights.ctypes.data,
                          float(sky_luminance), int(sky_texture))

    @staticmethod
    def from_flat(flat) -> "Scene":
        """flat: any object with .spheres/.materials/.textures/.lights/.sky_luminance/.sky_texture (same byte layout)."""
        return Scene(flat.spheres, flat.materials, flat.textures, flat.lights, flat.sky_luminance, flat.sky_texture)


    @staticmethod
    def from_rtsc(path: str) -> "Scene":
        """Reads a flat scene file "RTSC" v1 (apps/spheres --dump-scene; byte layout of rt_scene_export) with numpy only --
        the reference arm of bench.py takes its input from the committed oracle/scenes/*.rtsc and never maps the product library."""
        b = open(path, "rb").read()
        if len(b) < 32 or b[:4] != b"RTSC":
            raise ValueError(f"{path}: not an RTSC scene file")
        ver, ns, nm, nt, nl = np.frombuffer(b, "<i4", 5, 4)
        if ver != 1 or min(ns, nm, nt, nl) < 0:
            raise ValueError(f"{path}: unsupported RTSC version/counts")
        sky_lum = float(np.frombuffer(b, "<f4", 1, 24)[0])
        sky_tex = int(np.frombuffer(b, "<i4", 1, 28)[0])
        p, out = 32, []
        for dt, n in ((SPHERE_DTYPE, ns), (MATERIAL_DTYPE, nm), (TEXTURE_DTYPE, nt), (LIGHT_DTYPE, nl)):
            if p + n * dt.itemsize > len(b):
                raise ValueError(f"{path}: truncated RTSC file")
            out.append(np.frombuffer(b, dt, n, p).copy())
            p += n * dt.itemsize
        return Scene(out[0], out[1], out[2], out[3], sky_lum, sky_tex)

    def to_rtsc(self, path: str) -> None:
        hdr = b"RTSC" + np.array([1, len(self.spheres), len(self.materials), len(self.textures), len(self.lights)], "<i4").tobytes()
        hdr += np.array([self.c.sky_luminance], "<f4").tobytes() + np.array([self.c.sky_texture], "<i4").tobytes()
        with open(path, "wb") as f:
            f.write(hdr + self.spheres.tobytes() + self.materials.tobytes() + self.textures.tobytes() + self.lights.tobytes())


SCENES_DIR = os.path.join(HERE, "scenes")


def frozen_scene(name: str) -> Scene:
    """oracle/scenes/<name>.rtsc: the bench / golden scenes frozen as files (oracle/scenes/freeze_scenes.py wrote them from
    cpuraytracer_b200.scenes; tests/test_host_api.py checks they still agree byte for byte)."""
    return Scene.from_rtsc(os.path.join(SCENES_DIR, name + ".rtsc"))


def camera_params(origin, look_at, vfov_deg, aspect, focus, aperture) -> OrcCameraParams:
    return OrcCameraParams(f3(*origin), f3(*look_at), vfov_deg, aspect, focus, aperture)


def default_camera_params(aspect: float, aperture: float = 0.4) -> OrcCameraParams:
    """SpheresApp::InitCamera (spheres-app.cpp:35-49): origin (12, 2, -2.5), look-at (0, 1, 0), vfov 25 (spheres-app.h:10),
    focus = |origin - lookAt| in float32 (XMVector3Length order), aperture 0.4 (spheres-app.h:11)."""
    f = np.float32
    d = np.array([12.0, 2.0, -2.5], f) - np.array([0.0, 1.0, 0.0], f)
    focus = np.sqrt(f(f(d[0] * d[0] + d[1] * d[1]) + d[2] * d[2]))
    return camera_params((12.0, 2.0, -2.5), (0.0, 1.0, 0.0), 25.0, float(f(aspect)), float(focus), aperture)


def config(width, height, spp=1, max_depth=50, first_sample=1, flags=0, rng_mode=RNG_COUNTER, accel=ACCEL_BRUTE,
           integrator=INTEGRATOR_ITERATIVE, threads=0, exposure_stops=-15.0, seed=0, math_mode=0) -> OrcConfig:
    return OrcConfig(width, height, first_sample, spp, max_depth, flags, rng_mode, accel, integrator, threads,
                     exposure_stops, seed, math_mode)


def render(scene: Scene, cam: OrcCameraParams, cfg: OrcConfig, hdr: np.ndarray | None = None):
    """Returns (hdr running sum (H, W, 4) float32, OrcStats)."""
    if hdr is None:
        hdr = np.zeros((cfg.height, cfg.width, 4), np.float32)
    assert hdr.dtype == np.float32 and hdr.flags.c_contiguous and hdr.size == 4 * cfg.width * cfg.height
    st = OrcStats()
    rc = load().orc_render(C.byref(scene.c), C.byref(cam), C.byref(cfg), hdr.ctypes.data, C.byref(st))
    if rc != 0:
        raise ValueError(f"orc_render: bad argument ({rc})")
    return hdr, st


def primary_hits(scene: Scene, cam: OrcCameraParams, cfg: OrcConfig):
    ids = np.empty((cfg.height, cfg.width), np.int32)
    mids = np.empty((cfg.height, cfg.width), np.int32)
    rc = load().orc_primary_hits(C.byref(scene.c), C.byref(cam), C.byref(cfg), ids.ctypes.data, mids.ctypes.data)
    if rc != 0:
        raise ValueError(f"orc_primary_hits: bad argument ({rc})")
    return ids, mids


def resolve_ldr(hdr: np.ndarray, sample_count: int) -> np.ndarray:
    hdr = np.ascontiguousarray(hdr, np.float32)
    h, w = hdr.shape[:2]
    out = np.empty((h, w), np.uint32)
    load().orc_resolve_ldr(hdr.ctypes.data, h * w, sample_count, out.ctypes.data)
    return out


# ------------------------------------------------------------------------------------------------ tier A: oracle/_ref
REF_LIB_PATH = os.path.join(HERE, "_ref", "libref.so")
_ref = None


def load_ref():
    """oracle/_ref/libref.so = the reference's own common-lib .cpp files compiled against dxm_shim/ + ref_driver.cpp.
    Built by `make -C oracle` where /root/reference exists; travels to the GPU box as a prebuilt file.  None if absent."""
    global _ref
    if _ref is not None:
        return _ref
    if not os.path.exists(REF_LIB_PATH):
        if os.path.isdir("/root/reference/src/common-lib"):
            subprocess.run(["make", "-C", HERE, "ref"], capture_output=True, text=True)
        if not os.path.exists(REF_LIB_PATH):
            return None
    lib = C.CDLL(REF_LIB_PATH)
    F, P = C.c_float, C.c_void_p
    FP = C.POINTER(C.c_float)
    lib.ref_render.restype = C.c_int
    lib.ref_render.argtypes = [C.POINTER(OrcScene), C.POINTER(OrcCameraParams), C.POINTER(OrcConfig), P, C.POINTER(OrcStats)]
    lib.ref_primary_hits.restype = C.c_int
    lib.ref_primary_hits.argtypes = [C.POINTER(OrcScene), C.POINTER(OrcCameraParams), C.POINTER(OrcConfig), P, P]
    lib.ref_resolve_ldr.restype = None
    lib.ref_resolve_ldr.argtypes = [P, C.c_int64, C.c_uint64, P]
    lib.ref_halton.restype = F
    lib.ref_halton.argtypes = [C.c_uint64, C.c_uint32]
    lib.ref_halton_disk.restype = None
    lib.ref_halton_disk.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, FP]
    lib.ref_halton_hemisphere.restype = None
    lib.ref_halton_hemisphere.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, FP]
    lib.ref_camera_get_ray.restype = None
    lib.ref_camera_get_ray.argtypes = [C.POINTER(OrcCameraParams), F, F, F, F, FP, FP]
    lib.ref_sphere_intersect.restype = C.c_int
    lib.ref_sphere_intersect.argtypes = [P, FP, FP, FP, FP, FP, FP]
    lib.ref_closest_hit.restype = C.c_int
    lib.ref_closest_hit.argtypes = [C.POINTER(OrcScene), FP, FP, FP]
    lib.ref_texture_eval.restype = None
    lib.ref_texture_eval.argtypes = [P, F, F, FP]
    lib.ref_fresnel_term.restype = F
    lib.ref_fresnel_term.argtypes = [F, F]
    lib.ref_reflect.restype = None
    lib.ref_reflect.argtypes = [FP, FP, FP]
    lib.ref_refract.restype = None
    lib.ref_refract.argtypes = [FP, FP, F, FP]
    lib.ref_xmcolor_quantize.restype = C.c_uint8
    lib.ref_xmcolor_quantize.argtypes = [F]
    lib.ref_xmcolor_load.restype = F
    lib.ref_xmcolor_load.argtypes = [C.c_uint8]
    lib.ref_scatter_sequence.restype = C.c_int
    lib.ref_scatter_sequence.argtypes = [C.POINTER(OrcScene), C.c_int, C.c_int, P, P, P, P, P, P]
    lib.ref_shade_hit.restype = C.c_int
    lib.ref_shade_hit.argtypes = [C.POINTER(OrcScene), FP, FP, FP, FP]
    lib.ref_build_info.restype = C.c_char_p
    lib.ref_build_info.argtypes = []
    lib.ref_world_create.restype = P
    lib.ref_world_create.argtypes = [C.POINTER(OrcScene), C.POINTER(OrcCameraParams), C.c_int]
    lib.ref_world_destroy.restype = None
    lib.ref_world_destroy.argtypes = [P]
    lib.ref_world_threads.restype = C.c_int
    lib.ref_world_threads.argtypes = [P]
    lib.ref_world_render.restype = C.c_int
    lib.ref_world_render.argtypes = [P, C.POINTER(OrcConfig), P, C.POINTER(OrcStats)]
    _ref = lib
    return lib


class RefWorld:
    """SpheresApp after InitScene / InitCamera on the reference's own classes: scene, BVH, per-material Halton counters and
    the worker pool persist across render() calls, as they do between the reference's DrawBitmap calls."""

    def __init__(self, scene: Scene, cam: OrcCameraParams, threads: int = 0, lib=None):
        self.lib = lib if lib is not None else load_ref()
        if self.lib is None:
            raise RuntimeError("oracle/_ref/libref.so is not built (needs /root/reference)")
        self._scene = scene
        self.h = self.lib.ref_world_create(C.byref(scene.c), C.byref(cam), threads)
        if not self.h:
            raise ValueError("ref_world_create: bad argument")
        self.threads = self.lib.ref_world_threads(self.h)

    def render(self, cfg: OrcConfig, hdr: np.ndarray | None = None):
        if hdr is None:
            hdr = np.zeros((cfg.height, cfg.width, 4), np.float32)
        assert hdr.dtype == np.float32 and hdr.flags.c_contiguous and hdr.size == 4 * cfg.width * cfg.height
        st = OrcStats()
        rc = self.lib.ref_world_render(self.h, C.byref(cfg), hdr.ctypes.data, C.byref(st))
        if rc != 0:
            raise ValueError(f"ref_world_render: bad argument ({rc})")
        return hdr, st

    def close(self):
        if self.h:
            self.lib.ref_world_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def ref_render(scene: Scene, cam: OrcCameraParams, cfg: OrcConfig, hdr: np.ndarray | None = None):
    """The reference's own classes (atomic Halton counters + BVH + recursion).  Returns (hdr sum, OrcStats)."""
    lib = load_ref()
    if lib is None:
        raise RuntimeError("oracle/_ref/libref.so is not built (needs /root/reference)")
    if hdr is None:
        hdr = np.zeros((cfg.height, cfg.width, 4), np.float32)
    st = OrcStats()
    rc = lib.ref_render(C.byref(scene.c), C.byref(cam), C.byref(cfg), hdr.ctypes.data, C.byref(st))
    if rc != 0:
        raise ValueError(f"ref_render: bad argument ({rc})")
    return hdr, st


def ref_primary_hits(scene: Scene, cam: OrcCameraParams, cfg: OrcConfig):
    lib = load_ref()
    ids = np.empty((cfg.height, cfg.width), np.int32)
    mids = np.empty((cfg.height, cfg.width), np.int32)
    rc = lib.ref_primary_hits(C.byref(scene.c), C.byref(cam), C.byref(cfg), ids.ctypes.data, mids.ctypes.data)
    if rc != 0:
        raise ValueError(f"ref_primary_hits: bad argument ({rc})")
    return ids, mids
