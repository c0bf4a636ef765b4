"""ctypes binding of include/rt_b200.h (csrc/librt_b200.so).

There is no fallback: if the CUDA library is missing or cannot be loaded, importing the compute API
raises.  PyTorch is not required by this module.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# RT_B200_LIB: developer override used by tools/variants.py to A/B experimental builds of the same source
LIB_PATH = os.environ.get("RT_B200_LIB") or os.path.join(HERE, "csrc", "librt_b200.so")

RT_OK = 0
STATUS_NAMES = {0: "RT_OK", 1: "RT_ERR_INVALID_ARGUMENT", 2: "RT_ERR_CUDA", 3: "RT_ERR_OUT_OF_MEMORY",
                4: "RT_ERR_NOT_READY", 5: "RT_ERR_UNSUPPORTED"}
RT_MAT_OPAQUE, RT_MAT_METAL, RT_MAT_TRANSPARENT, RT_MAT_SKY = 0, 1, 2, 3
RT_TEX_CONST, RT_TEX_CHECKER = 0, 1
RT_FLAG_JITTER_OFF, RT_FLAG_LENS_OFF, RT_FLAG_WAVEFRONT, RT_FLAG_ACCEL, RT_FLAG_HALTON, RT_FLAG_DEBUG_NO_CULL = 1, 2, 4, 8, 16, 0x100
RT_FLAG_OVERWRITE, RT_FLAG_DEBUG_ONE_KERNEL = 32, 0x200
RT_STATIC_SCRATCH_BYTES, RT_STATIC_BATCH = 64 << 20, 16

# flat scene description v1 (byte-identical to the structs in include/rt_b200.h)
SPHERE_DTYPE = np.dtype([("cx", "<f4"), ("cy", "<f4"), ("cz", "<f4"), ("radius", "<f4"), ("material", "<i4")])
MATERIAL_DTYPE = np.dtype([("kind", "<i4"), ("texture", "<i4"), ("smoothness", "<f4"), ("ior", "<f4")])
TEXTURE_DTYPE = np.dtype([("kind", "<i4"), ("tiling", "<f4"), ("c0", "<f4", (4,)), ("c1", "<f4", (4,))])
LIGHT_DTYPE = np.dtype([("dir", "<f4", (3,)), ("color", "<f4", (4,)), ("luminance", "<f4")])
assert SPHERE_DTYPE.itemsize == 20 and MATERIAL_DTYPE.itemsize == 16 and TEXTURE_DTYPE.itemsize == 40 and LIGHT_DTYPE.itemsize == 32


class CameraParams(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("look_at", C.c_float * 3), ("vfov_deg", C.c_float),
                ("aspect", C.c_float), ("focus", C.c_float), ("aperture", C.c_float)]


class CameraBasis(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("x", C.c_float * 3), ("y", C.c_float * 3),
                ("image_origin", C.c_float * 3), ("aperture", C.c_float), ("focus", C.c_float)]


class AccelImageInfo(C.Structure):
    _fields_ = [("bytes", C.c_size_t), ("off_B", C.c_int32), ("off_R", C.c_int32), ("off_I", C.c_int32), ("off_box", C.c_int32),
                ("off_top", C.c_int32), ("n_pairs", C.c_int32), ("n_big_groups", C.c_int32), ("n_big_pairs", C.c_int32),
                ("n_groups", C.c_int32), ("group_slots", C.c_int32), ("n_l2_nodes", C.c_int32), ("node_floats", C.c_int32),
                ("bound_center", C.c_float * 3), ("bound_radius", C.c_float), ("r_min", C.c_float), ("r_max", C.c_float),
                ("c_max", C.c_float)]


class CullImageInfo(C.Structure):
    _fields_ = [("bytes", C.c_size_t), ("n_spheres", C.c_int32), ("n_pairs", C.c_int32), ("off_B", C.c_int32), ("off_G", C.c_int32),
                ("rot", C.c_float * 9), ("light_dir", C.c_float * 3)]


class Stats(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("closest_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("sphere_tests", C.c_uint64), ("passes", C.c_uint64), ("render_ms", C.c_double),
                ("resolve_ms", C.c_double), ("sm_count", C.c_int32), ("clock_khz", C.c_int32),
                ("launches", C.c_int32), ("reserved", C.c_int32)]


class RtError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"{STATUS_NAMES.get(code, code)}: {message}")
        self.code = code


# every symbol include/rt_b200.h declares: (name, restype, argtypes)
_P = C.c_void_p
_F = C.POINTER(C.c_float)
_I32 = C.POINTER(C.c_int32)
SYMBOLS = [
    ("rt_context_create", C.c_int, [C.c_int, C.POINTER(_P)]),
    ("rt_context_destroy", None, [_P]),
    ("rt_last_error", C.c_char_p, []),
    ("rt_set_stream", C.c_int, [_P, _P]),
    ("rt_get_stream", _P, [_P]),
    ("rt_synchronize", C.c_int, [_P]),
    ("rt_scene_create", C.c_int, [C.POINTER(_P)]),
    ("rt_scene_destroy", None, [_P]),
    ("rt_scene_add_texture_const", C.c_int, [_P, _F]),
    ("rt_scene_add_texture_checker", C.c_int, [_P, _F, _F, C.c_float]),
    ("rt_scene_add_material_opaque", C.c_int, [_P, C.c_int, C.c_float]),
    ("rt_scene_add_material_metal", C.c_int, [_P, C.c_int, C.c_float]),
    ("rt_scene_add_material_transparent", C.c_int, [_P, C.c_float, C.c_float]),
    ("rt_scene_add_sphere", C.c_int, [_P, _F, C.c_float, C.c_int]),
    ("rt_scene_set_sky", C.c_int, [_P, C.c_float, C.c_int]),
    ("rt_scene_add_directional_light", C.c_int, [_P, _F, _F, C.c_float]),
    ("rt_scene_counts", C.c_int, [_P, _I32, _I32, _I32, _I32]),
    ("rt_scene_export", C.c_int, [_P, _P, _P, _P, _P, _F, _I32]),
    ("rt_scene_import", C.c_int, [_P, C.c_int32, _P, C.c_int32, _P, C.c_int32, _P, C.c_int32, _P, C.c_float, C.c_int32]),
    ("rt_set_scene", C.c_int, [_P, _P]),
    ("rt_set_camera", C.c_int, [_P, C.POINTER(CameraParams)]),
    ("rt_camera_basis_of", C.c_int, [C.POINTER(CameraParams), C.POINTER(CameraBasis)]),
    ("rt_set_exposure", C.c_int, [_P, C.c_float]),
    ("rt_set_seed", C.c_int, [_P, C.c_uint32]),
    ("rt_sampler_reset", C.c_int, [_P]),
    ("rt_set_chunk_spp", C.c_int, [_P, C.c_int]),
    ("rt_set_accel_group_slots", C.c_int, [_P, C.c_int]),
    ("rt_framebuffer_resize", C.c_int, [_P, C.c_int, C.c_int]),
    ("rt_framebuffer_clear", C.c_int, [_P]),
    ("rt_render", C.c_int, [_P, C.c_uint64, C.c_int, C.c_int, C.c_uint32]),
    ("rt_render_static", C.c_int, [_P, C.c_uint64, C.c_int, C.c_int, C.c_uint32]),
    ("rt_render_static_scratch_bytes", C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    ("rt_read_hdr_sum", C.c_int, [_P, _P, C.c_size_t]),
    ("rt_write_hdr_sum", C.c_int, [_P, _P, C.c_size_t]),
    ("rt_resolve_ldr", C.c_int, [_P, C.c_uint64, _P, C.c_size_t]),
    ("rt_primary_hits", C.c_int, [_P, C.c_uint64, C.c_uint32, _P, _P, C.c_size_t]),
    ("rt_hdr_device_ptr", _P, [_P]),
    ("rt_hdr_bind_external", C.c_int, [_P, _P]),
    ("rt_scene_accel_image", C.c_int, [_P, _P, C.c_size_t, C.POINTER(AccelImageInfo)]),
    ("rt_scene_accel_image_slots", C.c_int, [_P, C.c_int, _P, C.c_size_t, C.POINTER(AccelImageInfo)]),
    ("rt_scene_cull_image", C.c_int, [_P, C.c_int, _P, C.c_size_t, C.POINTER(CullImageInfo)]),
    ("rt_get_stats", C.c_int, [_P, C.POINTER(Stats)]),
    ("rt_stats_reset", C.c_int, [_P]),
    ("rt_build_info", C.c_char_p, []),
]

_lib = None


def load() -> C.CDLL:
    """Loads csrc/librt_b200.so.  Raises (never falls back) when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run `python -m cpuraytracer_b200.build` (needs nvcc). "
                          "cpuraytracer_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, res, args in SYMBOLS:
        fn = getattr(lib, name)          # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(code: int) -> None:
    if code != RT_OK:
        raise RtError(code, load().rt_last_error().decode("utf-8", "replace"))


def check_id(value: int) -> int:
    if value < 0:
        raise RtError(-value, load().rt_last_error().decode("utf-8", "replace"))
    return value
