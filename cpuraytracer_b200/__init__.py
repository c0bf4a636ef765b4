"""cpuraytracer_b200 -- B200-native (sm_100a) implementation of the SakibSaikia/CPURayTracer
`src/spheres` per-pixel Monte Carlo render loop behind the reference's scene / camera / render API.

The compute path is csrc/librt_b200.so (hand-written CUDA, C ABI in include/rt_b200.h); this package
is the thin host-side mirror of the reference interface.  There is no CPU fallback.
"""
from ._lib import (LIB_PATH, RT_FLAG_DEBUG_NO_CULL, RT_FLAG_JITTER_OFF, RT_FLAG_LENS_OFF, RT_FLAG_WAVEFRONT, RT_FLAG_ACCEL, RT_FLAG_HALTON, RT_FLAG_OVERWRITE, RT_FLAG_DEBUG_ONE_KERNEL, RT_MAT_METAL, RT_MAT_OPAQUE,
                   RT_MAT_SKY, RT_MAT_TRANSPARENT, RtError, load)
from .api import (Camera, CheckerTexture, ConstTexture, DielectricOpaque, DielectricTransparent, DirectionalLight,
                  Emissive, FlatScene, Metal, Renderer, Scene, Sphere, XMCOLOR)
from . import scenes
from .imageio import write_pfm, write_ppm

__all__ = [n for n in dir() if not n.startswith("_")]
