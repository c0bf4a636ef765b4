"""Builds the in-tree native artefacts: csrc/librt_b200.so (nvcc, sm_100a only).

The .so is git-ignored but travels to the GPU box with the gpurun snapshot.  nvcc cross-compiles
without a GPU, so this runs anywhere the CUDA toolkit is installed.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(CSRC, "librt_b200.so")
SOURCES = ["rt_b200.cu", "rt_kernels.cuh", "rt_device.cuh", "rt_bvh.cuh", "rt_wavefront.cuh"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-fmad=false",                       # pinned operation order: FMAs only where written (rt_device.cuh)
    "-prec-div=true", "-prec-sqrt=true",
    "-Xcompiler", "-fPIC,-ffp-contract=off,-fvisibility=hidden,-mfma",
    "-shared",
]


GROUP_LIB = os.path.join(CSRC, "librt_b200_group.so")
APPS = os.path.normpath(os.path.join(HERE, "..", "apps"))
CLI = os.path.join(APPS, "spheres")
INCLUDE = os.path.normpath(os.path.join(HERE, "..", "include"))
CUDA_HOME = os.environ.get("CUDA_HOME", "/usr/local/cuda")
HOST_FLAGS = ["-std=c++17", "-O2", "-Wall", "-ffp-contract=off", "-fPIC", f"-I{CUDA_HOME}/include"]
CUDART_STATIC = [f"-L{CUDA_HOME}/lib64", "-lcudart_static", "-ldl", "-lrt", "-lpthread"]


def _stale(target: str, deps: list[str]) -> bool:
    return not os.path.exists(target) or any(os.path.getmtime(d) > os.path.getmtime(target) for d in deps)


def build_host(force: bool = False) -> tuple[str, str]:
    """The C++17 host side above the C ABI: librt_b200_group.so (single-process multi-GPU sharding + ncclReduce,
    include/rt_b200_group.h) and apps/spheres (headless replacement of the reference executable).  Host code only (g++)."""
    cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
    hdrs = [os.path.join(INCLUDE, h) for h in ("rt_b200.h", "rt_b200.hpp", "rt_b200_group.h", "rt_image_io.hpp")]
    src = os.path.join(CSRC, "rt_group.cpp")
    if force or _stale(GROUP_LIB, [src, LIB, *hdrs]):
        cmd = [cxx, *HOST_FLAGS, "-fvisibility=hidden", "-shared", "-o", GROUP_LIB, src, f"-L{CSRC}", "-lrt_b200",
               *CUDART_STATIC, "-Wl,-rpath,$ORIGIN"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("g++ (rt_group) failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    main = os.path.join(APPS, "spheres_main.cpp")
    if force or _stale(CLI, [main, GROUP_LIB, LIB, *hdrs]):
        cmd = [cxx, *HOST_FLAGS, "-o", CLI, main, f"-L{CSRC}", "-lrt_b200_group", "-lrt_b200",
               "-Wl,-rpath,$ORIGIN/../cpuraytracer_b200/csrc"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("g++ (apps/spheres) failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    return GROUP_LIB, CLI


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; cannot build librt_b200.so")


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES] + [os.path.join(HERE, "..", "include", "rt_b200.h"), __file__]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, out: str | None = None, extra: list[str] | None = None) -> str:
    """`out` / `extra` build an experimental variant (tools/variants.py) beside the product library."""
    if out is None and not force and not needs_build():
        return LIB
    cmd = [_nvcc(), *NVCC_FLAGS, *(extra or []), "-o", out or LIB, os.path.join(CSRC, "rt_b200.cu")]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd), file=sys.stderr)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stdout + r.stderr, file=sys.stderr)
    return out or LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
    print(*build_host(force="--force" in sys.argv))
