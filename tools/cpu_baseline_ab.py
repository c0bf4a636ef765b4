#!/usr/bin/env python
"""A/B of the CPU reference driver (oracle/ref_driver.cpp) on this host's cores, workload c3 (d488, 1920x1080, depth 50):
  new   = per-thread ray counters, persistent world (what bench.py --impl reference runs),
  new/drawbitmap = same + the reference's frame structure (serial GenerateRays, parallel trace, parallel resolve),
  old   = round-1 instrumentation: two shared std::atomic counters bumped per ray (make -C oracle ref_ab).
Usage: python tools/cpu_baseline_ab.py [frames]   -> prints one line per variant and thread count."""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import orc  # noqa: E402

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 4
w, h, depth = 1920, 1080, 50
sc, cp = orc.frozen_scene("d488"), orc.default_camera_params(w / h)
new = orc.load_ref()
old_path = os.path.join(ROOT, "oracle", "_ref", "libref_r01counters.so")
libs = [("new (per-thread counters)", new, 0), ("new, DrawBitmap structure", new, orc.FLAG_REF_DRAWBITMAP)]
if os.path.exists(old_path):
    old = C.CDLL(old_path)
    for fn in ("ref_world_create", "ref_world_render", "ref_world_threads", "ref_world_destroy"):
        getattr(old, fn).restype, getattr(old, fn).argtypes = getattr(new, fn).restype, getattr(new, fn).argtypes
    libs.append(("old (shared atomic counters per ray)", old, 0))
ncpu = os.cpu_count() or 1
print(f"host cpus {ncpu}; {frames} timed frames of {w}x{h} depth {depth} after 1 untimed frame")
for threads in sorted({ncpu, max(1, ncpu // 2), min(ncpu, 16)}):
    for name, lib, flags in libs:
        world = orc.RefWorld(sc, cp, threads, lib=lib)
        hdr = np.zeros((h, w, 4), np.float32)
        world.render(orc.config(w, h, spp=1, max_depth=depth, first_sample=1, flags=flags), hdr)
        t0 = time.perf_counter()
        _, st = world.render(orc.config(w, h, spp=frames, max_depth=depth, first_sample=2, flags=flags), hdr)
        el = time.perf_counter() - t0
        world.close()
        print(f"threads {threads:3d}  {name:40s} {st.paths / el:.4e} path samples/s", flush=True)
