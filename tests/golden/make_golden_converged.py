#!/usr/bin/env python
"""Regenerates tests/golden/golden_converged.npz: converged (4096 spp) linear mean-radiance images of BASELINE.json's configs
at 1/8 scale (240x135, same 16:9 aspect and camera), rendered by oracle/_ref -- the reference's OWN common-lib sources
(atomic Halton sampler, BVH, recursive integrator; all host threads) -- plus the Monte Carlo noise floor of each config
(relRMSE between two counter-RNG seeds of the oracle at the same spp).  Needs /root/reference (build container):

    python tests/golden/make_golden_converged.py          # ~25 min on 8 cores

c2 = configs[1] (direct light only), c3 = configs[2] (= configs[3] at this scale: same camera, aperture 0.4), c5 = configs[4]
(4096 spheres).  Stored as float16 of the mean radiance / 1024 (values up to ~40000 would overflow binary16; its 2^-11 relative
precision is 20x below the 1e-2 bar).
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import cpuraytracer_b200 as rt  # noqa: E402
from oracle import orc  # noqa: E402

assert orc.load_ref() is not None, "oracle/_ref/libref.so missing: run `make -C oracle` where /root/reference exists"
W, H, SPP = 240, 135, 4096
CONFIGS = {"c2": ("d488", 0), "c3": ("d488", 50), "c5": ("s4096", 50)}


def rel_rmse(g, o):
    g = np.asarray(g, np.float64); o = np.asarray(o, np.float64)
    return float(np.sqrt(np.mean(((g - o) / (o + 1e-3 * o.mean())) ** 2)))


out = {}
cam = rt.scenes.default_camera(aspect=W / H)
cp = orc.camera_params(cam.origin, cam.look_at, cam.vertical_fov, cam.aspect_ratio, cam.focal_length, cam.aperture)
for key, (scene, depth) in CONFIGS.items():
    t0 = time.time()
    osc = orc.Scene.from_flat(rt.scenes.NAMED[scene]().export())
    hdr, _ = orc.ref_render(osc, cp, orc.config(W, H, spp=SPP, max_depth=depth, threads=0))
    ref = hdr[..., :3] / SPP
    # noise floor: two counter-RNG seeds of the oracle (its restated BVH: same hits as brute force, affordable at 4096 spheres)
    a, _ = orc.render(osc, cp, orc.config(W, H, spp=SPP, max_depth=depth, seed=101, accel=orc.ACCEL_BVH))
    b, _ = orc.render(osc, cp, orc.config(W, H, spp=SPP, max_depth=depth, seed=202, accel=orc.ACCEL_BVH))
    floor = rel_rmse(a[..., :3] / SPP, b[..., :3] / SPP)
    ex = float(2.0 ** 15)                                   # undo the exposure factor 2^-15 the running sums carry
    out[key + "_mean"] = (ref * ex / 1024.0).astype(np.float16)
    out[key + "_floor"] = np.float64(floor)
    out[key + "_ref_vs_counter"] = np.float64(rel_rmse(a[..., :3] / SPP, ref))
    np.savez_compressed(os.path.join(HERE, "golden_converged.npz"), **out)      # incremental: a later config may take long
    print(key, scene, depth, f"floor {floor:.4f} ref-vs-counter {out[key + '_ref_vs_counter']:.4f} max {float((ref * ex).max()):.1f} {time.time() - t0:.0f}s", flush=True)
np.savez_compressed(os.path.join(HERE, "golden_converged.npz"), **out)
print("wrote golden_converged.npz", os.path.getsize(os.path.join(HERE, "golden_converged.npz")))
