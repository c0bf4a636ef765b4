#!/usr/bin/env python
"""Regenerates tests/golden/golden_converged.npz: converged linear mean-radiance images of BASELINE.json's configs at 1/8
scale (240x135, same 16:9 aspect and camera), rendered by oracle/_ref -- the reference's OWN common-lib sources (atomic Halton
sampler, BVH, recursive integrator) in ONE persistent world per config -- at 65,536 spp, so that the Monte Carlo noise of the
golden is far below BASELINE.json's bars (per-pixel relRMSE < 1e-2, image-mean relative error < 0.5 %) and the GPU test can
assert the bars themselves.  Needs /root/reference (build container); input scenes are the committed oracle/scenes/*.rtsc:

    nice python tests/golden/make_golden_converged.py [threads]        # ~70 min on 6 threads

Per config the render runs as two halves of 32,768 frames (frames 1..32768 and 32769..65536 of the same world: sample
count, frame jitter and the materials' Halton indices continue); `<key>_floor` = relRMSE(half A, half B) / 2 is the standard
estimate of the noise left in their mean (two independent halves with per-pixel sigma s each differ by sqrt(2) s, their mean
carries s / sqrt(2)).  It is a printed diagnostic in the tests, not part of any bar.

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
from oracle import orc  # noqa: E402

assert orc.load_ref() is not None, "oracle/_ref/libref.so missing: run `make -C oracle` where /root/reference exists"
W, H, HALF = 240, 135, 32768
THREADS = int(sys.argv[1]) if len(sys.argv) > 1 else 0
CONFIGS = {"c2": ("d488", 0), "c3": ("d488", 50), "c5": ("s4096", 50)}
OUT = os.environ.get("GOLDEN_OUT", os.path.join(HERE, "golden_converged.npz"))


def rel_rmse(g, o):
    g = np.asarray(g, np.float64); o = np.asarray(o, np.float64)
    return float(np.sqrt(np.mean(((g - o) / (o + 1e-3 * o.mean())) ** 2)))


out = {"spp": np.int64(2 * HALF)}
cp = orc.default_camera_params(W / H)
for key, (scene, depth) in CONFIGS.items():
    t0 = time.time()
    world = orc.RefWorld(orc.frozen_scene(scene), cp, THREADS)
    halves = []
    for k in range(2):
        hdr = np.zeros((H, W, 4), np.float32)
        done = 0
        while done < HALF:                                   # blocks of 4096 frames: progress lines, same world throughout
            n = min(4096, HALF - done)
            world.render(orc.config(W, H, spp=n, max_depth=depth, first_sample=1 + k * HALF + done), hdr)
            done += n
            print(f"  {key} half {k}: {done}/{HALF} frames, {time.time() - t0:.0f} s", flush=True)
        assert np.all(hdr[..., 3] == HALF)
        halves.append(hdr[..., :3].astype(np.float64) / HALF)
    world.close()
    ref = 0.5 * (halves[0] + halves[1])
    floor = 0.5 * rel_rmse(halves[0], halves[1])
    ex = float(2.0 ** 15)                                   # undo the exposure factor 2^-15 the running sums carry
    out[key + "_mean"] = (ref * ex / 1024.0).astype(np.float16)
    out[key + "_floor"] = np.float64(floor)
    np.savez_compressed(OUT, **out)                         # incremental: a later config may take long
    print(key, scene, depth, f"spp {2 * HALF} golden-noise estimate {floor:.5f} max {float((ref * ex).max()):.1f} {time.time() - t0:.0f}s", flush=True)
np.savez_compressed(OUT, **out)
print("wrote golden_converged.npz", os.path.getsize(OUT))
