#!/usr/bin/env python
"""Small renders through every kernel of the library (brute force, RT_FLAG_ACCEL, wavefront, RT_FLAG_HALTON, resolve, primary
hits) with cross-checks between the modes.  Written as the compute-sanitizer target; compute-sanitizer is closed on this GPU pool
(r01x: "runs under it have left GPUs needing a reset"), so it serves as a quick all-kernels self-check instead."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import cpuraytracer_b200 as rt

W, H = 72, 40
for scene in ("d488", "s4096"):
    r = rt.Renderer(0, W, H)
    r.set_scene(rt.scenes.NAMED[scene]())
    r.set_camera(rt.scenes.default_camera(aspect=W / H))
    ref = None
    for flags, name in ((0, "brute"), (8, "accel"), (4, "wavefront"), (16, "halton")):
        r.clear(); r.reset_stats()
        r.render(3, 6, flags)
        hdr = r.read_hdr_sum()
        if flags in (0, 8, 4):
            if ref is None: ref = hdr.copy()
            else: assert np.array_equal(ref, hdr), f"{scene} {name}: differs from brute force"
        ldr = r.resolve_ldr()
        print(scene, name, "mean", float(hdr[..., :3].mean()), "ldr", ldr.shape, flush=True)
    a = r.primary_hits(1, 3)
    b = r.primary_hits(1, 3 | 8)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    print(scene, "primary ok", flush=True)
    r.close()
print("allkernels_case done")
