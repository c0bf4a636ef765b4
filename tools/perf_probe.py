#!/usr/bin/env python
"""Quick A/B probe on a GPU box: kernel-time throughput of the render kernel on the two named scenes."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cpuraytracer_b200 as rt

FLAGS = int(os.environ.get("RT_PROBE_FLAGS", "0"), 0)     # e.g. 4 = RT_FLAG_WAVEFRONT

CASES = [("d488", 1920, 1080, 64, 50), ("d488", 1920, 1080, 64, 0), ("s4096", 1920, 1080, 8, 50)]
if os.environ.get("RT_PROBE_RES"):                       # e.g. 3840x2160
    rw, rh = (int(v) for v in os.environ["RT_PROBE_RES"].split("x"))
    CASES = [(n, rw, rh, s, d) for n, _, _, s, d in CASES]
if os.environ.get("RT_PROBE_SPP"):                       # e.g. 256: the default bench step (several batches per call)
    CASES = [(n, w, h, int(os.environ["RT_PROBE_SPP"]) // (8 if n == "s4096" else 1), d) for n, w, h, _, d in CASES]
if len(sys.argv) > 1:
    CASES = [c for c in CASES if c[0] in sys.argv[1:]]
for name, w, h, spp, depth in CASES:
    r = rt.Renderer(0, w, h)
    r.set_scene(rt.scenes.NAMED[name]())
    r.set_camera(rt.scenes.default_camera(aspect=w / h))
    r.render(spp, depth, FLAGS)
    r.synchronize()
    best = None
    for _ in range(3):
        r.clear()
        r.reset_stats()
        r.render(spp, depth, FLAGS)
        st = r.stats()
        if best is None or st.render_ms < best.render_ms:
            best = st
    s = best.render_ms * 1e-3
    print(f"{name} {w}x{h} spp {spp} depth {depth}: {best.render_ms:8.2f} ms  {best.paths / s:.3e} paths/s  "
          f"{best.sphere_tests / s:.3e} tests/s  {best.sphere_tests * 17 / s / 1e12:.1f} TF-eq  rays/path {(best.closest_rays + best.shadow_rays) / best.paths:.2f} "
          f"passes/path {best.passes * 32 / best.paths:.2f}", flush=True)
    r.close()
