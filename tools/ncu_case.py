#!/usr/bin/env python
"""One short render for ncu: ncu_case.py <scene> <w> <h> <spp> <depth> [flags] (two launches: warm-up + profiled)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cpuraytracer_b200 as rt
name, w, h, spp, depth = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
flags = int(sys.argv[6], 0) if len(sys.argv) > 6 else 0          # 4 = RT_FLAG_WAVEFRONT
r = rt.Renderer(0, w, h)
r.set_scene(rt.scenes.NAMED[name]())
r.set_camera(rt.scenes.default_camera(aspect=w / h))
for _ in range(2):
    r.clear(); r.reset_stats(); r.render(spp, depth, flags)
    st = r.stats()
print(f"{name} {w}x{h} spp {spp} depth {depth}: {st.render_ms:.2f} ms {st.paths / st.render_ms * 1e3:.3e} paths/s")
