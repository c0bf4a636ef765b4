#!/usr/bin/env python
"""Regenerates tests/golden/*.npz from oracle/_ref -- the reference's own camera/light/material/quasi-random/ray-tracing/
texture .cpp files compiled unmodified against oracle/dxm_shim (needs /root/reference; run in the build container):

    python tests/golden/make_golden.py

golden_primary.npz : primary-hit sphere-index / material-id buffers of the default 488-sphere scene and the 4096-sphere scene
golden_render.npz  : running-sum HDR buffers of short single-threaded reference renders (atomic Halton counters, BVH, recursion)
golden_scenes.npz  : SHA-256 of the flat scene bytes the generator in cpuraytracer_b200/scenes.py produces
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import cpuraytracer_b200 as rt  # noqa: E402
from oracle import orc  # noqa: E402

assert orc.load_ref() is not None, "oracle/_ref/libref.so missing: run `make -C oracle` where /root/reference exists"


def cam_of(w, h):
    c = rt.scenes.default_camera(aspect=w / h)
    return orc.camera_params(c.origin, c.look_at, c.vertical_fov, c.aspect_ratio, c.focal_length, c.aperture)


def scene_hash(flat):
    h = hashlib.sha256()
    for a in (flat.spheres, flat.materials, flat.textures, flat.lights):
        h.update(np.ascontiguousarray(a).tobytes())
    h.update(np.array([flat.sky_luminance, flat.sky_texture], np.float64).tobytes())
    return h.hexdigest()


PRIMARY = [("d488", 320, 180, 3, 1), ("d488", 320, 180, 0, 7), ("s4096", 320, 180, 3, 1), ("small", 97, 33, 3, 1)]
RENDER = [("small", 64, 36, 3, 50), ("d488", 64, 36, 2, 50), ("d488", 64, 36, 2, 0)]

prim, rend, hashes = {}, {}, {}
for name in ("d488", "s4096", "small"):
    hashes[name] = scene_hash(rt.scenes.NAMED[name]().export())
for name, w, h, flags, first in PRIMARY:
    osc = orc.Scene.from_flat(rt.scenes.NAMED[name]().export())
    ids, mids = orc.ref_primary_hits(osc, cam_of(w, h), orc.config(w, h, flags=flags, first_sample=first))
    key = f"{name}_{w}x{h}_flags{flags}_n{first}"
    prim[key + "_ids"] = ids.astype(np.int16)
    prim[key + "_mids"] = mids.astype(np.int8)
for name, w, h, spp, depth in RENDER:
    osc = orc.Scene.from_flat(rt.scenes.NAMED[name]().export())
    hdr, st = orc.ref_render(osc, cam_of(w, h), orc.config(w, h, spp=spp, max_depth=depth, threads=1))
    key = f"{name}_{w}x{h}_spp{spp}_depth{depth}"
    rend[key + "_hdr"] = hdr
    rend[key + "_rays"] = np.array([st.closest_rays, st.shadow_rays], np.int64)
np.savez_compressed(os.path.join(HERE, "golden_primary.npz"), **prim)
np.savez_compressed(os.path.join(HERE, "golden_render.npz"), **rend)
np.savez(os.path.join(HERE, "golden_scenes.npz"), **{k: np.array(v) for k, v in hashes.items()})
print("wrote", sorted(prim), sorted(rend), hashes)
