#!/usr/bin/env python
"""Writes oracle/scenes/{d488,s4096,small}.rtsc: the named synthetic scenes (cpuraytracer_b200/scenes.py)
frozen as flat RTSC v1 files, so that the CPU arm of bench.py (`--impl reference`, `cpu_baseline`) and the golden generators
take their input from committed bytes and never load the product library.  tests/test_host_api.py checks that the files
still equal what the product's recorder exports.

    python oracle/scenes/freeze_scenes.py
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import cpuraytracer_b200 as rt  # noqa: E402
from oracle import orc  # noqa: E402

for name in ("d488", "s4096", "small"):
    sc = orc.Scene.from_flat(rt.scenes.NAMED[name]().export())
    path = os.path.join(HERE, name + ".rtsc")
    sc.to_rtsc(path)
    back = orc.Scene.from_rtsc(path)
    assert back.spheres.tobytes() == sc.spheres.tobytes() and back.lights.tobytes() == sc.lights.tobytes()
    print(name, len(sc.spheres), "spheres", os.path.getsize(path), "bytes")
