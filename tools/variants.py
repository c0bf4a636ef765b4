#!/usr/bin/env python
"""Builds compile-time variants of the CUDA library and runs tools/perf_probe.py on each (GPU box).
usage: variants.py name=flag,flag ...   e.g.  base=  pf=-DRT_PREFETCH=1  r64=-maxrregcount=64"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cpuraytracer_b200 import build  # noqa: E402

scenes = [a for a in sys.argv[1:] if "=" not in a]
for spec in [a for a in sys.argv[1:] if "=" in a]:
    name, flags = spec.split("=", 1)
    env = dict(os.environ)
    fl = []
    for f in flags.split(","):
        if f.startswith("env:"):
            k, v = f[4:].split(":")
            env[k] = v
        elif f:
            fl.append(f)
    out = os.path.join(ROOT, "cpuraytracer_b200", "csrc", f"librt_b200_{name}.so")
    build.build(out=out, extra=fl)
    env["RT_B200_LIB"] = out
    print(f"=== {name}: {flags}", flush=True)
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "perf_probe.py"), *scenes], env=env)
