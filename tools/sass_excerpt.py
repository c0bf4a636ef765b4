#!/usr/bin/env python
"""Writes profiles/r02_cull_loop.sass: the cull-loop bodies (one unrolled group of 16 sphere pairs = 32 spheres) of the
brute-force path kernel (closest + shadow) and first-hit kernel (closest only), cut out of `cuobjdump -sass` of the in-tree
librt_b200.so, with the instruction counts per sphere pair that DESIGN.md section 3 argues from.
usage: python tools/sass_excerpt.py [out]"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "cpuraytracer_b200", "csrc", "librt_b200.so")
out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02_cull_loop.sass")
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
kernels, cur = collections.OrderedDict(), None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1); kernels[cur] = []
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4})\*/\s+(.*?)\s*;", line)
    if m and cur:
        kernels[cur].append((m.group(1), m.group(2)))


def cls(op):
    o = op.split()[0]
    if o.startswith("@"):
        o = op.split()[1]
    return "F" if o.startswith(("FFMA2", "FADD2")) else "S" if o.startswith("SHF") else "L" if o.startswith("LDS") else "o"


text = ["# cuobjdump -sass of cpuraytracer_b200/csrc/librt_b200.so (sm_100a), tools/sass_excerpt.py",
        "# One unrolled group (16 sphere pairs = 32 spheres x the rays of the pass) of the conservative cull, rt_device.cuh cull_pair().",
        "# FFMA2 / FADD2 = fma.rn.f32x2 / add.rn.f32x2 (Blackwell packed FP32: two spheres per instruction, 2 FMA-pipe cycles each);",
        "# SHF.L.W.U32.HI = funnel shift collecting one sign bit into the 32-sphere candidate mask; LDS.128 = broadcast load of a pair.", ""]
for want, title in (("rt_path_kernelILi0ELb0ELi5ELb0", "rt_path_kernel<MODE_CULL> -- closest + shadow ray from one origin"),
                    ("rt_first_hit_kernelILi0ELi5", "rt_first_hit_kernel<MODE_CULL> -- closest ray only")):
    name = next(k for k in kernels if want in k)
    ins = kernels[name]
    pat = "".join(cls(op) for _, op in ins)
    best = max(re.finditer(r"[FSL]{1}[FSLo]{60,}?(?=o{3,}|$)", pat), key=lambda m: m.group(0).count("F"))
    a = best.start() + best.group(0).index("F"); b = best.start() + len(best.group(0).rstrip("o"))
    while a > 0 and pat[a - 1] == "L":                       # the loads that feed the first FMAs
        a -= 1
    body = ins[a:b]
    c = collections.Counter(cls(op) for _, op in body)
    n_pairs = 16
    text += [f"## {title}", f"## {name}: {len(ins)} SASS instructions in the kernel; loop body below = {len(body)} instructions for {n_pairs} pairs:",
             f"##   per pair: {c['F'] / n_pairs:.2f} FFMA2/FADD2 + {c['S'] / n_pairs:.2f} SHF + {c['L'] / n_pairs:.2f} LDS + {c['o'] / n_pairs:.2f} other"
             f"  = {len(body) / n_pairs:.2f} issue slots, {2 * c['F'] / n_pairs:.1f} FMA-pipe cycles", f"##   order: {pat[a:b]}", ""]
    text += [f"        /*{addr}*/  {op} ;" for addr, op in body] + [""]
open(out, "w").write("\n".join(text))
print("\n".join(l for l in text if l.startswith("##")))
