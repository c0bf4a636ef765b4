#!/usr/bin/env python
"""Markdown tables from the bench lines of tools/gpu_round2_multi.sh: weak scaling (c3) and strong scaling (c2, c3, c5/accel).
usage: scaling_table.py <dir with bench_weak_c3_n*.json / strong_*_n*.json> [prefix of the file names]"""
import glob
import json
import os
import re
import sys

d, pre = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
rows = {}
for f in glob.glob(os.path.join(d, pre + "*.json")):
    m = re.match(pre + r"(bench_weak_c3|strong_c3|strong_c2|strong_c5_accel)_n(\d+)\.json", os.path.basename(f))
    if not m:
        continue
    try:
        rows.setdefault(m.group(1), {})[int(m.group(2))] = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception:
        pass
title = {"bench_weak_c3": "c3, brute force, WEAK scaling (256 spp per GPU and step; the driver's contract)",
         "strong_c3": "c3, brute force, STRONG scaling (one step = the config's 1024 spp, split over the ranks)",
         "strong_c2": "c2 (direct only), brute force, STRONG scaling (one step = 256 spp)",
         "strong_c5_accel": "c5 (4096 spheres, 3840x2160), RT_FLAG_ACCEL, STRONG scaling (one step = 4096 spp = 3.4e10 paths)"}
for key in ("bench_weak_c3", "strong_c3", "strong_c2", "strong_c5_accel"):
    if key not in rows:
        continue
    r = rows[key]
    base_n = min(r)
    print(f"\n**{title[key]}**\n")
    print("| GPUs | path samples/s (device) | e2e (host buffers) | ms per step | x of N=%d | efficiency | parity_n: max rel diff xyz (brute / accel), counts, hit buffers |" % base_n)
    print("|---|---|---|---|---|---|---|")
    for n in sorted(r):
        x = r[n]
        sp = x["value"] / r[base_n]["value"]
        pn = x.get("parity_n")
        ptxt = "—" if not pn else "%.1e / %.1e, equal, identical on all ranks" % (pn["brute"]["max_rel_diff_xyz"], pn["accel"]["max_rel_diff_xyz"])
        print(f"| {n} | {x['value']:.4e} | {x['e2e']['value']:.4e} | {x['ms_per_step']:.2f} | {sp:.3f} | {sp / (n / base_n):.4f} | {ptxt} |")
