#!/usr/bin/env python
"""profiles/traffic.json from an ncu CSV (`--metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv`) of
`tools/ncu_case.py <scene> <w> <h> <spp> <depth> [flags]`, which renders twice: the launches of the SECOND render are one bench
step.  usage: ncu_traffic.py <csv> <key e.g. c3:brute> <n_pix> <spp> [traffic.json]
bytes = read + write summed over every launch of the step (tables, first-hit, path, combine); algorithmic bytes of a step =
running sum read + written once per batch (32 B per pixel and batch) + first-hit records written and read once (4 B per path)."""
import collections
import csv
import json
import os
import sys

src, key, n_pix, spp = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
out = sys.argv[5] if len(sys.argv) > 5 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "traffic.json")
rows = [r for r in csv.reader(open(src)) if len(r) > 5]
hdr = rows[0]
iid, ik, im, iu, iv = hdr.index("ID"), hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Unit"), hdr.index("Metric Value")
launch = collections.OrderedDict()
for r in rows[1:]:
    v = float(r[iv].replace(",", ""))
    unit = r[iu].lower()
    scale = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "ns": 1, "us": 1e3, "usecond": 1e3, "nsecond": 1, "msecond": 1e6, "ms": 1e6}.get(unit, 1)
    launch.setdefault(int(r[iid]), {"kernel": r[ik].split("(")[0].replace("void ", "")})[r[im]] = v * scale
ids = sorted(launch)
second = ids[len(ids) // 2:]
per = collections.defaultdict(lambda: {"launches": 0, "bytes": 0.0, "ns": 0.0})
for i in second:
    l = launch[i]
    p = per[l["kernel"]]
    p["launches"] += 1
    p["bytes"] += l.get("dram__bytes_read.sum", 0) + l.get("dram__bytes_write.sum", 0)
    p["ns"] += l.get("gpu__time_duration.sum", 0)
total = sum(p["bytes"] for p in per.values())
n_batches = max(1, sum(p["launches"] for k, p in per.items() if "rt_path_kernel" in k))
algorithmic = 32 * n_pix * n_batches + 4 * n_pix * spp
data = json.load(open(out)) if os.path.exists(out) else {}
data[key] = {"bytes": int(total), "algorithmic_bytes": int(algorithmic), "ratio": total / algorithmic, "batches": n_batches,
             "per_kernel": {k: {"launches": p["launches"], "bytes": int(p["bytes"]), "ms": p["ns"] * 1e-6} for k, p in per.items()},
             "source": os.path.basename(src)}
json.dump(data, open(out, "w"), indent=1)
print(json.dumps(data[key], indent=1))
