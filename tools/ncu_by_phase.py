#!/usr/bin/env python
"""Sums a tools/ncu_by_line.py listing over named source-line ranges.  usage: ncu_by_phase.py <byline.txt> name:file:lo-hi ..."""
import re, sys
rows = []
for l in open(sys.argv[1]).read().splitlines()[1:]:
    m = re.match(r'(\S+):(\d+)\s+samples\s+([\d.]+)%\s+instr\s+([\d.]+)%\s+threads/instr\s+([\d.]+)', l)
    if m: rows.append([m.group(1), int(m.group(2)), float(m.group(3)), float(m.group(4)), float(m.group(5)), False])
for spec in sys.argv[2:]:
    name, f, rng = spec.split(':'); lo, hi = map(int, rng.split('-'))
    sel = [r for r in rows if r[0].startswith(f) and lo <= r[1] <= hi and not r[5]]
    for r in sel: r[5] = True
    s = sum(r[2] for r in sel); i = sum(r[3] for r in sel); ti = sum(r[3] * r[4] for r in sel)
    print(f"{name:28s} samples {s:6.2f}%  instr {i:6.2f}%  threads/instr {ti / max(i, 1e-9):5.1f}")
rest = [r for r in rows if not r[5]]
s = sum(r[2] for r in rest); i = sum(r[3] for r in rest); ti = sum(r[3] * r[4] for r in rest)
print(f"{'(rest)':28s} samples {s:6.2f}%  instr {i:6.2f}%  threads/instr {ti / max(i, 1e-9):5.1f}")
