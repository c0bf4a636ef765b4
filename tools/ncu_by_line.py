#!/usr/bin/env python
"""Joins an `ncu --page source --csv` dump with `nvdisasm --print-line-info` so samples / executed instructions can be
read per CUDA source line.  usage: ncu_by_line.py <ncu_source.csv> <nvdisasm.sass> <mangled kernel name> [top]"""
import collections
import csv
import re
import sys

src_csv, sass, kname = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
rows = list(csv.reader(open(src_csv)))
hdr, data = rows[1], rows[2:]
iS, iI, iT = hdr.index('# Samples'), hdr.index('Instructions Executed'), hdr.index('Thread Instructions Executed')
lines = open(sass).read().splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('.text.' + kname + ':'))
cur, tags = ('?', 0), []
for l in lines[start + 1:]:
    if l.startswith('//----') or l.startswith('\t.section'):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2)))
        continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+\S', l):
        tags.append(cur)
assert len(tags) == len(data), (len(tags), len(data))
agg = collections.defaultdict(lambda: [0, 0, 0])
for t, r in zip(tags, data):
    a = agg[t]
    a[0] += int(r[iS]); a[1] += int(r[iI]); a[2] += int(r[iT])
ts = sum(a[0] for a in agg.values()); ti = sum(a[1] for a in agg.values())
print(f'total samples {ts} warp-instructions {ti}')
for t, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f'{t[0]}:{t[1]:<5d} samples {100*a[0]/ts:6.2f}%  instr {100*a[1]/ti:6.2f}%  threads/instr {a[2]/max(a[1],1):5.1f}')
