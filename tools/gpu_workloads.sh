#!/bin/bash
# One bench line per BASELINE config (c1..c5) on one GPU; c5 with a reduced step so it stays within minutes.
set -u
TAG=${1:-wl}; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
for wl in c1 c2 c4 c5; do
  extra=""; [ $wl = c5 ] && extra="--spp-step 16 --steps 2"
  [ $wl = c4 ] && extra="--spp-step 64"
  python bench.py --workload $wl $extra > "$OUT/bench_$wl.json" 2> "$OUT/bench_$wl.err"; echo "$wl exit $?"
  python -c "
import json
d=json.loads(open('$OUT/bench_$wl.json').read().strip().splitlines()[-1]); print('$wl value %.4e e2e %.4e frac %.3f ms/step %.1f rays/path %.2f cpu %.3e (%s, %d cores)'%(d['value'],d['e2e']['value'],d['roofline']['frac'],d['ms_per_step'],d['rays_per_path'],d['cpu_baseline']['value'],d['cpu_baseline']['kind'],d['cpu_baseline']['cores']))"
done
