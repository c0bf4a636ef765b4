#!/bin/bash
# Final sanity of a round: whole GPU suite, smoke, bench lines of every workload (each carries both query modes), probe.
set -u
TAG=${1:-final}; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
python -m pytest tests -m gpu -x -q > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?"; tail -3 "$OUT/pytest_gpu.log"
python -c "import __graft_entry__ as g; g.smoke()" > "$OUT/smoke.log" 2>&1; echo "smoke exit $?"; tail -1 "$OUT/smoke.log"
python bench.py > "$OUT/bench_c3.json" 2> "$OUT/bench_c3.err"; echo "bench c3 exit $?"
for w in c1 c2 c4 c5; do python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline > "$OUT/bench_$w.json" 2> "$OUT/bench_$w.err"; echo "bench $w exit $?"; done
python bench.py --workload c5 --query accel --steps 3 --warmup 3 --no-cpu-baseline > "$OUT/bench_c5_accel.json" 2> "$OUT/bench_c5_accel.err"; echo "bench c5 accel exit $?"
for f in "$OUT"/bench_*.json; do python -c "
import json
d=json.loads(open('$f').read().strip().splitlines()[-1]); a=d.get('accel',{})
print('$f', 'value %.4e'%d['value'], 'frac %.3f'%d['roofline']['frac'], 'e2e %.4e'%d['e2e']['value'], 'accel', a.get('value'), 'traffic', d['roofline']['traffic'])"; done
python tools/perf_probe.py > "$OUT/probe_brute.log" 2>&1; cat "$OUT/probe_brute.log"
RT_PROBE_FLAGS=8 python tools/perf_probe.py > "$OUT/probe_accel.log" 2>&1; cat "$OUT/probe_accel.log"
