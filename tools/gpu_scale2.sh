#!/bin/bash
# Weak scaling of the default bench workload in both query modes (gpurun --gpus N): bench.py at 1,2,4,..N ranks via torchrun.
set -u
TAG=${1:-scale2}; N=${2:-8}; WL=${3:-c3}
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
nvidia-smi -L > "$OUT/smi.txt" 2>&1
for Q in brute accel; do
  python bench.py --workload $WL --query $Q --no-cpu-baseline > "$OUT/bench_${WL}_${Q}_n1.json" 2> "$OUT/bench_${WL}_${Q}_n1.err"; echo "bench $Q n1 exit $?"
  n=2
  while [ $n -le $N ]; do
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --workload $WL --query $Q --gpus $n > "$OUT/bench_${WL}_${Q}_n$n.json" 2> "$OUT/bench_${WL}_${Q}_n$n.err"; echo "bench $Q n$n exit $?"
    n=$((n*2))
  done
done
for f in "$OUT"/bench_${WL}_*_n*.json; do python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1]); print('$f', d['n_gpus'], 'value %.4e'%d['value'], 'e2e %.4e'%d['e2e']['value'], 'frac %.3f'%d['roofline']['frac'], 'ms/step %.1f'%d['ms_per_step'], 'rays/s %.4e'%d['rays_per_s'])"; done
