#!/bin/bash
# Scaling pass (gpurun --gpus N): bench.py at 1,2,4,..N ranks via torchrun + the C++ CLI (single process, rt_group) at N.
set -u
TAG=${1:-scale}; N=${2:-8}; WL=${3:-c3}; Q=${4:-brute}
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
nvidia-smi -L > "$OUT/smi.txt" 2>&1
python bench.py --workload $WL --query $Q --no-cpu-baseline > "$OUT/bench_${WL}_${Q}_n1.json" 2> "$OUT/bench_${WL}_${Q}_n1.err"; echo "bench n1 exit $?"
n=2
while [ $n -le $N ]; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --workload $WL --query $Q --gpus $n > "$OUT/bench_${WL}_${Q}_n$n.json" 2> "$OUT/bench_${WL}_${Q}_n$n.err"; echo "bench n$n exit $?"
  n=$((n*2))
done
for f in "$OUT"/bench_${WL}_${Q}_n*.json; do python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1]); print('$f', d['n_gpus'], 'value %.4e'%d['value'], 'e2e %.4e'%d['e2e']['value'], 'frac %.3f'%d['roofline']['frac'], 'ms/step %.1f'%d['ms_per_step'])"; done
[ -n "${RT_SKIP_CLI:-}" ] || ./apps/spheres --width 1920 --height 1080 --spp 2048 --frames-per-call 2048 --gpus $N --out "$OUT/cli" --format ppm > "$OUT/cli_n$N.json" 2> "$OUT/cli_n$N.err"; echo "cli exit $?"; cat "$OUT/cli_n$N.json"; tail -2 "$OUT/cli_n$N.err"
rm -f "$OUT"/cli.ppm
