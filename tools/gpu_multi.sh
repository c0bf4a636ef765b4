#!/bin/bash
# Multi-GPU pass (gpurun --gpus N): GPU tests (incl. the 2-GPU CLI test), bench at 1..N ranks via torchrun, the C++ CLI on N GPUs.
# usage: bash tools/gpu_multi.sh <tag> <N>
set -u
TAG=${1:-multi}; N=${2:-2}
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
nvidia-smi -L > "$OUT/smi.txt" 2>&1
python -m pytest tests -m gpu -x -q > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?"; tail -3 "$OUT/pytest_gpu.log"
python bench.py --no-cpu-baseline > "$OUT/bench_n1.json" 2> "$OUT/bench_n1.err"; echo "bench n1 exit $?"
n=2
while [ $n -le $N ]; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n > "$OUT/bench_n$n.json" 2> "$OUT/bench_n$n.err"; echo "bench n$n exit $?"
  n=$((n*2))
done
for f in "$OUT"/bench_n*.json; do python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1]); print('$f', d['n_gpus'], '%.4e'%d['value'], 'e2e %.4e'%d['e2e']['value'], 'frac %.3f'%d['roofline']['frac'])"; done
./apps/spheres --width 1920 --height 1080 --spp 256 --gpus $N --out "$OUT/cli" --format ppm > "$OUT/cli_n$N.json" 2> "$OUT/cli_n$N.err"; echo "cli exit $?"; cat "$OUT/cli_n$N.json"
./apps/spheres --width 1920 --height 1080 --spp 256 --gpus 1 --out "$OUT/cli1" --format ppm --quiet > "$OUT/cli_n1.json" 2> "$OUT/cli_n1.err"; cat "$OUT/cli_n1.json"
ls -la "$OUT"
