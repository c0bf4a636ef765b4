#!/bin/bash
# Round-2 multi-GPU pass (gpurun --gpus N): GPU suite incl. the 2-GPU CLI test, weak-scaling bench at 2..N ranks (parity_n), strong scaling
# of c2 / c3 (brute force) and c5 (RT_FLAG_ACCEL) at 1..N ranks.   usage: bash tools/gpu_round2_multi.sh <tag> <N>
set -u
TAG=${1:-r02m}; N=${2:-2}
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
nvidia-smi -L > "$OUT/smi.txt" 2>&1
if [ -z "${SKIP_PYTEST:-}" ]; then timeout 1500 python -m pytest tests -m gpu -x -q -rs > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?" | tee -a "$OUT/pytest_gpu.log"; tail -4 "$OUT/pytest_gpu.log"; fi
run() {  # run <n> <outfile> <bench args...>
  n=$1; f=$2; shift 2
  if [ "$n" -eq 1 ]; then python bench.py --gpus 1 "$@" > "$f" 2> "${f%.json}.err"
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n "$@" > "$f" 2> "${f%.json}.err"; fi
  echo "$f exit $?"
}
n=2; while [ $n -le $N ]; do run $n "$OUT/bench_weak_c3_n$n.json" --no-extras; n=$((n*2)); done
n=1; while [ $n -le $N ]; do
  run $n "$OUT/strong_c3_n$n.json" --scaling strong --workload c3 --steps 3 --warmup 1 --no-extras --no-cpu-baseline
  run $n "$OUT/strong_c2_n$n.json" --scaling strong --workload c2 --steps 6 --warmup 2 --no-extras --no-cpu-baseline
  run $n "$OUT/strong_c5_accel_n$n.json" --scaling strong --workload c5 --query accel --steps 1 --warmup 1 --no-extras --no-cpu-baseline
  n=$((n*2))
done
python - <<PY
import glob, json
for f in sorted(glob.glob("$OUT/*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split("/")[-1], d["n_gpus"], d["scaling"], "%.4e" % d["value"], "e2e %.4e" % d["e2e"]["value"], "ms/step %.2f" % d["ms_per_step"], "parity_n" in d)
    except Exception as e:
        print(f, "ERR", e)
PY
