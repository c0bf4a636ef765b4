#!/bin/bash
# Final kernels on 8 GPUs, trimmed (gpurun --gpus 8): weak scaling of c3 at 2 / 4 / 8 ranks, strong scaling of c3, c2 and c5 (RT_FLAG_ACCEL) at 8 ranks
# (their N = 1 references: profiles/r02/multi/r02s_strong_*_n1.json).   usage: bash tools/gpu_round2_n8.sh <tag>
set -u
TAG=${1:-r02w}
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
nvidia-smi -L > "$OUT/smi.txt" 2>&1
run() { n=$1; f=$2; shift 2
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n "$@" > "$f" 2> "${f%.json}.err"; echo "$f exit $?"; }
for n in 2 4 8; do run $n "$OUT/bench_weak_c3_n$n.json" --no-extras --steps 6 --warmup 3; done
run 8 "$OUT/strong_c3_n8.json" --scaling strong --workload c3 --steps 3 --warmup 1 --no-extras --no-cpu-baseline
run 8 "$OUT/strong_c2_n8.json" --scaling strong --workload c2 --steps 6 --warmup 2 --no-extras --no-cpu-baseline
run 8 "$OUT/strong_c5_accel_n8.json" --scaling strong --workload c5 --query accel --steps 1 --warmup 1 --no-extras --no-cpu-baseline
python - <<PY
import glob, json
for f in sorted(glob.glob("$OUT/*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split("/")[-1], d["n_gpus"], d["scaling"], "%.4e" % d["value"], "e2e %.4e" % d["e2e"]["value"], "ms/step %.2f" % d["ms_per_step"], "parity_n" in d)
    except Exception as e:
        print(f, "ERR", e)
PY
