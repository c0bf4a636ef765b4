#!/bin/bash
# Final kernels (128-frame batches) on 8 GPUs, minimal: weak c3, strong c3, strong c2 at 8 ranks.   usage: bash tools/gpu_round2_n8_mini.sh <tag>
set -u
OUT=gpurun_out/${1:-r02v}; mkdir -p "$OUT"
run() { f=$1; shift
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 "$@" > "$f" 2> "${f%.json}.err"; echo "$f exit $?"; }
run "$OUT/bench_weak_c3_n8.json" --no-extras --steps 6 --warmup 3
run "$OUT/strong_c3_n8.json" --scaling strong --workload c3 --steps 4 --warmup 2 --no-extras --no-cpu-baseline
run "$OUT/strong_c2_n8.json" --scaling strong --workload c2 --steps 8 --warmup 3 --no-extras --no-cpu-baseline
python - <<PY
import glob, json
for f in sorted(glob.glob("$OUT/*.json")):
    d = json.loads(open(f).read().strip().splitlines()[-1])
    print(f.split("/")[-1], d["n_gpus"], d["scaling"], "%.4e" % d["value"], "e2e %.4e" % d["e2e"]["value"], "ms/step %.2f" % d["ms_per_step"], d.get("parity_n", {}).get("brute"))
PY
