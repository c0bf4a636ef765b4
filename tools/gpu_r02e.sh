#!/bin/bash
# Round 2 (1 GPU): after the tail-chunk change -- schedule/kernel parity tests, bench with both schedules.
set -u
OUT=gpurun_out/r02e; mkdir -p "$OUT"
timeout 900 python -m pytest tests -m gpu -x -q -rs -k "two_stage or batches or overwrite or deterministic or chunked or several_lights or path_for_path or asynchronous" > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?"; tail -4 "$OUT/pytest_gpu.log"
python bench.py --no-cpu-baseline --no-extras > "$OUT/bench.json" 2> "$OUT/bench.err"; echo "bench exit $?"; tail -3 "$OUT/bench.err"
python bench.py --no-cpu-baseline --no-extras --query accel > "$OUT/bench_accel.json" 2> "$OUT/bench_accel.err"; echo "bench accel exit $?"
python bench.py --no-cpu-baseline --no-extras --query accel --schedule one-kernel > "$OUT/bench_accel_1k.json" 2> "$OUT/bench_accel_1k.err"; echo "bench accel 1k exit $?"
python - <<'PY'
import json
for f in ("bench","bench_accel","bench_accel_1k"):
    d=json.loads(open(f"gpurun_out/r02e/{f}.json").read().strip().splitlines()[-1])
    print(f, "%.4e"%d["value"], "ms/step %.1f"%d["ms_per_step"], {k:"%.4e"%d[k]["value"] for k in ("accel","one_kernel_schedule") if k in d})
PY
