#!/bin/bash
set -u
OUT=gpurun_out/r02i; mkdir -p "$OUT"
timeout 900 python -m pytest tests -m gpu -x -q -rs -k "two_stage or batches or overwrite or deterministic or chunked or several_lights or path_for_path or ragged" > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?"; tail -3 "$OUT/pytest_gpu.log"
python bench.py --no-cpu-baseline > "$OUT/bench.json" 2> "$OUT/bench.err"; echo "bench exit $?"; tail -3 "$OUT/bench.err"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02i/bench.json").read().strip().splitlines()[-1])
print("c3 %.4e ms/step %.1f"%(d["value"],d["ms_per_step"]), {k:"%.4e"%d[k]["value"] for k in ("accel","one_kernel_schedule")})
for k,v in d["configs"].items(): print(k, "brute %.3e (%.2f)  accel %.3e"%(v["brute"]["value"],v["brute"]["roofline_frac"],v["accel"]["value"]))
PY
