#!/bin/bash
# RT_FLAG_ACCEL group size A/B (8 / 16 / 32 slots per group) + parity tests per size + all-kernels self-check (one GPU).
set -u
TAG=${1:-gshift}
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
for g in 5 4 3; do
  echo "=== RT_ACCEL_GSHIFT=$g" | tee -a "$OUT/probe.log"
  RT_ACCEL_GSHIFT=$g RT_PROBE_FLAGS=8 python tools/perf_probe.py >> "$OUT/probe.log" 2>&1
done
echo "=== brute force" >> "$OUT/probe.log"
python tools/perf_probe.py >> "$OUT/probe.log" 2>&1
cat "$OUT/probe.log"
for g in 3 4; do
  RT_ACCEL_GSHIFT=$g python -m pytest tests -m gpu -x -q -k "accel" > "$OUT/pytest_accel_g$g.log" 2>&1; echo "pytest accel g=$g exit $?"; tail -2 "$OUT/pytest_accel_g$g.log"
done
python -m pytest tests -m gpu -x -q > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?"; tail -3 "$OUT/pytest_gpu.log"
python tools/allkernels_case.py > "$OUT/allkernels.log" 2>&1; echo "allkernels exit $?"; tail -3 "$OUT/allkernels.log"
