#!/bin/bash
# RT_FLAG_ACCEL pass on a GPU box: parity tests of the hierarchy against brute force, then the A/B probe.
set -u
TAG=${1:-accel}; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "accel" > "$OUT/pytest_accel.log" 2>&1; echo "pytest exit $?" | tee -a "$OUT/pytest_accel.log"
tail -5 "$OUT/pytest_accel.log"
if [ -z "${RT_SKIP_BRUTE:-}" ]; then timeout 300 python tools/perf_probe.py > "$OUT/probe_brute.log" 2>&1; echo "probe brute exit $?"; cat "$OUT/probe_brute.log"; fi
RT_PROBE_FLAGS=8 timeout 300 python tools/perf_probe.py > "$OUT/probe_accel.log" 2>&1; echo "probe accel exit $?"
cat "$OUT/probe_accel.log"
