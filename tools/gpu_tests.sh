#!/bin/bash
# Runs a subset of the GPU tests: bash tools/gpu_tests.sh <tag> <pytest -k expression>
set -u
TAG=${1:-t}; K=${2:-accel}; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
timeout 1500 python -m pytest tests -m gpu -x -q -s -k "$K" > "$OUT/pytest.log" 2>&1; echo "pytest exit $?" | tee -a "$OUT/pytest.log"
tail -25 "$OUT/pytest.log"
