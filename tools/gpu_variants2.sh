#!/bin/bash
# r01y A/B: RT_FLAG_ACCEL entry-list / CTA-size / flat-level variants at the new group size, direction filter in the brute-force cull.
set -u
TAG=${1:-var2}; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
RT_PROBE_FLAGS=8 python tools/variants.py base= cap12=-DRT_ACCEL_ENTRY_CAP=12 cap16=-DRT_ACCEL_ENTRY_CAP=16 tpb512=-DRT_ACCEL_MAX_TPB=512 tpb768=-DRT_ACCEL_MAX_TPB=768 flat8=-DRT_ACCEL_FLAT_NODES=8 > "$OUT/accel_variants.log" 2>&1
cat "$OUT/accel_variants.log"
python tools/variants.py base= dirf=-DRT_CULL_DIRFILTER=1 > "$OUT/brute_variants.log" 2>&1
cat "$OUT/brute_variants.log"
RT_B200_LIB=$PWD/cpuraytracer_b200/csrc/librt_b200_dirf.so python -m pytest tests -m gpu -x -q > "$OUT/pytest_dirf.log" 2>&1; echo "pytest dirf exit $?"; tail -3 "$OUT/pytest_dirf.log"
