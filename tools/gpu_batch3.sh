#!/bin/bash
# r01z batch: group-size parity test, RT_FLAG_ACCEL with the direction filter in its cull.
set -u
TAG=${1:-b3}; shift; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
python -m pytest tests -m gpu -x -q -k "group_sizes or random_scenes" > "$OUT/pytest_new.log" 2>&1; echo "pytest new exit $?"; tail -3 "$OUT/pytest_new.log"
RT_PROBE_FLAGS=8 python tools/variants.py base= dirf=-DRT_CULL_DIRFILTER=1 "$@" > "$OUT/accel_variants.log" 2>&1
grep -v "^$" "$OUT/accel_variants.log"
