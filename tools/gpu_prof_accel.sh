#!/bin/bash
# ncu full set of the RT_FLAG_ACCEL render kernel (d488, 16 spp) + s4096.
set -u
TAG=${1:-prof}; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
python tools/ncu_case.py d488 1920 1080 16 50 8 > "$OUT/accel_plain.log" 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rt_render_kernel -s 1 -c 1 -o "$OUT/prof_accel_d488" \
    python tools/ncu_case.py d488 1920 1080 16 50 8 > "$OUT/ncu_full.log" 2>&1
echo "ncu full exit $?"
cat "$OUT/accel_plain.log"
