#!/bin/bash
# ncu evidence for the two layouts: full set on the megakernel render kernel, launch list of a wavefront render.
set -u
TAG=${1:-prof}; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
python tools/ncu_case.py d488 1920 1080 16 50 > "$OUT/mega_plain.log" 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rt_render_kernel -s 1 -c 1 -o "$OUT/prof_render" \
    python tools/ncu_case.py d488 1920 1080 16 50 > "$OUT/ncu_full.log" 2>&1
echo "ncu full exit $?"
python tools/ncu_case.py d488 1920 1080 16 50 4 > "$OUT/wf_plain.log" 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file "$OUT/wf_launches.csv" \
    python tools/ncu_case.py d488 1920 1080 16 50 4 > "$OUT/ncu_wf.log" 2>&1
echo "ncu wf exit $?"
cat "$OUT/mega_plain.log" "$OUT/wf_plain.log"
