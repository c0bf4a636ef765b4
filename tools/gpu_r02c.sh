#!/bin/bash
# Round 2 (1 GPU): parity suite on the two-stage schedule, launch list (per-kernel times) and full ncu captures of both stages.
set -u
OUT=gpurun_out/r02c; mkdir -p "$OUT"
timeout 1200 python -m pytest tests -m gpu -x -q -rs > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?" | tee -a "$OUT/pytest_gpu.log"; tail -6 "$OUT/pytest_gpu.log"
python tools/ncu_case.py d488 1920 1080 64 50 > "$OUT/case_plain.log" 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 100 --csv --log-file "$OUT/launches_d488_64spp.csv" python tools/ncu_case.py d488 1920 1080 64 50 > "$OUT/ncu_l.log" 2>&1; echo "launch list exit $?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 100 --csv --log-file "$OUT/launches_d488_64spp_accel.csv" python tools/ncu_case.py d488 1920 1080 64 50 8 > "$OUT/ncu_la.log" 2>&1
ncu --set full --clock-control none --import-source on -k regex:rt_path_kernel -s 1 -c 1 -o "$OUT/prof_path" python tools/ncu_case.py d488 1920 1080 16 50 > "$OUT/ncu_full_path.log" 2>&1; echo "ncu path exit $?"
ncu --set full --clock-control none --import-source on -k regex:rt_first_hit_kernel -s 1 -c 1 -o "$OUT/prof_first" python tools/ncu_case.py d488 1920 1080 16 50 > "$OUT/ncu_full_first.log" 2>&1; echo "ncu first exit $?"
ncu --set full --clock-control none --import-source on -k regex:rt_path_kernel -s 1 -c 1 -o "$OUT/prof_path_accel" python tools/ncu_case.py d488 1920 1080 16 50 8 > "$OUT/ncu_full_path_accel.log" 2>&1; echo "ncu path accel exit $?"
cat "$OUT/case_plain.log"; ls -la "$OUT"
