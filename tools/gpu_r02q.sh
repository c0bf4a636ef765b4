#!/bin/bash
# Round 2 (1 GPU): light-aligned cull frame -- whole parity suite, kernel-time probe, launch list.
set -u
OUT=gpurun_out/r02q; mkdir -p "$OUT"
timeout 1500 python -m pytest tests -m gpu -x -q -rs > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?" | tee -a "$OUT/pytest_gpu.log"; tail -5 "$OUT/pytest_gpu.log"
python tools/perf_probe.py > "$OUT/probe.log" 2>&1; cat "$OUT/probe.log"
RT_PROBE_FLAGS=8 python tools/perf_probe.py > "$OUT/probe_accel.log" 2>&1; cat "$OUT/probe_accel.log"
ncu --metrics gpu__time_duration.sum --clock-control none -c 100 --csv --log-file "$OUT/launches.csv" python tools/ncu_case.py d488 1920 1080 64 50 > "$OUT/ncu_l.log" 2>&1
grep -h "first_hit\|path_kernel" "$OUT"/launches.csv | tail -2 | cut -c60-260
