#!/bin/bash
# Round 2 (1 GPU): full parity suite (incl. big-scene, async, golden bars when the new golden is in), HBM traffic of one bench step.
set -u
OUT=gpurun_out/r02f; mkdir -p "$OUT"
timeout 1500 python -m pytest tests -m gpu -x -q -rs > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?" | tee -a "$OUT/pytest_gpu.log"; tail -6 "$OUT/pytest_gpu.log"
for q in "brute 0" "accel 8"; do
  set -- $q
  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:rt_ --csv \
      --log-file "$OUT/traffic_c3_$1.csv" python tools/ncu_case.py d488 1920 1080 256 50 $2 > "$OUT/ncu_traffic_$1.log" 2>&1
  echo "ncu traffic $1 exit $?"
done
RT_PROBE_FLAGS=0 python tools/perf_probe.py > "$OUT/probe.log" 2>&1; cat "$OUT/probe.log"
