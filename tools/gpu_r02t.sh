#!/bin/bash
# Round 2 (1 GPU): streaming (evict-first) access to the first-hit records: HBM traffic and kernel time, rungs of 8 and 16 frames.
set -u
OUT=gpurun_out/r02t; mkdir -p "$OUT"
for v in base recstream recstream16; do
  lib=cpuraytracer_b200/csrc/librt_b200_$v.so; [ $v = base ] && lib=cpuraytracer_b200/csrc/librt_b200.so
  echo "=== $v" | tee -a "$OUT/probe.log"
  RT_B200_LIB=$PWD/$lib python tools/perf_probe.py d488 >> "$OUT/probe.log" 2>&1
  RT_B200_LIB=$PWD/$lib ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:rt_ --csv \
      --log-file "$OUT/traffic_c3_brute_$v.csv" python tools/ncu_case.py d488 1920 1080 256 50 0 > "$OUT/ncu_traffic_$v.log" 2>&1; echo "ncu traffic $v exit $?"
done
cat "$OUT/probe.log"
