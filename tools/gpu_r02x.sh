#!/bin/bash
# Round 2 (1 GPU): batch size of the two-stage schedule (first-hit records of 256 / 512 / 1024 MiB = 64 / 128 / 256 frames at 1920x1080), 256 spp per call.
set -u
OUT=gpurun_out/r02x; mkdir -p "$OUT"
for v in cap256 cap512 cap1024; do
  lib=cpuraytracer_b200/csrc/librt_b200_$v.so; [ $v = cap256 ] && lib=cpuraytracer_b200/csrc/librt_b200.so
  echo "=== $v" | tee -a "$OUT/probe.log"
  RT_PROBE_SPP=256 RT_B200_LIB=$PWD/$lib python tools/perf_probe.py d488 >> "$OUT/probe.log" 2>&1
done
cat "$OUT/probe.log"
