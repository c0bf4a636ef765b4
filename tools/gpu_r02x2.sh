#!/bin/bash
# Round 2 (1 GPU): batch size at 3840x2160 -- 16 frames (256 MiB rule) vs 32 frames (512 MiB for every size), 64 frames per call.
set -u
OUT=gpurun_out/r02x2; mkdir -p "$OUT"
for v in base capall; do
  lib=cpuraytracer_b200/csrc/librt_b200_$v.so; [ $v = base ] && lib=cpuraytracer_b200/csrc/librt_b200.so
  echo "=== $v" | tee -a "$OUT/probe.log"
  RT_PROBE_RES=3840x2160 RT_PROBE_SPP=64 RT_B200_LIB=$PWD/$lib python tools/perf_probe.py >> "$OUT/probe.log" 2>&1
done
cat "$OUT/probe.log"
