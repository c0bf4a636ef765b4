#!/bin/bash
# A/B of variant libraries built BEFORE the call (cpuraytracer_b200/csrc/librt_b200_<name>.so travel with the snapshot):
# usage: bash tools/gpu_prebuilt_variants.sh <tag> <name> [<name> ...]   ("base" = the product library)
set -u
TAG=${1:-pv}; shift; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
for v in "$@"; do
  lib=$PWD/cpuraytracer_b200/csrc/librt_b200_$v.so; [ "$v" = base ] && lib=$PWD/cpuraytracer_b200/csrc/librt_b200.so
  echo "=== $v accel" | tee -a "$OUT/variants.log"; RT_B200_LIB=$lib RT_PROBE_FLAGS=8 python tools/perf_probe.py 2>&1 | tee -a "$OUT/variants.log"
  echo "=== $v brute" | tee -a "$OUT/variants.log"; RT_B200_LIB=$lib python tools/perf_probe.py d488 2>&1 | tee -a "$OUT/variants.log"
done
