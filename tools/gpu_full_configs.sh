#!/bin/bash
# BASELINE.json's configs rendered IN FULL (their own resolution, spp and depth) by the headless CLI on N GPUs of one box
# (sample-range sharding + ncclReduce, include/rt_b200_group.h); brute force and RT_FLAG_ACCEL images compared byte for byte.
set -u
TAG=${1:-full}; N=${2:-8}; SKIP_C5=${3:-}
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
run() {  # name, then CLI arguments
  local name=$1; shift
  ./apps/spheres "$@" --gpus $N --out "$OUT/$name" --format ppm > "$OUT/$name.json" 2> "$OUT/$name.err"; echo "$name exit $?"; cat "$OUT/$name.json"
  sha256sum "$OUT/$name.ppm" | cut -c1-64 > "$OUT/$name.sha256"
}
run c2_brute --scene d488 --width 1920 --height 1080 --spp 256 --max-depth 0
run c2_accel --scene d488 --width 1920 --height 1080 --spp 256 --max-depth 0 --accel
run c3_brute --scene d488 --width 1920 --height 1080 --spp 1024
run c3_accel --scene d488 --width 1920 --height 1080 --spp 1024 --accel
run c4_brute --scene d488 --width 3840 --height 2160 --spp 1024
run c4_accel --scene d488 --width 3840 --height 2160 --spp 1024 --accel
[ -n "$SKIP_C5" ] || run c5_accel --scene s4096 --width 3840 --height 2160 --spp 4096 --accel
[ -n "$SKIP_C5" ] || run c5_brute --scene s4096 --width 3840 --height 2160 --spp 4096
for c in c2 c3 c4 c5; do [ -f "$OUT/${c}_brute.ppm" ] && cmp "$OUT/${c}_brute.ppm" "$OUT/${c}_accel.ppm" && echo "$c: brute force and RT_FLAG_ACCEL images identical ($(cat $OUT/${c}_brute.sha256))"; done | tee "$OUT/compare.txt"
rm -f "$OUT"/*.ppm
python -m pytest tests/test_cpp_host.py -m gpu -x -q > "$OUT/pytest_cpp_host.log" 2>&1; echo "pytest cpp host exit $?"; tail -2 "$OUT/pytest_cpp_host.log"
