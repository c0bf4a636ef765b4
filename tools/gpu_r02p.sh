#!/bin/bash
# Round 2 (1 GPU): multi-ray first-hit stage -- parity suite on the default build (4 rays per lane, 512-thread bound), then kernel-time A/B of the
# variants (rays per lane x register budget) and the ncu launch list (per-kernel times) of the default and the 1-ray build.
set -u
OUT=gpurun_out/r02p; mkdir -p "$OUT"
timeout 1500 python -m pytest tests -m gpu -x -q -rs > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?" | tee -a "$OUT/pytest_gpu.log"; tail -5 "$OUT/pytest_gpu.log"
for v in r1 r2_768 r2_512 r4_640 r4_512 r4_384; do
  lib=cpuraytracer_b200/csrc/librt_b200_$v.so; [ $v = r4_512 ] && lib=cpuraytracer_b200/csrc/librt_b200.so
  echo "=== $v" | tee -a "$OUT/probe.log"
  RT_B200_LIB=$PWD/$lib python tools/perf_probe.py >> "$OUT/probe.log" 2>&1
done
cat "$OUT/probe.log"
for v in r1 r4_512; do
  lib=cpuraytracer_b200/csrc/librt_b200_$v.so; [ $v = r4_512 ] && lib=cpuraytracer_b200/csrc/librt_b200.so
  RT_B200_LIB=$PWD/$lib ncu --metrics gpu__time_duration.sum --clock-control none -c 100 --csv --log-file "$OUT/launches_$v.csv" python tools/ncu_case.py d488 1920 1080 64 50 > "$OUT/ncu_l_$v.log" 2>&1
done
grep -h "first_hit" "$OUT"/launches_r1.csv | tail -1 | cut -c1-200; grep -h "first_hit" "$OUT"/launches_r4_512.csv | tail -1 | cut -c1-200
