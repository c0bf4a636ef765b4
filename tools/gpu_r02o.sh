#!/bin/bash
# Round 2 (1 GPU): rt_render_static test alone (fresh process) and after the whole GPU suite; record-sector-load variants A/B.
set -u
OUT=gpurun_out/r02o; mkdir -p "$OUT"
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > "$OUT/pytest_no_alloc_alone.log" 2>&1; echo "no_alloc alone exit $?" | tee -a "$OUT/pytest_no_alloc_alone.log"; tail -5 "$OUT/pytest_no_alloc_alone.log"
timeout 1500 python -m pytest tests -m gpu -x -q -rs > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?" | tee -a "$OUT/pytest_gpu.log"; tail -5 "$OUT/pytest_gpu.log"
for v in "base cpuraytracer_b200/csrc/librt_b200.so" "recsec cpuraytracer_b200/csrc/librt_b200_recsec.so" "recsec16 cpuraytracer_b200/csrc/librt_b200_recsec16.so"; do
  set -- $v
  echo "=== $1" | tee -a "$OUT/probe.log"
  RT_B200_LIB=$PWD/$2 python tools/perf_probe.py d488 >> "$OUT/probe.log" 2>&1
  RT_B200_LIB=$PWD/$2 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:rt_ --csv \
      --log-file "$OUT/traffic_c3_brute_$1.csv" python tools/ncu_case.py d488 1920 1080 256 50 0 > "$OUT/ncu_traffic_$1.log" 2>&1; echo "ncu traffic $1 exit $?"
done
cat "$OUT/probe.log"
