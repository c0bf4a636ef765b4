#!/bin/bash
# Round 2 (1 GPU): chunk-ladder granularity A/B (8 vs 16 frames per rung unit): kernel time and HBM traffic of one bench step.
set -u
OUT=gpurun_out/r02n; mkdir -p "$OUT"
for v in "gran8 cpuraytracer_b200/csrc/librt_b200.so" "gran16 cpuraytracer_b200/csrc/librt_b200_gran16.so"; do
  set -- $v
  echo "=== $1" | tee -a "$OUT/probe.log"
  RT_B200_LIB=$PWD/$2 python tools/perf_probe.py d488 >> "$OUT/probe.log" 2>&1
  RT_B200_LIB=$PWD/$2 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:rt_ --csv \
      --log-file "$OUT/traffic_c3_brute_$1.csv" python tools/ncu_case.py d488 1920 1080 256 50 0 > "$OUT/ncu_traffic_$1.log" 2>&1; echo "ncu traffic $1 exit $?"
  RT_B200_LIB=$PWD/$2 python bench.py --no-cpu-baseline --no-extras --steps 4 --warmup 3 > "$OUT/bench_$1.json" 2> "$OUT/bench_$1.err"; echo "bench $1 exit $?"
done
cat "$OUT/probe.log"
python - <<'PY'
import json
for v in ("gran8","gran16"):
    d=json.loads(open(f"gpurun_out/r02n/bench_{v}.json").read().strip().splitlines()[-1]); print(v, "%.4e"%d["value"], "ms/step %.1f"%d["ms_per_step"], "accel %.4e"%d["accel"]["value"])
PY
