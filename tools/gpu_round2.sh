#!/bin/bash
# Round-2 evidence pass on ONE GPU: parity suite, smoke, default bench + reference arm, ncu launch list of the bench command,
# HBM traffic of one step (both query modes), full ncu captures of both stages (brute force and RT_FLAG_ACCEL).
# usage (under gpurun): bash tools/gpu_round2.sh <tag>
set -u
TAG=${1:-r02z}
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv > "$OUT/smi.txt" 2>&1; nproc >> "$OUT/smi.txt"
timeout 1500 python -m pytest tests -m gpu -x -q -rs > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?" | tee -a "$OUT/pytest_gpu.log"; tail -5 "$OUT/pytest_gpu.log"
python -c "import __graft_entry__ as g; g.smoke()" > "$OUT/smoke.log" 2>&1; echo "smoke exit $?" | tee -a "$OUT/smoke.log"
python bench.py > "$OUT/bench.json" 2> "$OUT/bench.err"; echo "bench exit $?"; tail -3 "$OUT/bench.err"
python bench.py --impl reference --steps 4 --warmup 1 > "$OUT/bench_reference.json" 2> "$OUT/bench_reference.err"; echo "reference exit $?"
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > "$OUT/bench_short.json" 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file "$OUT/launches.csv" \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > "$OUT/ncu_launches.log" 2>&1
echo "ncu launches exit $?"
for q in "brute 0" "accel 8"; do
  set -- $q
  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:rt_ --csv \
      --log-file "$OUT/traffic_c3_$1.csv" python tools/ncu_case.py d488 1920 1080 256 50 $2 > "$OUT/ncu_traffic_$1.log" 2>&1
  echo "ncu traffic $1 exit $?"
done
python tools/ncu_case.py d488 1920 1080 16 50 > "$OUT/ncu_case_plain.log" 2>&1
for spec in "path rt_path_kernel 0" "first rt_first_hit_kernel 0" "path_accel rt_path_kernel 8" "first_accel rt_first_hit_kernel 8"; do
  set -- $spec
  ncu --set full --clock-control none --import-source on -k regex:$2 -s 1 -c 1 -o "$OUT/prof_$1" python tools/ncu_case.py d488 1920 1080 16 50 $3 > "$OUT/ncu_full_$1.log" 2>&1
  echo "ncu full $1 exit $?"
done
cat "$OUT/ncu_case_plain.log"; ls -la "$OUT"
