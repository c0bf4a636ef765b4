#!/bin/bash
# One GPU-box pass: parity tests, smoke, default bench + reference arm, the other workloads / query modes, then the ncu
# captures the judge reads (launch list of the bench command; --set full of both render kernels on a short case).
# usage (from the repo root, under gpurun): bash tools/gpu_round.sh <tag>
set -u
TAG=${1:-run}
OUT=gpurun_out/$TAG
mkdir -p "$OUT"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv > "$OUT/smi.txt" 2>&1
python -m pytest tests -m gpu -x -q > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?" | tee -a "$OUT/pytest_gpu.log"
tail -3 "$OUT/pytest_gpu.log"
python -c "import __graft_entry__ as g; g.smoke()" > "$OUT/smoke.log" 2>&1; echo "smoke exit $?" | tee -a "$OUT/smoke.log"
python bench.py > "$OUT/bench.json" 2> "$OUT/bench.err"; echo "bench exit $?"
tail -c 900 "$OUT/bench.json"
python bench.py --impl reference --steps 4 --warmup 1 > "$OUT/bench_reference.json" 2> "$OUT/bench_reference.err"; echo "reference exit $?"
for spec in "c3 accel" "c2 brute" "c5 brute" "c5 accel" "c1 brute" "c4 brute"; do
  set -- $spec
  python bench.py --workload $1 --query $2 --steps 3 --warmup 3 --cpu-seconds 6 > "$OUT/bench_$1_$2.json" 2> "$OUT/bench_$1_$2.err"; echo "bench $1 $2 exit $?"
done
# ncu: launch list of the bench command (device time of every launch)
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > "$OUT/bench_short.json" 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file "$OUT/launches.csv" \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > "$OUT/ncu_launches.log" 2>&1
echo "ncu launches exit $?"
# ncu: HBM traffic of one render-kernel launch of the default bench step (c3: 256 spp, depth 50) in both query modes -> roofline.traffic
for q in "brute 0" "accel 8"; do
  set -- $q
  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:rt_render_kernel -s 1 -c 1 --csv \
      --log-file "$OUT/traffic_c3_$1.csv" python tools/ncu_case.py d488 1920 1080 256 50 $2 > "$OUT/ncu_traffic_$1.log" 2>&1
  echo "ncu traffic $1 exit $?"
done
# ncu: full set on the render kernel, short case (second launch = warm): brute force, then RT_FLAG_ACCEL on both scenes
python tools/ncu_case.py d488 1920 1080 16 50 > "$OUT/ncu_case_plain.log" 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rt_render_kernel -s 1 -c 1 -o "$OUT/prof_render" \
    python tools/ncu_case.py d488 1920 1080 16 50 > "$OUT/ncu_full.log" 2>&1
echo "ncu full exit $?"
python tools/ncu_case.py d488 1920 1080 16 50 8 > "$OUT/ncu_case_accel_plain.log" 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rt_render_kernel -s 1 -c 1 -o "$OUT/prof_render_accel_d488" \
    python tools/ncu_case.py d488 1920 1080 16 50 8 > "$OUT/ncu_full_accel.log" 2>&1
echo "ncu full accel exit $?"
python tools/ncu_case.py s4096 1920 1080 8 50 8 > "$OUT/ncu_case_accel_s4096_plain.log" 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rt_render_kernel -s 1 -c 1 -o "$OUT/prof_render_accel_s4096" \
    python tools/ncu_case.py s4096 1920 1080 8 50 8 > "$OUT/ncu_full_accel_s4096.log" 2>&1
echo "ncu full accel s4096 exit $?"
cat "$OUT"/ncu_case*plain.log
ls -la "$OUT"
