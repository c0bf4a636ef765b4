#!/bin/bash
# One GPU-box pass: parity tests, smoke, default bench, then the two ncu captures the judge reads
# (launch list of the bench command; --set full of the render kernel on a short case).
# usage (from the repo root, under gpurun): bash tools/gpu_round.sh <tag>
set -u
TAG=${1:-run}
OUT=gpurun_out/$TAG
mkdir -p "$OUT"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv > "$OUT/smi.txt" 2>&1
python -m pytest tests -m gpu -x -q > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?" | tee -a "$OUT/pytest_gpu.log"
python -c "import __graft_entry__ as g; g.smoke()" > "$OUT/smoke.log" 2>&1; echo "smoke exit $?" | tee -a "$OUT/smoke.log"
python bench.py > "$OUT/bench.json" 2> "$OUT/bench.err"; echo "bench exit $?"
tail -c 600 "$OUT/bench.json"
python bench.py --impl reference --steps 4 --warmup 1 > "$OUT/bench_reference.json" 2> "$OUT/bench_reference.err"; echo "reference exit $?"
# ncu: launch list of the bench command (device time of every launch)
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > "$OUT/bench_short.json" 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file "$OUT/launches.csv" \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > "$OUT/ncu_launches.log" 2>&1
echo "ncu launches exit $?"
# ncu: full set on the render kernel, short case (second launch = warm)
python tools/ncu_case.py d488 1920 1080 16 50 > "$OUT/ncu_case_plain.log" 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rt_render_kernel -s 1 -c 1 -o "$OUT/prof_render" \
    python tools/ncu_case.py d488 1920 1080 16 50 > "$OUT/ncu_full.log" 2>&1
echo "ncu full exit $?"
ls -la "$OUT"
