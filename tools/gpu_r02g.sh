#!/bin/bash
# Round 2 (1 GPU): cull-loop scheduling microbenchmark; schedule tests after the 32-byte record groups; HBM traffic re-capture.
set -u
OUT=gpurun_out/r02g; mkdir -p "$OUT"
tools/microbench/cull_sched > "$OUT/cull_sched.txt" 2>&1; cat "$OUT/cull_sched.txt"
timeout 900 python -m pytest tests -m gpu -x -q -rs -k "two_stage or batches or overwrite or deterministic or chunked or several_lights or path_for_path or asynchronous or larger_than or ragged or metal" > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?"; tail -4 "$OUT/pytest_gpu.log"
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:rt_ --csv \
      --log-file "$OUT/traffic_c3_brute.csv" python tools/ncu_case.py d488 1920 1080 256 50 0 > "$OUT/ncu_traffic_brute.log" 2>&1; echo "ncu traffic exit $?"
RT_PROBE_FLAGS=0 python tools/perf_probe.py > "$OUT/probe.log" 2>&1; cat "$OUT/probe.log"
