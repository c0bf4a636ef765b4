#!/bin/bash
# Final sanity of a round: whole GPU suite, smoke, probe of both query modes, one default bench line.
set -u
TAG=${1:-final}; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
python -m pytest tests -m gpu -x -q > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?"; tail -3 "$OUT/pytest_gpu.log"
python -c "import __graft_entry__ as g; g.smoke()" > "$OUT/smoke.log" 2>&1; echo "smoke exit $?"; tail -1 "$OUT/smoke.log"
python tools/perf_probe.py > "$OUT/probe_brute.log" 2>&1; cat "$OUT/probe_brute.log"
RT_PROBE_FLAGS=8 python tools/perf_probe.py > "$OUT/probe_accel.log" 2>&1; cat "$OUT/probe_accel.log"
python bench.py > "$OUT/bench.json" 2> "$OUT/bench.err"; echo "bench exit $?"; tail -c 1200 "$OUT/bench.json"
