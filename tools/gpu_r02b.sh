#!/bin/bash
# Round 2, second pass (1 GPU): two-stage schedule -- parity suite, then kernel-time A/B against the round-1 one-kernel schedule.
set -u
OUT=gpurun_out/r02b; mkdir -p "$OUT"
timeout 900 python -m pytest tests -m gpu -x -q -rs > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?" | tee -a "$OUT/pytest_gpu.log"; tail -15 "$OUT/pytest_gpu.log"
for f in 0 0x200 8 0x208; do
  echo "=== flags $f" | tee -a "$OUT/probe.log"
  RT_PROBE_FLAGS=$f timeout 600 python tools/perf_probe.py >> "$OUT/probe.log" 2>&1
done
cat "$OUT/probe.log"
