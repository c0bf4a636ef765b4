#!/bin/bash
# Round 2, first pass (gpurun --gpus 2): GPU suite with no skipped multi-GPU test, N=2 bench with parity_n, clean CPU arm + A/B.
set -u
OUT=gpurun_out/r02a; mkdir -p "$OUT"
nvidia-smi -L > "$OUT/smi.txt" 2>&1; nproc >> "$OUT/smi.txt"
python -m pytest tests -m gpu -x -q -rs > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?" | tee -a "$OUT/pytest_gpu.log"; tail -4 "$OUT/pytest_gpu.log"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > "$OUT/bench_n2.json" 2> "$OUT/bench_n2.err"; echo "bench n2 exit $?"
tail -c 1500 "$OUT/bench_n2.json"; tail -5 "$OUT/bench_n2.err"
python bench.py --impl reference --steps 4 --warmup 1 > "$OUT/bench_reference.json" 2> "$OUT/bench_reference.err"; echo "reference exit $?"; cat "$OUT/bench_reference.json"
python tools/cpu_baseline_ab.py 4 > "$OUT/cpu_baseline_ab.txt" 2>&1; cat "$OUT/cpu_baseline_ab.txt"
python bench.py > "$OUT/bench_n1.json" 2> "$OUT/bench_n1.err"; echo "bench n1 exit $?"; tail -c 1200 "$OUT/bench_n1.json"
