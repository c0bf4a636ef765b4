#!/bin/bash
# Last check of HEAD on one GPU: the no-alloc test alone in a fresh process, the whole GPU suite, smoke, the default bench line.
set -u
OUT=gpurun_out/${1:-r02y}; mkdir -p "$OUT"
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > "$OUT/pytest_no_alloc_alone.log" 2>&1; echo "no_alloc alone exit $?" | tee -a "$OUT/pytest_no_alloc_alone.log"
timeout 1500 python -m pytest tests -m gpu -x -q -rs > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?" | tee -a "$OUT/pytest_gpu.log"; tail -5 "$OUT/pytest_gpu.log"
python -c "import __graft_entry__ as g; g.smoke()" > "$OUT/smoke.log" 2>&1; echo "smoke exit $?" | tee -a "$OUT/smoke.log"; tail -2 "$OUT/smoke.log"
python bench.py > "$OUT/bench.json" 2> "$OUT/bench.err"; echo "bench exit $?"; tail -3 "$OUT/bench.err"
python - <<PY
import json
d=json.loads(open("$OUT/bench.json").read().strip().splitlines()[-1])
print("c3 %.4e e2e %.4e ms/step %.1f frac %.3f"%(d["value"],d["e2e"]["value"],d["ms_per_step"],d["roofline"]["frac"]), {k:"%.3e"%v["brute"]["value"] for k,v in d["configs"].items()}, {k:round(v["rel_rmse"],4) for k,v in d["rmse"].items() if k.startswith("c")})
PY
