#!/bin/bash
# A/B of compile-time variants of the render kernel + the cull-loop microbenchmarks (one GPU).
set -u
TAG=${1:-var}; shift
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
python tools/variants.py "$@" > "$OUT/variants.log" 2>&1; echo "variants exit $?"
cat "$OUT/variants.log"
if [ -n "${RT_MICRO:-}" ]; then
  for m in cull_rays mix_variants; do
    nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench/$m tools/microbench/$m.cu && ./tools/microbench/$m > "$OUT/$m.log" 2>&1
    cat "$OUT/$m.log"
  done
fi
