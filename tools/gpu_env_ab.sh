#!/bin/bash
# A/B of environment settings on one box: bash tools/gpu_env_ab.sh tag "VAR=1" "VAR=2 OTHER=x" ...  (first = baseline "")
tag=$1; shift
mkdir -p gpurun_out
for r in 1 2; do for e in "" "$@"; do
  env $e python bench.py --steps 40 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.readline())
print('[$e]', round(d['ms_per_step'], 4), {k: round(x['ms'], 4) for k, x in d['roofline']['per_kernel'].items()})"
done; done | tee gpurun_out/${tag}_envab.txt
