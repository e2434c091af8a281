#!/bin/bash
# library variants x environment: bash tools/gpu_ab_env.sh tag "ENV..." sfx1 sfx2 ...   (A = in-tree build, always with and without ENV)
tag=$1; envs=$2; shift 2
mkdir -p gpurun_out
for r in 1 2; do for v in A "$@"; do for e in "" "$envs"; do
  if [ $v = A ]; then unset QUADE_LIB_PATH; else export QUADE_LIB_PATH=$PWD/quade_b200/libquade_b200_$v.so; fi
  env $e python bench.py --steps 40 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.readline())
print('$v [$e]', round(d['ms_per_step'], 4), {k: round(x['ms'], 4) for k, x in d['roofline']['per_kernel'].items() if k in ('frame', 'compact')})"
done; done; done | tee gpurun_out/${tag}_abenv.txt
