#!/bin/bash
# A/B of two builds of the library on one box: A = in-tree libquade_b200.so, B = $1 (built with QUADE_LIB_PATH /
# QUADE_BUILD_DEFS, see quade_b200/build.py).  Bench lines alternate A B A B, then the parity tests run on B.
# usage (under gpurun): bash tools/gpu_ab_lib.sh quade_b200/libquade_b200_nl3.so tag
B=$PWD/$1; tag=${2:-ab}
mkdir -p gpurun_out
for r in 1 2; do for v in A B; do
  if [ $v = B ]; then export QUADE_LIB_PATH=$B; else unset QUADE_LIB_PATH; fi
  python bench.py --steps 40 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.readline())
print('$v', round(d['ms_per_step'], 4), {k: round(x['ms'], 4) for k, x in d['roofline']['per_kernel'].items()})"
done; done | tee gpurun_out/${tag}.txt
QUADE_LIB_PATH=$B timeout 150 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2 | tee -a gpurun_out/${tag}.txt
