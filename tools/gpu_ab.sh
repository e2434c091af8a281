#!/bin/bash
# A/B/C... of library builds on one box: A = in-tree libquade_b200.so, the others = quade_b200/libquade_b200_<sfx>.so
# (tools/build_variant.sh).  Kernel-only bench lines alternate over two rounds; the parity tests run on every variant.
# usage (under gpurun): bash tools/gpu_ab.sh tag sfx1 [sfx2 ...]
tag=$1; shift
mkdir -p gpurun_out
for r in $(seq ${ROUNDS:-2}); do for v in A "$@"; do
  if [ $v = A ]; then unset QUADE_LIB_PATH; else export QUADE_LIB_PATH=$PWD/quade_b200/libquade_b200_$v.so; fi
  python bench.py --steps 40 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.readline())
print('$v', round(d['ms_per_step'], 4), {k: round(x['ms'], 4) for k, x in d['roofline']['per_kernel'].items()})"
done; done | tee gpurun_out/${tag}_ab.txt
[ -n "$SKIP_PARITY" ] && exit 0
for v in "$@"; do
  QUADE_LIB_PATH=$PWD/quade_b200/libquade_b200_$v.so timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2 | sed "s/^/$v: /" | tee -a gpurun_out/${tag}_ab.txt
done
