#!/bin/bash
# full ncu capture of kernels matching $1 (regex) under env $3.., report to gpurun_out/$2.ncu-rep
# usage (under gpurun): bash tools/gpu_ncu.sh 'k_frame|k_compact' tag [ENV=VAL ...]
pat=$1; tag=$2; shift 2
mkdir -p gpurun_out
env "$@" python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/${tag}_plain.log 2>&1 &&
env "$@" ncu --set full --clock-control none --import-source on -k regex:"$pat" -s ${NCU_SKIP:-40} -c ${NCU_COUNT:-10} \
    -f -o gpurun_out/${tag} python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/${tag}_ncu.log 2>&1
echo "ncu exit $?"; tail -2 gpurun_out/${tag}_ncu.log
