#!/bin/bash
# per-launch device times of one bench run (ncu launch list).  usage: bash tools/gpu_launches.sh tag [ENV=VAL ...]
tag=$1; shift
mkdir -p gpurun_out
env "$@" python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/${tag}_plain.log 2>&1 &&
env "$@" ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/${tag}_ncu1.log 2>&1
echo "exit $?"
