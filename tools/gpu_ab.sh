#!/bin/bash
# A/B of kernel variants on the GPU box: parity tests + per-kernel timings under each environment setting.
# usage (under gpurun): bash tools/gpu_ab.sh "QUADE_FRAME=3" ["QUADE_FRAME=2" ...]
mkdir -p gpurun_out
for envs in "$@"; do
  echo "=== $envs"
  env $envs timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
  env $envs timeout 300 python tools/probe_steps.py 2>&1 | tail -2
done
