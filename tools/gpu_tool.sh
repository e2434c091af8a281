#!/bin/bash
# whole-tool timings (files -> files) on the GPU box.  usage: bash tools/gpu_tool.sh tag ["pairs chunks plain|gz level dir" ...]
tag=${1:-tool}; shift
mkdir -p gpurun_out
if [ $# -eq 0 ]; then set -- "1000000 4 plain 0 /dev/shm" "1000000 4 plain 1 /dev/shm" "1000000 4 gz 1 /dev/shm" "1000000 4 plain 0 /tmp"; fi
{
df -h /dev/shm /tmp | head -5
for args in "$@"; do
  QUADE_PROFILE=1 timeout 900 python tools/tool_e2e.py $args 2>&1 | grep -E "^inputs|^tool|^profile|Error|error|Traceback"
done
} | tee gpurun_out/${tag}_e2e.txt
