#!/bin/bash
# parity tests + a short kernel-only bench line.  usage (under gpurun): bash tools/gpu_quick.sh <tag> [pytest args]
tag=${1:-q}; shift
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q "$@" > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?"
tail -40 gpurun_out/${tag}_pytest.log
timeout 300 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"
tail -5 gpurun_out/${tag}_bench.err
python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${tag}_bench.json").readline())
    print("ms_per_step", round(d["ms_per_step"], 4), {k: round(x["ms"], 4) for k, x in d["roofline"]["per_kernel"].items()})
except Exception as e:
    print("no bench line", e)
PY
