#!/bin/bash
# bench lines of the other BASELINE configs (parity-test cases; the headline line is cfg3) + the cfg5 threshold sweep
mkdir -p gpurun_out
: > gpurun_out/r02_configs.jsonl
for c in cfg2 cfg4 cfg5; do python bench.py --config $c --steps 10 --warmup 3 --no-cpu-baseline >> gpurun_out/r02_configs.jsonl; done
for q in 0 10 20 30 40; do python bench.py --config cfg5 --min-qual $q --steps 10 --warmup 3 --no-cpu-baseline --no-e2e >> gpurun_out/r02_configs.jsonl; done
python - <<'PY'
import json
for l in open("gpurun_out/r02_configs.jsonl"):
    d = json.loads(l)
    print(d["config"]["config"], "min_qual", d["config"]["min_qual"], "ms/step %.3f" % d["ms_per_step"], "value %.0fM" % (d["value"] / 1e6),
          "e2e", d["e2e"] and "%.1fM" % (d["e2e"]["value"] / 1e6), d["roofline"]["kernel"], "%.3f" % d["roofline"]["frac"],
          "und %.3f" % d["workload_stats"]["undetermined_frac"], {k: round(v["ms"], 3) for k, v in d["roofline"]["per_kernel"].items()})
PY
