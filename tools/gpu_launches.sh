#!/bin/bash
# per-launch device times of one bench run (ncu launch list).  usage: bash tools/gpu_launches.sh tag [ENV=VAL ...]
tag=$1; shift
mkdir -p gpurun_out
env "$@" python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/${tag}_plain.log 2>&1 &&
env "$@" ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/${tag}_ncu1.log 2>&1
echo "exit $?"
python - <<PY
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/${tag}_launches.csv")) if len(r) > 10 and r[0].isdigit()]
agg = collections.OrderedDict()
for r in rows[-60:]:
    name = r[4].split("(")[0].split("::")[-1]
    agg.setdefault(name, []).append(float(r[-1].replace(",", "")))
for k, v in agg.items():
    print(k, len(v), "launches, median", sorted(v)[len(v) // 2], r[-2] if False else "")
PY
