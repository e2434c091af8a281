#!/bin/bash
# The N-GPU round (run under `gpurun --gpus N`): host topology, copy-bandwidth probe with and without rank placement,
# multi-GPU parity tests, the scaling bench line (cfg3) and BASELINE config 4 at its stated scale (2 B pairs over 8 GPUs).
N=${1:-8}; tag=${2:-r02_n$N}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
{ nproc; lscpu | grep -E "Model name|Socket|NUMA|Core|Thread"; nvidia-smi topo -m 2>&1 | head -24
  for d in /sys/bus/pci/devices/*; do if [ "$(cat $d/class 2>/dev/null)" = "0x030200" ]; then echo "$d numa_node=$(cat $d/numa_node) local_cpulist=$(cat $d/local_cpulist)"; fi; done
  free -g | head -2; } > gpurun_out/${tag}_host.txt 2>&1
$TR --master-port 29601 tools/probe_pcie.py > gpurun_out/${tag}_pcie.txt 2>gpurun_out/${tag}_pcie.err
PROBE_AFFINITY=1 $TR --master-port 29602 tools/probe_pcie.py >> gpurun_out/${tag}_pcie.txt 2>>gpurun_out/${tag}_pcie.err
cat gpurun_out/${tag}_pcie.txt
timeout 600 python -m pytest tests/test_gpu_host.py tests/test_gpu_parity.py -m gpu -q -k "nccl or allreduce" > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${tag}_pytest.log
$TR --master-port 29603 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"
$TR --master-port 29604 bench.py --gpus $N --steps 20 --warmup 3 --no-bind > gpurun_out/${tag}_bench_nobind.json 2>> gpurun_out/${tag}_bench.err; echo "bench nobind exit $?"
$TR --master-port 29605 bench.py --gpus $N --config cfg4 --steps 125 --warmup 3 --distinct 32 --distinct-gb 60 --slots 3 --no-e2e > gpurun_out/${tag}_bench_cfg4.json 2>> gpurun_out/${tag}_bench.err; echo "bench cfg4 exit $?"
python - <<PY
import json
for f in ("bench", "bench_nobind", "bench_cfg4"):
    try:
        d = json.loads(open("gpurun_out/${tag}_%s.json" % f).readline())
        e = d.get("e2e") or {}
        print(f, "value %.4g pairs/s" % d["value"], "ms/step %.4f" % d["ms_per_step"], "e2e %.4g" % (e.get("value") or 0), "pcie/gpu %.1f GB/s" % (e.get("pcie_GBps_per_gpu") or 0), (e.get("placement") or {}).get("mode"), d["workload_stats"]["distinct_chunks_per_gpu"])
    except Exception as ex:
        print(f, "no line", ex)
PY
tail -5 gpurun_out/${tag}_bench.err
