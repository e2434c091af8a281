#!/bin/bash
# baseline of round 2 after the prune: parity tests, bench line, host topology of the box
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02a_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r02a_pytest.log
python bench.py --steps 20 --warmup 3 > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err; echo "bench exit $?"; cat gpurun_out/r02a_bench.json
{ nproc; lscpu | head -25; numactl -H 2>&1 | head; nvidia-smi topo -m 2>&1 | head -20; for d in /sys/bus/pci/devices/*; do if [ "$(cat $d/class 2>/dev/null)" = "0x030200" ]; then echo "$d numa_node=$(cat $d/numa_node) local_cpulist=$(cat $d/local_cpulist)"; fi; done; python2.7 -c "import numpy" 2>&1 | head -2; free -g | head -3; } > gpurun_out/r02a_host.txt 2>&1
cat gpurun_out/r02a_host.txt
