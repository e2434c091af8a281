#!/bin/bash
# The parity tests and the generated-FASTQ fuzz on the QUADE_CHECKS build (tools/build_variant.sh checks -DQUADE_CHECKS=1):
# every shared / global store index and window read of k_frame, k_tilescan, k_classify and k_compact is tested
# against its buffer on the device; quade_ctx_destroy prints the tally.  compute-sanitizer is closed on this pool.
tag=${1:-r02}
mkdir -p gpurun_out
export QUADE_LIB_PATH=$PWD/quade_b200/libquade_b200_checks.so
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py tests/test_gpu_host.py -m gpu -q -s > gpurun_out/${tag}_checks_pytest.log 2>&1
echo "pytest exit $?"
grep -E "passed|failed" gpurun_out/${tag}_checks_pytest.log | tail -1
{ echo "contexts destroyed: $(grep -c 'QUADE_CHECKS:' gpurun_out/${tag}_checks_pytest.log)"
  echo "tallies:"; grep -o 'QUADE_CHECKS: [0-9]* violations (first code [0-9]*)' gpurun_out/${tag}_checks_pytest.log | sort | uniq -c; } | tee gpurun_out/${tag}_checks.txt
