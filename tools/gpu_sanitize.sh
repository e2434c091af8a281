#!/bin/bash
# ONE compute-sanitizer tool per gpurun call (B200_PROFILING.md) over the small-input parity tests.
# usage (under gpurun): bash tools/gpu_sanitize.sh memcheck|racecheck|initcheck|synccheck [tag]
tool=$1; tag=${2:-r02}
mkdir -p gpurun_out
sel='adversarial or tiny or eof or blank_header or error_codes or golden_dataset'
# the same selection without the tool first: it must pass on its own
QUADE_FUZZ_EXAMPLES=25 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py -m gpu -q -x -k "$sel or generated" > gpurun_out/${tag}_san_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${tag}_san_plain.log; exit 1; }
tail -1 gpurun_out/${tag}_san_plain.log
QUADE_FUZZ_EXAMPLES=25 timeout 1500 compute-sanitizer --tool $tool --error-exitcode 77 --launch-timeout 0 --log-file gpurun_out/${tag}_san_${tool}.log \
    python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py -m gpu -q -x -k "$sel or generated" > gpurun_out/${tag}_san_${tool}_pytest.log 2>&1
echo "compute-sanitizer $tool exit $?"
tail -2 gpurun_out/${tag}_san_${tool}_pytest.log
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|Error:|Hazard" gpurun_out/${tag}_san_${tool}.log | sort | uniq -c | head -20
