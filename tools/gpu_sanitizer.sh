#!/bin/bash
# one compute-sanitizer tool per gpurun call (B200_PROFILING.md): tools/gpu_sanitizer.sh memcheck|racecheck
TOOL=${1:-memcheck}
mkdir -p gpurun_out
K='cube5 or square9 or in_a_row or no_alloc_limits'
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -k "$K" -x -q > gpurun_out/sanitizer_plain_$TOOL.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitizer_plain_$TOOL.log; exit 1; }
tail -2 gpurun_out/sanitizer_plain_$TOOL.log
timeout 1200 compute-sanitizer --tool $TOOL --error-exitcode 7 --log-file gpurun_out/sanitizer_$TOOL.txt python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -k "$K" -x -q > gpurun_out/sanitizer_pytest_$TOOL.log 2>&1
echo "compute-sanitizer $TOOL rc=$?"
tail -3 gpurun_out/sanitizer_pytest_$TOOL.log; tail -5 gpurun_out/sanitizer_$TOOL.txt
