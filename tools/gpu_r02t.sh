#!/bin/bash
# micro-variants on top of the new default: 4 partial row sums, 64-row tiles, first START peeled (its two gathers issued together)
mkdir -p gpurun_out
BENCH_ARGS="--no-parity" bash tools/gpu_variants.sh base rs4 rows64 peel peelrs4 base rs4 rows64 peel peelrs4
B2F_LIBRARY=$PWD/arcanefem_b200/lib/libb200fem_peelrs4.so timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_gpu_heat_steps.py tests/test_gpu_pattern_layout.py tests/test_gpu_reorder.py -x -q > gpurun_out/pytest_var_peelrs4.log 2>&1; echo "pytest peelrs4 rc=$?"; tail -2 gpurun_out/pytest_var_peelrs4.log
