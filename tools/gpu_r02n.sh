#!/bin/bash
# L1-data-pipe variants of the fan kernel: bulk-store write-out (bs), 16+8-byte gathers (v2), both; parity suite on the candidates
mkdir -p gpurun_out
BENCH_ARGS="--no-parity" bash tools/gpu_variants.sh base aae bs v2 v2bs base aae bs v2 v2bs
for n in bs v2bs; do
  B2F_LIBRARY=$PWD/arcanefem_b200/lib/libb200fem_$n.so timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_gpu_heat_steps.py tests/test_gpu_pattern_layout.py tests/test_gpu_reorder.py -x -q > gpurun_out/pytest_var_$n.log 2>&1; echo "pytest $n rc=$?"; tail -2 gpurun_out/pytest_var_$n.log
done
