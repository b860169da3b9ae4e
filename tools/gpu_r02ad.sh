#!/bin/bash
# per-warp bulk copies + mbarriers in the prologue of the fan kernel (wcopy) against the default; parity on the variant
mkdir -p gpurun_out
BENCH_ARGS="--no-parity" bash tools/gpu_variants.sh base wcopy base wcopy
for lib in base wcopy; do
  if [ "$lib" = "base" ]; then unset B2F_LIBRARY; else export B2F_LIBRARY=$PWD/arcanefem_b200/lib/libb200fem_$lib.so; fi
  for w in elasticity3d heat3d heat2d poisson2d; do
    timeout 300 python bench.py --steps 20 --warmup 3 --workload $w --cpu-cells 200000 --no-parity > gpurun_out/bench_r02ad_${w}_$lib.json 2> gpurun_out/bench_r02ad_${w}_$lib.err
    python -c "
import json; d=json.load(open('gpurun_out/bench_r02ad_${w}_$lib.json')); print('$lib $w', d['ms_per_step'], d['value'], d['roofline']['frac'])"
  done
done
B2F_LIBRARY=$PWD/arcanefem_b200/lib/libb200fem_wcopy.so timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_gpu_heat_steps.py tests/test_gpu_pattern_layout.py tests/test_gpu_reorder.py tests/test_user_operator.py -x -q > gpurun_out/pytest_var_wcopy.log 2>&1; echo "pytest wcopy rc=$?"; tail -2 gpurun_out/pytest_var_wcopy.log
