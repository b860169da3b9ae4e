#!/bin/bash
# heat: unconditional per-cell load (no select at the issue point) + uniform-conductivity instantiation; depth-2 coordinate gather for HeatTet (d2)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_heat_steps.py tests/test_gpu_golden.py tests/test_gpu_pattern_layout.py tests/test_gpu_reorder.py -x -q > gpurun_out/pytest_r02y.log 2>&1; echo "pytest base rc=$?"; tail -2 gpurun_out/pytest_r02y.log
for lib in base d2 base d2; do
  if [ "$lib" = "base" ]; then unset B2F_LIBRARY; else export B2F_LIBRARY=$PWD/arcanefem_b200/lib/libb200fem_$lib.so; fi
  for w in heat3d heat2d; do
    timeout 300 python bench.py --steps 20 --warmup 3 --workload $w --cpu-cells 200000 --no-parity > gpurun_out/bench_r02y_${w}_$lib.json 2> gpurun_out/bench_r02y_${w}_$lib.err
    python -c "
import json; d=json.load(open('gpurun_out/bench_r02y_${w}_$lib.json')); print('$lib $w', d['ms_per_step'], d['value'], d['roofline']['frac'], d['config']['values_checksum'])"
  done
done
B2F_LIBRARY=$PWD/arcanefem_b200/lib/libb200fem_d2.so timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_heat_steps.py tests/test_gpu_golden.py -x -q -k "heat" > gpurun_out/pytest_var_d2.log 2>&1; echo "pytest d2 rc=$?"; tail -2 gpurun_out/pytest_var_d2.log
