#!/bin/bash
mkdir -p gpurun_out
for lib in base d2post base d2post; do
  if [ "$lib" = "base" ]; then unset B2F_LIBRARY; else export B2F_LIBRARY=$PWD/arcanefem_b200/lib/libb200fem_$lib.so; fi
  timeout 300 python bench.py --steps 20 --warmup 3 --workload elasticity3d --cpu-cells 200000 --no-parity > gpurun_out/bench_r02ag_$lib.json 2> gpurun_out/bench_r02ag_$lib.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_r02ag_$lib.json')); print('$lib elasticity3d', d['ms_per_step'], d['value'], d['roofline']['frac'], d['config']['values_checksum'])"
done
B2F_LIBRARY=$PWD/arcanefem_b200/lib/libb200fem_d2post.so timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_gpu_pattern_layout.py tests/test_gpu_reorder.py tests/test_gpu_zz_fullsize.py -x -q -k "elasticity or 3-" > gpurun_out/pytest_var_d2post.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_var_d2post.log
