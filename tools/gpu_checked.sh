#!/bin/bash
# The GPU suite against the range-checked build (B2F_CHECKED: stands in for compute-sanitizer, closed on this pool).
mkdir -p gpurun_out
export B2F_LIBRARY=$PWD/arcanefem_b200/lib/libb200fem_checked.so
timeout 1500 python -m pytest tests/test_gpu_checked_build.py tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_gpu_heat_steps.py tests/test_gpu_pattern_layout.py tests/test_gpu_reorder.py tests/test_zz_no_alloc.py -x -q > gpurun_out/pytest_checked.log 2>&1; echo "checked suite rc=$?"; tail -3 gpurun_out/pytest_checked.log
timeout 600 python bench.py --steps 3 --warmup 3 --cpu-cells 100000 > gpurun_out/bench_checked.json 2> gpurun_out/bench_checked.err; python -c "
import json; d=json.load(open('gpurun_out/bench_checked.json')); print('checked build at full size:', d['ms_per_step'], 'ms, columns_not_found', d['config'].get('columns_not_found'), d['parity_sample']['max_rel'])"
