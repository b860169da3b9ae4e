#!/bin/bash
# SoA coordinate copy experiment (conversion before every launch, inside the timed step) against the default; parity on the variant
mkdir -p gpurun_out
BENCH_ARGS="--no-parity" bash tools/gpu_variants.sh base soa base soa
B2F_LIBRARY=$PWD/arcanefem_b200/lib/libb200fem_soa.so timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_gpu_heat_steps.py tests/test_gpu_pattern_layout.py tests/test_gpu_reorder.py -x -q > gpurun_out/pytest_var_soa.log 2>&1; echo "pytest soa rc=$?"; tail -2 gpurun_out/pytest_var_soa.log
for w in elasticity3d heat3d heat2d poisson2d; do
  B2F_LIBRARY=$PWD/arcanefem_b200/lib/libb200fem_soa.so timeout 300 python bench.py --steps 20 --warmup 3 --workload $w --cpu-cells 200000 --no-parity > gpurun_out/bench_soa_$w.json 2> gpurun_out/bench_soa_$w.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_soa_$w.json')); print('soa $w', d['ms_per_step'], d['value'], d['roofline']['frac'])"
done
