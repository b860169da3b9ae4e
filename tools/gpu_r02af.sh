#!/bin/bash
mkdir -p gpurun_out
for lib in base cd2 base cd2; do
  if [ "$lib" = "base" ]; then unset B2F_LIBRARY; else export B2F_LIBRARY=$PWD/arcanefem_b200/lib/libb200fem_$lib.so; fi
  for w in heat3d heat2d; do
    timeout 300 python bench.py --steps 20 --warmup 3 --workload $w --cpu-cells 200000 --no-parity > gpurun_out/bench_r02af_${w}_$lib.json 2> gpurun_out/bench_r02af_${w}_$lib.err
    python -c "
import json; d=json.load(open('gpurun_out/bench_r02af_${w}_$lib.json')); print('$lib $w', d['ms_per_step'], d['value'], d['roofline']['frac'], d['config']['values_checksum'])"
  done
done
B2F_LIBRARY=$PWD/arcanefem_b200/lib/libb200fem_cd2.so timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_heat_steps.py tests/test_gpu_golden.py tests/test_gpu_zz_fullsize.py -x -q -k "heat" > gpurun_out/pytest_var_cd2.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_var_cd2.log
