#!/bin/bash
# no-alloc entry point alone in a fresh process, then the whole GPU suite (it runs last there), then heat row-sum A/B
TAG=${1:-r02k}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/pytest_noalloc_$TAG.log 2>&1; echo "alone rc=$?"; tail -3 gpurun_out/pytest_noalloc_$TAG.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/pytest_gpu_$TAG.log
for lib in base heatsum base heatsum; do
  if [ "$lib" = "base" ]; then unset B2F_LIBRARY; else export B2F_LIBRARY=$PWD/arcanefem_b200/lib/libb200fem_$lib.so; fi
  for w in heat3d heat2d; do
    timeout 300 python bench.py --steps 20 --warmup 3 --workload $w --cpu-cells 200000 --no-parity > gpurun_out/bench_${TAG}_${w}_$lib.json 2> gpurun_out/bench_${TAG}_${w}_$lib.err
    python -c "
import json; d=json.load(open('gpurun_out/bench_${TAG}_${w}_$lib.json')); print('$lib $w', d['ms_per_step'], d['value'], d['roofline']['frac'])"
  done
done
