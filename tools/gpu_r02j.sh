#!/bin/bash
TAG=${1:-r02j}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_heat_steps.py tests/test_gpu_golden.py tests/test_gpu_pattern_layout.py tests/test_gpu_reorder.py -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
tail -4 gpurun_out/pytest_gpu_$TAG.log
BENCH_ARGS="--no-parity" bash tools/gpu_variants.sh base
for w in heat3d heat2d; do
  timeout 600 python bench.py --steps 10 --warmup 3 --workload $w --cpu-cells 3000000 > gpurun_out/bench_${TAG}_$w.json 2> gpurun_out/bench_${TAG}_$w.err; echo "bench $w rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/bench_${TAG}_$w.json')); print('$w', d['ms_per_step'], d['value'], d['roofline']['frac'])"
done
