#!/bin/bash
# new default fan kernel (early row loads, address chain ahead, bulk-store write-out): whole GPU suite + every workload line
TAG=${1:-r02o}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/pytest_gpu_$TAG.log
for w in poisson3d elasticity3d heat3d heat2d poisson2d; do
  timeout 600 python bench.py --steps 20 --warmup 3 --workload $w --cpu-cells 3000000 > gpurun_out/bench_${TAG}_$w.json 2> gpurun_out/bench_${TAG}_$w.err; echo "bench $w rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/bench_${TAG}_$w.json')); print('$w', d['ms_per_step'], d['value'], d['roofline']['frac'], d['e2e']['ms_per_step'] if d.get('e2e') else None, d.get('parity_sample'))"
done
