#!/bin/bash
# round-end pass on one GPU: whole GPU suite, smoke(), the bench line as the driver runs it (both arms), every workload line
TAG=${1:-r02z}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/pytest_gpu_$TAG.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke_$TAG.log
timeout 900 python bench.py --impl reference > gpurun_out/bench_${TAG}_reference.json 2> gpurun_out/bench_${TAG}_reference.err; echo "reference rc=$?"; cut -c1-300 gpurun_out/bench_${TAG}_reference.json
timeout 900 python bench.py > gpurun_out/bench_${TAG}_poisson3d.json 2> gpurun_out/bench_${TAG}_poisson3d.err; echo "bench rc=$?"
for w in elasticity3d heat3d heat2d poisson2d; do
  timeout 600 python bench.py --steps 20 --warmup 3 --workload $w --cpu-cells 3000000 > gpurun_out/bench_${TAG}_$w.json 2> gpurun_out/bench_${TAG}_$w.err; echo "bench $w rc=$?"
done
for w in poisson3d elasticity3d heat3d heat2d poisson2d; do
  python -c "
import json; d=json.load(open('gpurun_out/bench_${TAG}_$w.json')); print('$w', d['ms_per_step'], d['value'], d['roofline']['frac'], d['e2e']['ms_per_step'], d['cpu_baseline']['value'], d['cpu_baseline']['cores'], d.get('gpu_launches'), d['clocks'], (d.get('parity_sample') or {}).get('max_rel'))"
done
