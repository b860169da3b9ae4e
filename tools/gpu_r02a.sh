#!/bin/bash
# round 2, first GPU pass: parity tests, the bench line of every workload, plan statistics
TAG=${1:-r02a}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
tail -6 gpurun_out/pytest_gpu_$TAG.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_$TAG.err
cut -c1-600 gpurun_out/bench_$TAG.json
for w in elasticity3d heat3d heat2d poisson2d; do
  timeout 600 python bench.py --steps 10 --warmup 3 --workload $w --cpu-cells 3000000 > gpurun_out/bench_${TAG}_$w.json 2> gpurun_out/bench_${TAG}_$w.err; echo "bench $w rc=$?"; tail -2 gpurun_out/bench_${TAG}_$w.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_${TAG}_$w.json')); print('$w', d['ms_per_step'], d['value'], d['roofline']['frac'], d['e2e']['ms_per_step'], d['cpu_baseline']['value'])"
done
timeout 300 python tools/fan_plan_stats.py 203 > gpurun_out/fan_plan_stats_$TAG.json 2> gpurun_out/fan_plan_stats_$TAG.err; cat gpurun_out/fan_plan_stats_$TAG.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "ref rc=$?"; cut -c1-400 gpurun_out/bench_ref_$TAG.json
