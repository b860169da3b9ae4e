#!/bin/bash
# round 2, second GPU pass: parity of the keep-pivot plan, plan statistics, A/B of the persistent-CTA builds
TAG=${1:-r02b}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
tail -3 gpurun_out/pytest_gpu_$TAG.log
timeout 300 python tools/fan_plan_stats.py 203 > gpurun_out/fan_plan_stats_$TAG.json 2> gpurun_out/fan_plan_stats_$TAG.err; cat gpurun_out/fan_plan_stats_$TAG.json
BENCH_ARGS="--no-parity" bash tools/gpu_variants.sh base nokeep p1 p2 base
B2F_LIBRARY=$PWD/arcanefem_b200/lib/libb200fem_p1.so timeout 600 python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -2
