#!/bin/bash
# round 2: full GPU test-suite after the layout / adapter work + the bench line
TAG=${1:-r02d}
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
tail -30 gpurun_out/pytest_gpu_$TAG.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_$TAG.err
cut -c1-300 gpurun_out/bench_$TAG.json
