#!/bin/bash
# tools/gpu_final.sh <tag>: new tests first (results survive a clamp of the call), then the bench line, then the rest of the GPU suite
TAG=${1:-final}
mkdir -p gpurun_out
NEW="tests/test_gpu_zz_fullsize.py tests/test_user_operator.py tests/test_gpu_heat_steps.py"
timeout 420 python -m pytest $NEW -m gpu -q --tb=short > gpurun_out/pytest_new_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_new_$TAG.log
tail -25 gpurun_out/pytest_new_$TAG.log
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_$TAG.err
cut -c1-600 gpurun_out/bench_$TAG.json
IGN=""; for f in $NEW; do IGN="$IGN --ignore=$f"; done
timeout 900 python -m pytest tests -m gpu -x -q $IGN > gpurun_out/pytest_rest_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_rest_$TAG.log
tail -6 gpurun_out/pytest_rest_$TAG.log
