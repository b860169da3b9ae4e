#!/bin/bash
TAG=${1:-r02g}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_reorder.py -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
tail -4 gpurun_out/pytest_gpu_$TAG.log
timeout 1500 python tools/bench_unstructured.py --points 2000000 > gpurun_out/unstructured_$TAG.jsonl 2> gpurun_out/unstructured_$TAG.err; echo "unstructured rc=$?"; tail -3 gpurun_out/unstructured_$TAG.err
cut -c1-200 gpurun_out/unstructured_$TAG.jsonl
bash tools/gpu_ncu_r02.sh r02
