#!/bin/bash
TAG=${1:-r02e}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_reorder.py tests/test_gpu_pattern_layout.py tests/test_cpp_shim.py tests/test_gpu_parity.py -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
tail -25 gpurun_out/pytest_gpu_$TAG.log
timeout 1500 python tools/bench_unstructured.py --points 2000000 > gpurun_out/unstructured_$TAG.jsonl 2> gpurun_out/unstructured_$TAG.err; echo "unstructured rc=$?"; tail -3 gpurun_out/unstructured_$TAG.err
cut -c1-420 gpurun_out/unstructured_$TAG.jsonl
