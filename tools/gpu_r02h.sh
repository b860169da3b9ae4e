#!/bin/bash
TAG=${1:-r02h}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_reorder.py -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
tail -4 gpurun_out/pytest_gpu_$TAG.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; cut -c1-260 gpurun_out/bench_$TAG.json
timeout 300 ./tools/experiments/tile2_proto 203 > gpurun_out/tile2_proto_$TAG.jsonl 2>&1; echo "tile2 rc=$?"; cat gpurun_out/tile2_proto_$TAG.jsonl
timeout 1500 python tools/bench_unstructured.py --points 2000000 > gpurun_out/unstructured_$TAG.jsonl 2> gpurun_out/unstructured_$TAG.err; echo "unstructured rc=$?"; tail -3 gpurun_out/unstructured_$TAG.err
cut -c1-200 gpurun_out/unstructured_$TAG.jsonl
