#!/bin/bash
# first-visit marks of the plan (no clearing pass, first loads start from zero): parity first, then A/B against the same build without them
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r02q.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_r02q.log
BENCH_ARGS="--no-parity" bash tools/gpu_variants.sh base nofirst base nofirst
python bench.py --steps 10 --warmup 3 --cpu-cells 200000 > gpurun_out/bench_r02q.json 2> gpurun_out/bench_r02q.err; python -c "
import json; d=json.load(open('gpurun_out/bench_r02q.json')); print(d['ms_per_step'], d['roofline'], d['parity_sample'], d['config'].get('sparsity_build_ms'))"
