#!/bin/bash
# tools/gpu_profile.sh <tag> [kernel-regex] : parity tests, bench line, ncu launch list and one full capture of the top kernel
TAG=${1:-p}; KREGEX=${2:-k_node_fan}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
tail -4 gpurun_out/pytest_gpu_$TAG.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_$TAG.err
cut -c1-300 gpurun_out/bench_$TAG.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 2 --warmup 3 --cpu-cells 100000 > gpurun_out/ncu_launch_$TAG.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$KREGEX --launch-skip 3 -c 1 -f -o gpurun_out/top_$TAG python bench.py --steps 2 --warmup 3 --cpu-cells 100000 > gpurun_out/ncu_full_$TAG.log 2>&1; echo "ncu full rc=$?"
