#!/bin/bash
# GPU pass: parity tests, bench of the default variant (+ optional others), ncu launch list + one full capture.
# usage: tools/gpu_check.sh <tag> <kernel-regex> [extra variants...]
set -u
TAG=${1:-rXX}; KREGEX=${2:-k_node_fan}; shift 2 || true
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
tail -15 gpurun_out/pytest_gpu_$TAG.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_$TAG.err
cat gpurun_out/bench_$TAG.json
for v in "$@"; do
  timeout 400 python bench.py --steps 10 --warmup 3 --variant $v --cpu-cells 1000000 > gpurun_out/bench_${TAG}_v$v.json 2> gpurun_out/bench_${TAG}_v$v.err; echo "bench v$v rc=$?"
  cat gpurun_out/bench_${TAG}_v$v.json
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 2 --warmup 3 --cpu-cells 100000 > gpurun_out/ncu_launch_$TAG.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$KREGEX --launch-skip 3 -c 1 -f -o gpurun_out/top_$TAG python bench.py --steps 2 --warmup 3 --cpu-cells 100000 > gpurun_out/ncu_full_$TAG.log 2>&1; echo "ncu full rc=$?"
