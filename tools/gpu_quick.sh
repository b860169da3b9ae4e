#!/bin/bash
# quick GPU pass: parity tests + one bench line (+ optional ncu full capture when $2 is a kernel regex)
TAG=${1:-q}; KREGEX=${2:-}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
tail -12 gpurun_out/pytest_gpu_$TAG.log
timeout 600 python bench.py --steps 10 --warmup 3 --cpu-cells 2000000 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_$TAG.err
cut -c1-400 gpurun_out/bench_$TAG.json
if [ -n "$KREGEX" ]; then
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$KREGEX --launch-skip 3 -c 1 -f -o gpurun_out/top_$TAG python bench.py --steps 2 --warmup 3 --cpu-cells 100000 > gpurun_out/ncu_full_$TAG.log 2>&1; echo "ncu full rc=$?"
fi
