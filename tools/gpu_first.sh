#!/bin/bash
# First GPU pass: parity tests, bench of the three variants, ncu launch list + one full capture of the top kernel.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
nproc > gpurun_out/nproc.txt; free -g >> gpurun_out/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_v3.json 2> gpurun_out/bench_v3.err; echo "bench v3 rc=$?"
cat gpurun_out/bench_v3.json
for v in 1 2; do
  timeout 400 python bench.py --steps 10 --warmup 3 --variant $v --cpu-cells 1000000 > gpurun_out/bench_v$v.json 2> gpurun_out/bench_v$v.err; echo "bench v$v rc=$?"
  cat gpurun_out/bench_v$v.json
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r01.csv python bench.py --steps 2 --warmup 3 --cpu-cells 100000 > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_node_gather_slots --launch-skip 3 -c 1 -f -o gpurun_out/gather_r01 python bench.py --steps 2 --warmup 3 --cpu-cells 100000 > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out
