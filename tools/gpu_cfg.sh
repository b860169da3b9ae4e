#!/bin/bash
# secondary workloads + ncu launch list of the elasticity workload
mkdir -p gpurun_out
python tools/bench_configs.py --workloads ${1:-elasticity3d,heat3d,heat2d} > gpurun_out/configs_$2.jsonl 2> gpurun_out/configs_$2.err; tail -3 gpurun_out/configs_$2.err; cut -c1-200 gpurun_out/configs_$2.jsonl
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_elast_$2.csv python tools/bench_configs.py --workloads elasticity3d --steps 3 --warmup 1 > /dev/null 2>&1; echo ncu rc=$?
