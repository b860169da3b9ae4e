#!/bin/bash
# round 2, 2-GPU pass: N-rank parity tests, config #5 as written at N = 2 (strong) with the weak run nested, both arms
TAG=${1:-r02c}; NP=${2:-2}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_multirank.py -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
tail -4 gpurun_out/pytest_gpu_$TAG.log
timeout 300 python tools/fan_plan_stats.py 203 > gpurun_out/fan_plan_stats_$TAG.json 2> gpurun_out/fan_plan_stats_$TAG.err; cat gpurun_out/fan_plan_stats_$TAG.json
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $NP --steps 10 --warmup 3 > gpurun_out/bench_${TAG}_${NP}gpu.json 2> gpurun_out/bench_${TAG}_${NP}gpu.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_${TAG}_${NP}gpu.err
python -c "
import json; d=json.load(open('gpurun_out/bench_${TAG}_${NP}gpu.json')); print(d['scaling'], d['ms_per_step'], d['value'], d['config']['nnz_job'], d['parity_sample'], d['e2e'], d['config']['sparsity_build_ms'], d['config']['step_breakdown']); print(d.get('weak_scaling'))"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $NP --steps 2 --warmup 1 > gpurun_out/bench_ref_${TAG}_${NP}gpu.json 2> gpurun_out/bench_ref_${TAG}_${NP}gpu.err; echo "ref rc=$?"; cut -c1-500 gpurun_out/bench_ref_${TAG}_${NP}gpu.json
