#!/bin/bash
# tools/gpu_sparsity.sh <tag> : plan/sparsity parity tests, the bench line (sparsity_build_ms) and one full ncu capture per build kernel
TAG=${1:-s}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "plan or sparsity or edge or variant or golden" > gpurun_out/pytest_sp_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_sp_$TAG.log
timeout 600 python bench.py --steps 10 --warmup 3 --cpu-cells 100000 > gpurun_out/bench_sp_$TAG.json 2> gpurun_out/bench_sp_$TAG.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_sp_$TAG.json')); c=d['config']; print(d['ms_per_step'], c['sparsity_build_ms'], c['sparsity_build_first_call_ms'], c['values_checksum'])"
if [ -n "$NCU" ]; then
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$NCU" -c 3 -f -o gpurun_out/sp_$TAG python bench.py --steps 1 --warmup 3 --cpu-cells 100000 > gpurun_out/ncu_sp_$TAG.log 2>&1; echo "ncu rc=$?"
fi
