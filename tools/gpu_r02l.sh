#!/bin/bash
# fan-kernel prefetch-depth variants (A/B on one box), then one --set full capture of the default kernel for the source page
TAG=${1:-r02l}
mkdir -p gpurun_out
BENCH_ARGS="--no-parity" bash tools/gpu_variants.sh base a2w24 a2w20 w20 pf1 pf2a base
B="python bench.py --steps 2 --warmup 3 --cpu-cells 100000 --no-parity"
$B > gpurun_out/ncu_plain_$TAG.log 2>&1 || { echo "plain bench failed"; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_node_fan --launch-skip 4 -c 1 -f -o gpurun_out/ncu_${TAG}_fan $B > gpurun_out/ncu_${TAG}_fan.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/*.ncu-rep
