#!/bin/bash
mkdir -p gpurun_out
BENCH_ARGS="--no-parity" bash tools/gpu_variants.sh base early aa aae a2w24 a2w20 w20 pf1 pf2a base early aa aae
