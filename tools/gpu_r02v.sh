#!/bin/bash
# first-visit marks on the other workloads: same build with and without them, same box
mkdir -p gpurun_out
for lib in base nofirst base nofirst; do
  if [ "$lib" = "base" ]; then unset B2F_LIBRARY; else export B2F_LIBRARY=$PWD/arcanefem_b200/lib/libb200fem_$lib.so; fi
  for w in elasticity3d heat3d heat2d poisson2d; do
    timeout 300 python bench.py --steps 20 --warmup 3 --workload $w --cpu-cells 200000 --no-parity > gpurun_out/bench_r02v_${w}_$lib.json 2> gpurun_out/bench_r02v_${w}_$lib.err
    python -c "
import json; d=json.load(open('gpurun_out/bench_r02v_${w}_$lib.json')); print('$lib $w', d['ms_per_step'], d['value'], d['roofline']['frac'])"
  done
done
