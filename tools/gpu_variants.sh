#!/bin/bash
# bench each tuning build of the library: tools/gpu_variants.sh name1 name2 ...  ("base" = the in-tree default)
mkdir -p gpurun_out
for n in "$@"; do
  if [ "$n" = "base" ]; then unset B2F_LIBRARY; else export B2F_LIBRARY=$PWD/arcanefem_b200/lib/libb200fem_$n.so; fi
  timeout 300 python bench.py --steps 20 --warmup 3 --cpu-cells 200000 ${BENCH_ARGS:-} > gpurun_out/var_$n.json 2> gpurun_out/var_$n.err || tail -3 gpurun_out/var_$n.err
  python -c "
import json,sys
d=json.load(open('gpurun_out/var_$n.json')); print('$n', d['ms_per_step'], d['value'], d['roofline']['frac'], d['config']['values_checksum'])"
done
