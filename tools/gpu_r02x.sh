#!/bin/bash
# heat kernels with staged cell ids as the default: whole GPU suite + heat lines
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_r02x.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/pytest_gpu_r02x.log
for w in heat3d heat2d; do
  timeout 600 python bench.py --steps 20 --warmup 3 --workload $w --cpu-cells 3000000 > gpurun_out/bench_r02x_$w.json 2> gpurun_out/bench_r02x_$w.err; echo "bench $w rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/bench_r02x_$w.json')); print('$w', d['ms_per_step'], d['value'], d['roofline']['frac'], d['e2e']['ms_per_step'])"
done
H="python bench.py --steps 2 --warmup 3 --cpu-cells 100000 --workload heat3d"
$H > gpurun_out/ncu_plain_heat_r02x.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_node_fan --launch-skip 4 -c 1 -f -o gpurun_out/ncu_r02x_fan_heat3d $H > gpurun_out/ncu_r02x_fan_heat3d.log 2>&1; echo "ncu rc=$?"
