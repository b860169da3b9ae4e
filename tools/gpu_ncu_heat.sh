#!/bin/bash
mkdir -p gpurun_out
H="python bench.py --steps 2 --warmup 3 --cpu-cells 100000 --workload heat3d"
$H > gpurun_out/ncu_plain_heat_final.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_node_fan --launch-skip 4 -c 1 -f -o gpurun_out/ncu_final_fan_heat3d $H > gpurun_out/ncu_final_fan_heat3d.log 2>&1; echo "ncu rc=$?"
