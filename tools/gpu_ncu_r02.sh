#!/bin/bash
# round 2 ncu pass (one GPU): launch list of the bench command + one --set full capture per kernel of DESIGN.md's table.
# Every ncu line is preceded by the same command run plain (B200_PROFILING.md): the plain run must exit 0.
TAG=${1:-r02}
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --cpu-cells 100000 --no-parity"
$B > gpurun_out/ncu_plain_$TAG.log 2>&1 || { echo "plain bench failed"; tail -5 gpurun_out/ncu_plain_$TAG.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$TAG.csv $B > gpurun_out/ncu_launch_$TAG.log 2>&1; echo "ncu launches rc=$?"
WANT=${NCU_CAPS:-fan build_fan row_unique pack_fan spmv ij_fill fan_elasticity fan_heat3d}
cap() { # name regex skip cmd...
  local name=$1 rx=$2 skip=$3; shift 3
  [[ " $WANT " == *" $name "* ]] || return 0
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$rx --launch-skip $skip -c 1 -f -o gpurun_out/ncu_${TAG}_$name "$@" > gpurun_out/ncu_${TAG}_$name.log 2>&1; echo "ncu $name rc=$?"
}
cap fan k_node_fan 4 $B
cap build_fan k_build_fan 1 $B
cap row_unique k_row_unique 2 $B
cap pack_fan k_pack_fan 1 $B
cap spmv k_spmv 1 $B
cap ij_fill k_ij_fill 1 $B
E="python bench.py --steps 2 --warmup 3 --cpu-cells 100000 --workload elasticity3d"
$E > gpurun_out/ncu_plain_el_$TAG.log 2>&1 && cap fan_elasticity k_node_fan 4 $E
H="python bench.py --steps 2 --warmup 3 --cpu-cells 100000 --workload heat3d"
$H > gpurun_out/ncu_plain_heat_$TAG.log 2>&1 && cap fan_heat3d k_node_fan 4 $H
ls -la gpurun_out/*.ncu-rep
