#!/bin/bash
# plan build after the rewrite of k_fan_first_flags: parity of the marks, build time, launch list of one build
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "first_visits or fan_plan or medium_cube or long_rows or ragged" > gpurun_out/pytest_r02u.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_r02u.log
B="python bench.py --steps 5 --warmup 3 --cpu-cells 100000 --no-parity"
$B > gpurun_out/bench_r02u.json 2> gpurun_out/bench_r02u.err || { echo "plain bench failed"; exit 1; }
python -c "
import json; d=json.load(open('gpurun_out/bench_r02u.json')); print(d['ms_per_step'], d['config']['sparsity_build_ms'])"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_r02u.csv $B > gpurun_out/ncu_launch_r02u.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv,collections
rows=list(csv.reader(l for l in open('gpurun_out/launches_r02u.csv') if l.startswith('"')))
h=rows[0]; ix={k:i for i,k in enumerate(h)}
for r in rows[1:]:
    name=r[ix['Kernel Name']].split('(')[0][-40:]; v=float(r[ix['Metric Value']]); u=r[ix['Metric Unit']]
    v = v/1e3 if u.startswith('n') else v*1e3 if u.startswith('m') else v
    if v>300: print(f"{v:9.1f} us {name}")
PY
