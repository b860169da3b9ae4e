#!/bin/bash
# run-to-run spread of the headline line on one box: the driver's command three times, then 50 timed steps
mkdir -p gpurun_out
: > gpurun_out/bench_repeat.jsonl
for i in 1 2 3; do timeout 600 python bench.py --cpu-cells 200000 >> gpurun_out/bench_repeat.jsonl 2> gpurun_out/bench_repeat.err; done
timeout 600 python bench.py --steps 50 --warmup 5 --cpu-cells 200000 >> gpurun_out/bench_repeat.jsonl 2>> gpurun_out/bench_repeat.err
python - <<'PY'
import json
for l in open('gpurun_out/bench_repeat.jsonl'):
    d=json.loads(l); print(d['steps'], d['ms_per_step'], d['value'], d['roofline']['frac'], d['e2e']['ms_per_step'], d['clocks']['sm_mhz'], d['clocks']['reasons'])
PY
