#!/bin/bash
# bench lines after the parity_sample key change: N = 1 (driver's command) and N = 2 under torchrun
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_r02ab_1gpu.json 2> gpurun_out/bench_r02ab_1gpu.err; echo "N=1 rc=$?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_r02ab_2gpu.json 2> gpurun_out/bench_r02ab_2gpu.err; echo "N=2 rc=$?"; tail -3 gpurun_out/bench_r02ab_2gpu.err
python - <<'PY'
import json
for f in ('gpurun_out/bench_r02ab_1gpu.json','gpurun_out/bench_r02ab_2gpu.json'):
    t=open(f).read(); d=json.loads(t[t.index('{'):])
    print(f, d['ms_per_step'], d['value'], d['parity_sample'], (d.get('weak_scaling') or {}).get('parity_sample'))
PY
