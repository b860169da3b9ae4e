#!/bin/bash
# multi-GPU pass (gpurun --gpus N): 2-GPU parity test + bench at 1..N GPUs
N=${1:-2}; TAG=${2:-mg}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus_$TAG.txt
timeout 900 python -m pytest tests/test_gpu_multirank.py -m gpu -x -q > gpurun_out/pytest_multirank_$TAG.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_multirank_$TAG.log
for g in 1 2 4 8; do
  [ $g -le $N ] || continue
  if [ $g = 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 --cpu-cells 2000000 > gpurun_out/scale_${TAG}_$g.json 2> gpurun_out/scale_${TAG}_$g.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $g --steps 10 --warmup 3 --cpu-cells 2000000 > gpurun_out/scale_${TAG}_$g.json 2> gpurun_out/scale_${TAG}_$g.err
  fi
  echo "gpus=$g rc=$?"; tail -4 gpurun_out/scale_${TAG}_$g.err | cut -c1-300
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/scale_${TAG}_$g.json').read().strip().splitlines()[-1])
    print($g, d['ms_per_step'], d['value'], d['roofline']['frac'], d['e2e'])
except Exception as e: print('no json', e)
PY
done
