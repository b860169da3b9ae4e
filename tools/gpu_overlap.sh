#!/bin/bash
# gpurun --gpus N -- tools/gpu_overlap.sh N tag : multi-rank parity test, then the N-GPU bench with and without the overlapped exchange
N=${1:-2}; TAG=${2:-ov}
mkdir -p gpurun_out
if [ -z "$SKIP_TESTS" ]; then timeout 900 python -m pytest tests/test_gpu_multirank.py -m gpu -x -q > gpurun_out/pytest_multirank_$TAG.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_multirank_$TAG.log; fi
for mode in ${MODES:-overlap sequential}; do
  extra=""; [ $mode = sequential ] && extra="--no-overlap"
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 --cpu-cells 200000 $extra > gpurun_out/${TAG}_${mode}_$N.json 2> gpurun_out/${TAG}_${mode}_$N.err
  echo "$mode rc=$?"; tail -3 gpurun_out/${TAG}_${mode}_$N.err | cut -c1-300
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${TAG}_${mode}_$N.json').read().strip().splitlines()[-1])
    print('$mode', $N, d['ms_per_step'], d['value'], d['config']['step_breakdown'], d['e2e'], d['gpu_launches'])
except Exception as e: print('no json', e)
PY
done
