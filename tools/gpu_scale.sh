#!/bin/bash
# N-GPU pass of config #5: tools/gpu_scale.sh TAG NP  (run under `gpurun --gpus NP`).
# N-rank parity tests first, then the strong line (one 406^3 cube over NP ranks, the weak run nested) and the reference arm.
TAG=${1:-r02}; NP=${2:-8}
mkdir -p gpurun_out
timeout ${SCALE_PYTEST_TIMEOUT:-900} python -m pytest tests/test_gpu_multirank.py -x -q ${SCALE_K:+-k "$SCALE_K"} > gpurun_out/pytest_multirank_${TAG}_${NP}gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multirank_${TAG}_${NP}gpu.log
tail -4 gpurun_out/pytest_multirank_${TAG}_${NP}gpu.log
for n in ${SCALE_NS:-2 4 8}; do
  [ $n -le $NP ] || continue
  timeout ${SCALE_BENCH_TIMEOUT:-900} python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29520 + n)) bench.py --gpus $n --steps 10 --warmup 3 \
    > gpurun_out/scale_${TAG}_${n}gpu.json 2> gpurun_out/scale_${TAG}_${n}gpu.err; echo "bench N=$n rc=$?"; tail -2 gpurun_out/scale_${TAG}_${n}gpu.err
  python -c "
import json
t=open('gpurun_out/scale_${TAG}_${n}gpu.json').read(); d=json.loads(t[t.index('{'):])
print($n, d['scaling'], d['ms_per_step'], d['value'], d['parity_sample'], d['e2e'], d['config']['sparsity_build_ms'], d['config']['step_breakdown']); print(d.get('weak_scaling'))"
done
if [ -n "$SCALE_REF" ]; then
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1 --master-port 29531 bench.py --impl reference --gpus $NP --steps 2 --warmup 1 \
  > gpurun_out/scale_ref_${TAG}_${NP}gpu.json 2> gpurun_out/scale_ref_${TAG}_${NP}gpu.err; echo "ref rc=$?"; cut -c1-400 gpurun_out/scale_ref_${TAG}_${NP}gpu.json
fi
true
if [ -n "$SCALE_PROBE" ]; then
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1 --master-port 29541 tools/pcie_probe.py > gpurun_out/pcie_probe_${TAG}_${NP}gpu.json 2> gpurun_out/pcie_probe_${TAG}_${NP}gpu.err; echo "probe rc=$?"; grep probe gpurun_out/pcie_probe_${TAG}_${NP}gpu.json
fi
true
