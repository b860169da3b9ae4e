#!/bin/bash
# unstructured / shuffled meshes with the round's final kernels
TAG=${1:-r02zz}
mkdir -p gpurun_out
timeout 1500 python tools/bench_unstructured.py --points 2000000 > gpurun_out/unstructured_$TAG.jsonl 2> gpurun_out/unstructured_$TAG.err; echo "unstructured rc=$?"; tail -3 gpurun_out/unstructured_$TAG.err
python - <<PY
import json
for l in open('gpurun_out/unstructured_$TAG.jsonl'):
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print(d.get('name'), d.get('reorder'), d.get('ms_per_step'), d.get('gnnz_per_s'), d.get('frac'), d.get('sparsity_build_ms'), (d.get('parity_sample') or {}).get('max_rel'))
PY
