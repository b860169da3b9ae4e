#!/usr/bin/env python
"""Node locality and genuinely unstructured meshes on one GPU.  One JSON line per configuration:
  delaunay            scipy.spatial.Delaunay of N uniform random points of the unit cube (rows differ in local topology: 10 .. 60
                      neighbours, up to ~100 cells per node), nodes in their own RANDOM numbering: no locality at all
  delaunay + reorder  the same arrays with b2f_options.reorder = MORTON (locality supplied inside the library)
  delaunay_morton     the same mesh renumbered by the CALLER along a Morton curve (what a mesh generator / RCM pass delivers)
  cube_shuffled (+ reorder)   the headline Kuhn cube N = 203 with a random cell order and a random node renumbering
Each line: ms/step, GNNZ/s, algorithmic GB/s and fraction of the measured HBM peak, sparsity build time, a parity sample
(5000 rows against the extended-precision host recomputation of bench.py) and the fallback / not-found counters."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


def morton_renumber(cells, coords):
    q = np.minimum(((coords - coords.min(axis=0)) / np.ptp(coords, axis=0).clip(1e-300) * 1024).astype(np.int64), 1023)
    key = np.zeros(coords.shape[0], dtype=np.int64)
    for b in range(10):
        for d in range(3):
            key |= ((q[:, d] >> b) & 1) << (3 * b + d)
    order = np.argsort(key, kind="stable")  # new id i holds old node order[i]
    newid = np.empty_like(order)
    newid[order] = np.arange(order.shape[0])
    return np.ascontiguousarray(newid[cells].astype(np.int32)), np.ascontiguousarray(coords[order])


def measure(name, cells, coords, reorder, steps, warmup):
    import torch
    import arcanefem_b200 as fem
    from arcanefem_b200.bsr import Mesh
    import bench
    from bench_configs import peak
    bsr = fem.BSRFormat(1)
    bsr.initialize(Mesh.from_numpy(3, cells, coords), True, reorder=reorder)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    bsr.computeSparsity()
    torch.cuda.synchronize()
    e0.record()
    bsr.computeSparsity()
    e1.record()
    torch.cuda.synchronize()
    sparsity_ms = e0.elapsed_time(e1)
    op = fem.Operator.laplace()
    for _ in range(warmup):
        bsr.assembleBilinear(op)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        bsr.assembleBilinear(op)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    mat = bsr.matrix()
    nc, nn, nnzb = cells.shape[0], coords.shape[0], mat.nbCol()
    st = bsr.stats()
    b_alg = bench.algorithmic_bytes(4, nc, nn, nnzb, 1)
    parity = bench.parity_sample(cells, coords, mat, np.arange(nn, dtype=np.int64), None, nrows=5000)
    k = mat.nbNzPerRow()
    out = {"config": name, "reorder": reorder, "reordered": st["reordered"], "cells": nc, "nodes": nn, "nnz": nnzb, "row_length_mean": round(float(k.double().mean()), 2),
           "row_length_max": int(k.max()), "ms_per_step": round(ms, 4), "gnnz_per_s": round(nnzb / (ms * 1e-3) / 1e9, 2), "algorithmic_bytes": b_alg,
           "achieved_gbs": round(b_alg / (ms * 1e-3) / 1e9, 1), "frac_of_measured_hbm_peak": round(b_alg / (ms * 1e-3) / 1e9 / peak(), 4), "sparsity_build_ms": round(sparsity_ms, 2),
           "sparsity_by_lists": st["sparsity_by_lists"], "fan_steps": st["fan_steps"], "gather_fallback_tiles": st["gather_fallback_tiles"],
           "columns_not_found": st["columns_not_found"], "parity_sample": {q: parity[q] for q in ("rows", "entries", "max_rel", "over_1e-12", "vs_extended_precision_max_rel", "columns_missing")}}
    print(json.dumps(out), flush=True)
    del bsr
    torch.cuda.empty_cache()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=2_000_000)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--cube", type=int, default=203)
    args = ap.parse_args()
    from bench_configs import delaunay_mesh
    from arcanefem_b200 import mesh as gen
    t0 = time.time()
    cells, coords = delaunay_mesh(args.points)
    print(json.dumps({"delaunay_points": args.points, "cells": int(cells.shape[0]), "generation_s": round(time.time() - t0, 1)}), flush=True)
    measure("delaunay, random node numbering", cells, coords, 0, args.steps, args.warmup)
    measure("delaunay, random node numbering, reorder=MORTON", cells, coords, 1, args.steps, args.warmup)
    measure("delaunay, random node numbering, reorder=PENCIL", cells, coords, 2, args.steps, args.warmup)
    mc, mx = morton_renumber(cells, coords)
    measure("delaunay, nodes numbered along a Morton curve by the caller", mc, mx, 0, args.steps, args.warmup)
    del cells, coords, mc, mx
    if args.cube > 0:
        cells, coords = gen.cube_kuhn(args.cube, shuffle_cells=42, shuffle_nodes=7)
        measure(f"Kuhn cube {args.cube}^3, random cell order + random node numbering", cells, coords, 0, args.steps, args.warmup)
        measure(f"Kuhn cube {args.cube}^3, random cell order + random node numbering, reorder=MORTON", cells, coords, 1, args.steps, args.warmup)
        measure(f"Kuhn cube {args.cube}^3, random cell order + random node numbering, reorder=PENCIL", cells, coords, 2, args.steps, args.warmup)
        cells, coords = gen.cube_kuhn(args.cube)
        measure(f"Kuhn cube {args.cube}^3, lexicographic numbering (headline mesh), reorder=PENCIL", cells, coords, 2, args.steps, args.warmup)


if __name__ == "__main__":
    main()
