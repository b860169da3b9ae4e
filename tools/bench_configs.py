#!/usr/bin/env python
"""Secondary workloads of BASELINE.json (configs #3 and #4) on one GPU — the headline metric stays in bench.py.

  elasticity3d  config #3: linear elasticity 3D P1 tets, BSR 3x3, N=149 (19.8M tets), step = values assembly + Dirichlet
                row elimination on the plane z=0 (flag + apply)
  heat3d/heat2d config #4: heat, mass + stiffness re-assembled every time step on a fixed pattern (repeated-assembly throughput)
  poisson2d     2D twin of the headline (square, 2 triangles per quad)
  poisson3d_shuffled_cells / poisson3d_shuffled   the headline mesh with a random cell order / random cell order and node
                numbering (the unstructured-access cost); not in the default list
Prints one JSON line per workload: GNNZ/s, ms/step, algorithmic GB/s and the fraction of the measured HBM peak.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


def delaunay_mesh(npts, seed=11, order="random"):
    """Genuinely unstructured tetrahedra: scipy.spatial.Delaunay of `npts` uniform random points of the unit cube (rows differ
    in local topology, valences 10 .. 40).  order: "random" = the points' own (random) numbering, no locality at all;
    "morton" = the caller numbers its nodes along a Morton curve (what a mesh generator or an RCM pass gives)."""
    from scipy.spatial import Delaunay
    rng = np.random.default_rng(seed)
    pts = rng.uniform(size=(npts, 3))
    if order == "morton":
        q = np.minimum((pts * 1024).astype(np.int64), 1023)
        key = np.zeros(npts, dtype=np.int64)
        for b in range(10):
            for d in range(3):
                key |= ((q[:, d] >> b) & 1) << (3 * b + d)
        pts = pts[np.argsort(key, kind="stable")]
    tet = Delaunay(pts).simplices.astype(np.int32)
    # positive orientation like a mesh reader delivers
    a, b, c, d = (pts[tet[:, i]] for i in range(4))
    neg = np.einsum("ij,ij->i", np.cross(b - a, c - a), d - a) < 0
    tet[neg, 2], tet[neg, 3] = tet[neg, 3].copy(), tet[neg, 2].copy()
    vol = np.abs(np.einsum("ij,ij->i", np.cross(b - a, c - a), d - a))
    tet = tet[vol > 1e-12 * np.median(vol)]  # drop the exactly flat slivers of the hull (degenerate cells carry no stiffness)
    return np.ascontiguousarray(tet), np.ascontiguousarray(pts)


def run(workload, n, steps, warmup, variant, reorder=0):
    import torch
    import arcanefem_b200 as fem
    from arcanefem_b200 import mesh as gen
    from arcanefem_b200.bsr import Mesh
    if workload == "elasticity3d":
        dim, bs, (cells, coords), op = 3, 3, gen.cube_kuhn(n), fem.Operator.elasticity(21.0e5, 0.28)
    elif workload == "heat3d":
        dim, bs, (cells, coords), op = 3, 1, gen.cube_kuhn(n), None
    elif workload == "heat2d":
        dim, bs, (cells, coords), op = 2, 1, gen.square(n), None
    elif workload == "poisson2d":
        dim, bs, (cells, coords), op = 2, 1, gen.square(n), fem.Operator.laplace()
    elif workload == "poisson3d_shuffled_cells":  # random cell permutation (seed 42), node numbering kept
        dim, bs, (cells, coords), op = 3, 1, gen.cube_kuhn(n, shuffle_cells=42), fem.Operator.laplace()
    elif workload == "poisson3d_shuffled":  # random cell permutation and random node renumbering (seed 7): no locality at all
        dim, bs, (cells, coords), op = 3, 1, gen.cube_kuhn(n, shuffle_cells=42, shuffle_nodes=7), fem.Operator.laplace()
    elif workload in ("delaunay", "delaunay_morton"):  # n = number of points
        dim, bs, (cells, coords), op = 3, 1, delaunay_mesh(n, order="morton" if workload.endswith("morton") else "random"), fem.Operator.laplace()
    else:
        raise SystemExit(f"unknown workload {workload}")
    nc, nn, nv = cells.shape[0], coords.shape[0], cells.shape[1]
    keep = []
    if op is None:
        lam = torch.as_tensor(np.random.default_rng(5).uniform(0.5, 2.0, size=nc)).cuda()
        keep.append(lam)
        op = fem.Operator.heat(1.75, 0.1, cell_lambda=lam)
    bsr = fem.BSRFormat(bs)
    bsr.initialize(Mesh.from_numpy(dim, cells, coords), True, assembly_variant=variant, reorder=reorder)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    bsr.computeSparsity()
    e0.record()
    bsr.computeSparsity()
    e1.record()
    torch.cuda.synchronize()
    sparsity_ms = e0.elapsed_time(e1)
    mat = bsr.matrix()
    nnzb, nnz = mat.nbCol(), mat.nbNz()
    ls = fem.DoFLinearSystem()
    bsr.toLinearSystem(ls)
    group = torch.as_tensor(gen.boundary_nodes(coords, dim - 1, 0.0)).cuda()

    def step():
        bsr.assembleBilinear(op)
        if workload == "elasticity3d":
            ls.clearValues()
            for comp in range(bs):
                bsr.applyDirichletViaRowElimination(group, 0.0, dof_index=comp)
            ls.applyDirichlet()

    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    b_alg = 4 * nv * nc + 24 * nn + 4 * (nn + 1) + 4 * nnzb + 8 * bs * bs * nnzb
    if workload.startswith("heat"):
        b_alg += 8 * nc  # per-cell conductivity
    gbs = b_alg / (ms * 1e-3) / 1e9
    st = bsr.stats()
    parity = None
    if workload.startswith("delaunay") or workload.startswith("poisson3d_shuffled"):
        # a fixed sample of rows against the extended-precision host recomputation (bench.py parity_sample)
        import bench
        parity = bench.parity_sample(cells, coords, mat, np.arange(nn, dtype=np.int64), None, nrows=5000)
    return {"workload": workload, "reorder": reorder, "reordered": st["reordered"], "sparsity_build_ms": round(sparsity_ms, 3), "parity_sample": parity, "n": n, "cells": nc, "nodes": nn, "block_size": bs, "nnz": nnz, "ms_per_step": round(ms, 4), "gnnz_per_s": round(nnz / (ms * 1e-3) / 1e9, 2),
            "algorithmic_bytes": b_alg, "achieved_gbs": round(gbs, 1), "frac_of_measured_hbm_peak": round(gbs / peak(), 4), "steps": steps, "variant": variant,
            "gather_fallback_tiles": st["gather_fallback_tiles"], "columns_not_found": st["columns_not_found"],
            "checksum": float(mat.values().abs().sum().item())}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workloads", default="elasticity3d,heat3d,heat2d,poisson2d")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--variant", type=int, default=3)
    ap.add_argument("--reorder", type=int, default=0, help="b2f_options.reorder: 1 = Morton order inside the library")
    ap.add_argument("--points", type=int, default=2_000_000, help="points of the delaunay workloads")
    args = ap.parse_args()
    sizes = {"elasticity3d": 149, "heat3d": 203, "heat2d": 5000, "poisson2d": 5000, "poisson3d_shuffled_cells": 203, "poisson3d_shuffled": 203, "delaunay": args.points,
             "delaunay_morton": args.points}
    for w in args.workloads.split(","):
        print(json.dumps(run(w, sizes[w], args.steps, args.warmup, args.variant, args.reorder)), flush=True)


if __name__ == "__main__":
    main()
