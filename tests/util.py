"""Shared helpers of the test-suite: tolerance rule and oracle-side reference results."""
import numpy as np

from arcanefem_b200 import mesh as gen
from oracle import oracle as orc

RTOL = 1e-12  # north-star: matrix/RHS values within 1e-12 relative per entry in fp64


def row_scale(values, row_index, bs):
    """max |a| over each block row, broadcast to every scalar of that block row (either value layout)."""
    nn = row_index.shape[0] - 1
    counts = np.diff(row_index).astype(np.int64) * bs * bs
    owner = np.repeat(np.arange(nn), counts)
    scale = np.zeros(nn)
    np.maximum.at(scale, owner, np.abs(values))
    return scale[owner]


PURE_RTOL = 1e-11  # pure relative error allowed on the entries that do not cancel (|ref| > PURE_FLOOR * row scale)
PURE_FLOOR = 1e-3


def assert_values_close(got, ref, row_index, bs, rtol=RTOL, what="values"):
    """Two checks, both asserted, the measured figures printed and returned:
    (1) row-norm-relative: |d| <= rtol * max(|a|, |b|, ||block row||_inf) - the floor is what lets entries that cancel to ~0
        pass (SURVEY.md §7 hard parts); by itself it would also let a small genuine entry carry a large relative error, so
    (2) pure relative: |d| / |ref| <= PURE_RTOL on every entry with |ref| > PURE_FLOOR * row scale (entries that do not
        cancel).  The worst pure relative error over |ref| > 1e-6 * row scale is reported too (not asserted: an entry a
        million times below its row's scale legitimately carries eps * 1e6 of relative error in ANY fp64 summation)."""
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert got.shape == ref.shape, (got.shape, ref.shape)
    assert np.isfinite(got).all(), f"{what}: non-finite entries"
    d = np.abs(got - ref)
    scale = row_scale(ref, row_index, bs)
    tol = rtol * np.maximum(np.maximum(np.abs(got), np.abs(ref)), scale)
    bad = d > tol
    if bad.any():
        i = int(np.argmax(d / np.maximum(tol, 1e-300)))
        raise AssertionError(f"{what}: {int(bad.sum())} of {d.size} entries differ; worst idx {i}: got {got[i]!r} ref {ref[i]!r} |d|={d[i]:.3e} tol={tol[i]:.3e}")
    aref = np.abs(ref)
    stats = {"entries": int(d.size), "worst_row_norm_relative": float((d / np.maximum(scale, 1e-300)).max()) if d.size else 0.0}
    for name, floor in (("worst_pure_relative", PURE_FLOOR), ("worst_pure_relative_1e-6", 1e-6)):
        m = aref > floor * scale
        stats[name] = float((d[m] / aref[m]).max()) if m.any() else 0.0
        stats[name + "_entries"] = int(m.sum())
    stats["passing_only_through_the_floor"] = int((d > rtol * np.maximum(np.abs(got), aref)).sum())
    print(f"[parity] {what}: {stats}")
    assert stats["worst_pure_relative"] <= PURE_RTOL, f"{what}: pure relative error {stats['worst_pure_relative']:.3e} on an entry above {PURE_FLOOR} of its row scale"
    return stats


def assert_vec_close(got, ref, rtol=RTOL, what="vector"):
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert got.shape == ref.shape
    d = np.abs(got - ref)
    scale = np.abs(ref).max() if ref.size else 0.0
    tol = rtol * np.maximum(np.maximum(np.abs(got), np.abs(ref)), scale)
    assert (d <= tol).all(), f"{what}: max |d| {d.max():.3e} (scale {scale:.3e})"


def _mixed_orientation(dim, cells, coords, seed=11):
    """Swap two vertices of about half of the cells: negatively oriented cells, as a reader may deliver them."""
    rng = np.random.default_rng(seed)
    flip = rng.uniform(size=cells.shape[0]) < 0.5
    cells = cells.copy()
    a, b = (1, 2) if dim == 2 else (2, 3)
    cells[flip, a], cells[flip, b] = cells[flip, b].copy(), cells[flip, a].copy()
    # and rotate some others (even permutation) so that the row node sits at every local position
    rot = rng.uniform(size=cells.shape[0]) < 0.5
    cells[rot] = np.roll(cells[rot], 1, axis=1) if dim == 2 else cells[rot][:, [1, 0, 3, 2]]
    return dim, np.ascontiguousarray(cells), coords


MESHES = {
    "cube4_mixed_orientation": lambda: _mixed_orientation(3, *gen.cube_kuhn(4, shuffle_cells=3)),
    "square7_mixed_orientation": lambda: _mixed_orientation(2, *gen.square(7, shuffle_cells=3)),
    "cube1": lambda: (3,) + gen.cube_kuhn(1, perturb=0.0),
    "cube2": lambda: (3,) + gen.cube_kuhn(2),
    "cube5": lambda: (3,) + gen.cube_kuhn(5),
    "cube5_flat": lambda: (3,) + gen.cube_kuhn(5, perturb=0.0),
    "cube7_shuffled": lambda: (3,) + gen.cube_kuhn(7, shuffle_cells=42, shuffle_nodes=7),
    "cube12": lambda: (3,) + gen.cube_kuhn(12, L=3.0),
    "square1": lambda: (2,) + gen.square(1, perturb=0.0),
    "square3": lambda: (2,) + gen.square(3),
    "square9_shuffled": lambda: (2,) + gen.square(9, shuffle_cells=42, shuffle_nodes=7),
    "square20": lambda: (2,) + gen.square(20, L=2.0),
}


def op_params(kind, nb_cell=None, seed=5):
    """(oracle kwargs, Operator factory kwargs) for the three shipped operators."""
    if kind == "laplace":
        return dict(op=orc.OP_LAPLACE)
    if kind == "elasticity":
        E, nu = 21.0e5, 0.28  # modules/elasticity/inputs/bar.2D.traction.bsr.hypre.arc:24-25
        return dict(op=orc.OP_ELASTICITY, lambda_=E * nu / ((1 + nu) * (1 - 2 * nu)), mu2=(E / (2 * (1 + nu))) * 2)
    if kind == "heat":
        return dict(op=orc.OP_HEAT, lambda_=1.75, inv_dt=1.0 / 0.1)
    if kind == "heat_cell":
        rng = np.random.default_rng(seed)
        return dict(op=orc.OP_HEAT, lambda_=0.0, inv_dt=1.0 / 0.1, cell_lambda=rng.uniform(0.5, 2.0, size=nb_cell))
    raise KeyError(kind)
