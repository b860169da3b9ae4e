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


def assert_values_close(got, ref, row_index, bs, rtol=RTOL, what="values"):
    """|d| <= rtol*max(|a|,|b|)  or  |d| <= rtol*||block row||_inf (entries that cancel to ~0; SURVEY.md §7 hard parts)."""
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert got.shape == ref.shape, (got.shape, ref.shape)
    assert np.isfinite(got).all(), f"{what}: non-finite entries"
    d = np.abs(got - ref)
    tol = rtol * np.maximum(np.maximum(np.abs(got), np.abs(ref)), row_scale(ref, row_index, bs))
    bad = d > tol
    if bad.any():
        i = int(np.argmax(d / np.maximum(tol, 1e-300)))
        raise AssertionError(f"{what}: {int(bad.sum())} of {d.size} entries differ; worst idx {i}: got {got[i]!r} ref {ref[i]!r} |d|={d[i]:.3e} tol={tol[i]:.3e}")


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
