"""Parity at BASELINE.json's full sizes: configs #2 (Poisson 3D, 50.2 M tets), #3 (elasticity 3D BS = 3, 19.8 M tets, Dirichlet
row elimination) and #4 (heat 2D, 50 M triangles, re-assembled per time step), through the C ABI on one B200.

At these sizes the oracle's sequential value assembly still runs in 6-17 s, so the values are compared entry by entry
(1e-12 relative with the row-scale floor of tests/util.py, evaluated on the device); the pattern is compared bit for bit
with the closed-form lattice pattern (gen.kuhn_pattern / gen.square_pattern, pinned on the oracle's edge-key algorithm at
small sizes by tests/test_oracle.py).  On top: the size-independent properties of the operators - constants in the null
space, symmetry, bitwise repeatability, agreement of the atomic and the gather kernels, exact Dirichlet rows.
Runs last (file name) so that a memory problem here cannot hide the small-case results under `-x`.
"""
import numpy as np
import pytest

from tests.util import RTOL, op_params

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _fem():
    import arcanefem_b200 as fem
    return fem


def _row_owner(ri, bs):
    counts = torch.as_tensor(np.diff(ri).astype(np.int64) * bs * bs).cuda()
    return torch.repeat_interleave(torch.arange(counts.shape[0], device="cuda"), counts)


def _assert_close_on_device(got, ref, owner, nb_row, what, rtol=RTOL):
    """tests/util.py:assert_values_close on the device: |d| <= rtol * max(|a|, |b|, max |ref| over the block row)."""
    assert got.shape == ref.shape, (got.shape, ref.shape)
    assert bool(torch.isfinite(got).all()), f"{what}: non-finite entries"
    scale = torch.zeros(nb_row, dtype=torch.float64, device="cuda").scatter_reduce_(0, owner, ref.abs(), "amax")
    tol = rtol * torch.maximum(torch.maximum(got.abs(), ref.abs()), scale[owner])
    d = (got - ref).abs()
    bad = d > tol
    nbad = int(bad.sum())
    if nbad:
        i = int(torch.argmax(d / tol.clamp_min(1e-300)))
        raise AssertionError(f"{what}: {nbad} of {d.numel()} entries differ; worst idx {i}: got {got[i].item()!r} ref {ref[i].item()!r}")


def _assert_pattern(mat, ri, co, counts):
    nc, nn, e, nnzb = counts
    assert (mat.nbRow(), mat.nbEdge(), mat.nbCol()) == (nn, e, nnzb)
    assert torch.equal(mat.rowIndexWithEnd(), torch.as_tensor(ri).cuda())
    assert torch.equal(mat.columns(), torch.as_tensor(co).cuda())
    assert torch.equal(mat.nbNzPerRow(), torch.as_tensor(np.diff(ri).astype(np.int32)).cuda())


def _assert_symmetric(bsr, ndof, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.rand(ndof, dtype=torch.float64, device="cuda", generator=g) - 0.5
    y = torch.rand(ndof, dtype=torch.float64, device="cuda", generator=g) - 0.5
    ay, ax = bsr.spmv(y), bsr.spmv(x)
    s1, s2 = torch.dot(x, ay), torch.dot(y, ax)
    bound = torch.dot(x.abs(), ay.abs()) + torch.dot(y.abs(), ax.abs())
    assert abs(float(s1 - s2)) <= 1e-11 * float(bound), (float(s1), float(s2))


def test_config2_poisson3d_50m_tets():
    from oracle import oracle as orc
    from arcanefem_b200 import mesh as gen
    from arcanefem_b200.bsr import Mesh
    fem = _fem()
    N = 203
    cells, coords = gen.cube_kuhn(N)
    counts = gen.kuhn_counts(N)
    assert counts == (50192562, 8489664, 58929479, 126348622)
    ri, co = gen.kuhn_pattern(N)
    m = Mesh.from_numpy(3, cells, coords)
    bsr = fem.BSRFormat(1)
    bsr.initialize(m, True)
    bsr.computeSparsity()
    mat = bsr.matrix()
    _assert_pattern(mat, ri, co, counts)
    bsr.assembleBilinear(fem.Operator.laplace())
    torch.cuda.synchronize()
    st = bsr.stats()
    assert st["columns_not_found"] == 0 and st["gather_fallback_tiles"] == 0
    ref, not_found = orc.assemble(cells, coords, ri, co, bs=1, per_block=False)
    assert not_found == 0
    ref = torch.as_tensor(ref).cuda()
    owner = _row_owner(ri, 1)
    got = mat.values().clone()
    _assert_close_on_device(got, ref, owner, counts[1], "config #2 values (node gather)")
    # constants are in the null space of the Laplacian: every row sums to zero up to rounding of its own entries
    y = bsr.spmv(torch.ones(counts[1], dtype=torch.float64, device="cuda"))
    scale = torch.zeros(counts[1], dtype=torch.float64, device="cuda").scatter_reduce_(0, owner, got.abs(), "amax")
    assert bool((y.abs() <= 1e-12 * scale).all())
    _assert_symmetric(bsr, counts[1], 1)
    # bitwise repeatable (no atomics on the default path)
    bsr.assembleBilinear(fem.Operator.laplace())
    torch.cuda.synchronize()
    assert torch.equal(mat.values(), got)
    # the reference's own scheme (cell-wise fp64 atomics, femutils/BSRFormat.h:842-898) at the same size
    del bsr, mat
    b1 = fem.BSRFormat(1)
    b1.initialize(m, True, assembly_variant=fem.ASM_CELL_ATOMIC)
    b1.computeSparsity()
    _assert_pattern(b1.matrix(), ri, co, counts)
    b1.assembleBilinear(fem.Operator.laplace())
    torch.cuda.synchronize()
    assert b1.stats()["columns_not_found"] == 0
    _assert_close_on_device(b1.matrix().values(), ref, owner, counts[1], "config #2 values (cell atomic)")


def test_config3_elasticity3d_20m_tets_with_row_elimination():
    from oracle import oracle as orc
    from arcanefem_b200 import mesh as gen
    from arcanefem_b200.bsr import Mesh
    fem = _fem()
    N, bs = 149, 3
    cells, coords = gen.cube_kuhn(N)
    counts = gen.kuhn_counts(N)
    assert counts == (19847694, 3375000, 23355899, 50086798)
    ri, co = gen.kuhn_pattern(N)
    p = op_params("elasticity")
    m = Mesh.from_numpy(3, cells, coords)
    bsr = fem.BSRFormat(bs)
    bsr.initialize(m, True)
    bsr.computeSparsity()
    mat = bsr.matrix()
    _assert_pattern(mat, ri, co, counts)
    assert mat.nbNz() == 450781182
    op = fem.Operator.elasticity(21.0e5, 0.28)  # modules/elasticity/inputs/bar.2D.traction.bsr.hypre.arc:24-25
    assert (op.lambda_, op.mu2) == (p["lambda_"], p["mu2"])
    bsr.assembleBilinear(op)
    torch.cuda.synchronize()
    assert bsr.stats()["columns_not_found"] == 0
    vals, not_found = orc.assemble(cells, coords, ri, co, bs=bs, per_block=False, **p)
    assert not_found == 0
    owner = _row_owner(ri, bs)
    ref = torch.as_tensor(vals).cuda()
    _assert_close_on_device(mat.values(), ref, owner, counts[1], "config #3 values")
    ndof = counts[1] * bs
    # rigid translations are in the null space; the operator is symmetric
    for c in range(bs):
        x = torch.zeros(ndof, dtype=torch.float64, device="cuda")
        x[c::bs] = 1.0
        assert float(bsr.spmv(x).abs().max()) <= 1e-12 * float(ref.abs().max())
    _assert_symmetric(bsr, ndof, 2)
    del ref
    # Dirichlet by row elimination on z = 0, all three components (the "with Dirichlet elimination" of config #3)
    rows, nbcol, cols = orc.to_csr(bs, ri, co)
    group = gen.boundary_nodes(coords, 2, 0.0)
    assert group.size == (N + 1) ** 2
    rhs = np.zeros(ndof)
    finfo, fval = np.zeros(ndof, np.uint8), np.zeros(ndof)
    einfo, evalue = np.zeros(ndof, np.uint8), np.zeros(ndof)
    comps = [(0, 0.5), (1, 0.0), (2, -0.25)]
    for dof_index, g in comps:
        orc.dirichlet_flag_elimination(bs, dof_index, g, 1, group, einfo, evalue)
    orc.dirichlet_apply(rows, cols, vals, einfo, evalue, finfo, fval, rhs)
    ls = fem.DoFLinearSystem()
    bsr.toLinearSystem(ls)
    ls.clearValues()
    dgroup = torch.as_tensor(group).cuda()
    for dof_index, g in comps:
        bsr.applyDirichletViaRowElimination(dgroup, g, dof_index, mode=1)
    ls.applyDirichlet()
    torch.cuda.synchronize()
    assert torch.equal(ls.getEliminationInfo(), torch.as_tensor(einfo).cuda())  # masks bit-exact
    assert torch.equal(ls.getEliminationValue(), torch.as_tensor(evalue).cuda())
    assert torch.equal(ls.rhsVariable(), torch.as_tensor(rhs).cuda())  # exact stores of g
    _assert_close_on_device(mat.values(), torch.as_tensor(vals).cuda(), owner, counts[1], "config #3 values after row elimination")
    csr = mat.toCsr()
    assert torch.equal(csr.rowsWithEnd(), torch.as_tensor(rows).cuda())
    assert torch.equal(csr.columns(), torch.as_tensor(cols).cuda())
    assert torch.equal(csr.rowsNbColumn(), torch.as_tensor(nbcol).cuda())


def test_config4_heat2d_50m_triangles_reassembled_per_step():
    from oracle import oracle as orc
    from arcanefem_b200 import mesh as gen
    from arcanefem_b200.bsr import Mesh
    fem = _fem()
    N = 5000
    cells, coords = gen.square(N)
    nn, nc = (N + 1) ** 2, 2 * N * N
    ri, co = gen.square_pattern(N)
    e = 3 * N * N + 2 * N
    assert co.shape[0] == 2 * e + nn
    m = Mesh.from_numpy(2, cells, coords)
    bsr = fem.BSRFormat(1)
    bsr.initialize(m, True)
    bsr.computeSparsity()
    mat = bsr.matrix()
    _assert_pattern(mat, ri, co, (nc, nn, e, 2 * e + nn))
    owner = _row_owner(ri, 1)
    p = op_params("heat")  # lambda = 1.75, dt = 0.1 (modules/heat/inputs)
    ref, not_found = orc.assemble(cells, coords, ri, co, bs=1, per_block=False, **p)
    assert not_found == 0
    ref = torch.as_tensor(ref).cuda()
    first = None
    for step in range(3):  # the module re-assembles mass + stiffness every time step (modules/heat/FemModule.cc:113-131)
        bsr.assembleBilinear(fem.Operator.heat(p["lambda_"], 1.0 / p["inv_dt"]))
        torch.cuda.synchronize()
        if first is None:
            first = mat.values().clone()
            _assert_close_on_device(first, ref, owner, nn, "config #4 values")
        else:
            assert torch.equal(mat.values(), first)
    assert bsr.stats()["columns_not_found"] == 0
    # K 1 = 0, so the row sums are those of the mass matrix / dt: they add up to area / dt = 1 / dt
    y = bsr.spmv(torch.ones(nn, dtype=torch.float64, device="cuda"))
    assert abs(float(y.sum()) - p["inv_dt"]) <= 1e-9 * p["inv_dt"]
    _assert_symmetric(bsr, nn, 3)
