"""Parity at BASELINE.json's full sizes: configs #2 (Poisson 3D, 50.2 M tets), #3 (elasticity 3D BS = 3, 19.8 M tets, Dirichlet
row elimination) and #4 (heat 2D, 50 M triangles, re-assembled per time step), through the C ABI on one B200.

At these sizes the oracle's sequential value assembly still runs in 6-17 s, so the values are compared entry by entry
(1e-12 relative with the row-scale floor of tests/util.py, evaluated on the device; the few entries of almost flat
tetrahedra that no fp64 formula resolves to 1e-12 are settled against extended precision, see _adjudicate); the pattern is compared bit for bit
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


def _violations_on_device(got, ref, owner, nb_row, what, rtol=RTOL):
    """tests/util.py:assert_values_close evaluated on the device: returns the (host) indices of the entries with
    |d| > rtol * max(|a|, |b|, max |ref| over the block row)."""
    assert got.shape == ref.shape, (got.shape, ref.shape)
    assert bool(torch.isfinite(got).all()), f"{what}: non-finite entries"
    scale = torch.zeros(nb_row, dtype=torch.float64, device="cuda").scatter_reduce_(0, owner, ref.abs(), "amax")
    tol = rtol * torch.maximum(torch.maximum(got.abs(), ref.abs()), scale[owner])
    return torch.nonzero((got - ref).abs() > tol).flatten().cpu().numpy()


def _ke_extended(X, kind, p):
    """Element matrix of one tetrahedron in x87 extended precision (64-bit mantissa) from the exact fp64 vertex coordinates:
    true gradients G = cofactors / det, Laplace V G G^T, isotropic elasticity V (lambda G_ai G_bj + mu G_aj G_bi + mu d_ij G_a.G_b)."""
    X = X.astype(np.longdouble)
    e = X[1:] - X[0]
    g = np.zeros((4, 3), np.longdouble)
    g[1], g[2], g[3] = np.cross(e[1], e[2]), np.cross(e[2], e[0]), np.cross(e[0], e[1])
    det = np.dot(e[0], g[1])
    G = g / det
    G[0] = -(G[1] + G[2] + G[3])
    vol = abs(det) / 6
    if kind == "laplace":
        return vol * (G @ G.T)
    lam, mu = np.longdouble(p["lambda_"]), np.longdouble(p["mu2"]) / 2
    K = np.zeros((12, 12), np.longdouble)
    for a in range(4):
        for b in range(4):
            dab = np.dot(G[a], G[b])
            for i in range(3):
                for j in range(3):
                    K[3 * a + i, 3 * b + j] = vol * (lam * G[a, i] * G[b, j] + mu * G[a, j] * G[b, i] + (mu * dab if i == j else 0))
    return K


def _adjudicate(bad, got, ref, cells, coords, ri, co, bs, kind, p, what, max_fraction=1e-5):
    """Entries outside the 1e-12 rule.  On the jittered 50 M-tet cube a handful of tetrahedra are almost flat (|det| thousands of
    times below nominal): cancellation in the determinant alone costs them kappa * eps ~ 1e-12 relative in ANY fp64 formula,
    the oracle's included, so neither side is the truth there.  Such entries are recomputed in extended precision (exact to
    ~1e-16 here) and the CUDA value must be within 1e-12 of it or no further from it than 4x the oracle's own error.  They
    must stay rare (< 1e-5 of the entries)."""
    assert np.finfo(np.longdouble).eps < 1e-18, "needs x87 extended precision"
    assert bad.size <= max_fraction * ref.numel(), f"{what}: {bad.size} of {ref.numel()} entries differ"
    if bad.size == 0:
        print(f"[adjudicate] {what}: every one of the {ref.numel()} entries is within the 1e-12 rule of the oracle")
        return 0
    bs2 = bs * bs
    r = np.searchsorted(ri.astype(np.int64) * bs2, bad, side="right") - 1
    rb, k = ri[r].astype(np.int64), (ri[r + 1] - ri[r]).astype(np.int64)
    local = bad - rb * bs2
    i, rem = local // (bs * k), local % (bs * k)  # per-row layout, femutils/BSRFormat.h:162
    x, j = rem // bs, rem % bs
    c = co[rb + x]
    sub = cells[np.isin(cells, np.unique(r)).any(axis=1)]
    g, f = got[torch.as_tensor(bad).cuda()].cpu().numpy(), ref[torch.as_tensor(bad).cuda()].cpu().numpy()
    worst_ratio, worst_rel = 0.0, 0.0
    for q in range(bad.size):
        exact = np.longdouble(0)
        for cn in sub[(sub == r[q]).any(axis=1) & (sub == c[q]).any(axis=1)]:
            a, b = int(np.nonzero(cn == r[q])[0][0]), int(np.nonzero(cn == c[q])[0][0])
            exact += _ke_extended(coords[cn], kind, p)[bs * a + i[q], bs * b + j[q]]
        err_got, err_ref = abs(np.longdouble(g[q]) - exact), abs(np.longdouble(f[q]) - exact)
        assert err_got <= max(4 * err_ref, RTOL * abs(exact)), (what, int(bad[q]), g[q], f[q], float(exact))
        worst_ratio = max(worst_ratio, float(err_got / max(err_ref, np.longdouble(1e-300))))
        worst_rel = max(worst_rel, float(err_got / abs(exact)) if exact != 0 else 0.0)
    # the carve-out is reported, not only asserted: how many entries, how far from extended precision, relative to the oracle
    print(f"[adjudicate] {what}: {bad.size} of {ref.numel()} entries outside the 1e-12 rule vs the oracle; against extended precision the CUDA values are "
          f"within {worst_rel:.3e} relative, at most {worst_ratio:.2f}x the oracle's own error")
    return int(bad.size)


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
    bad = _violations_on_device(got, ref, owner, counts[1], "config #2 values (node gather)")
    _adjudicate(bad, got, ref, cells, coords, ri, co, 1, "laplace", None, "config #2 values (node gather)")
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
    got1 = b1.matrix().values()
    bad = _violations_on_device(got1, ref, owner, counts[1], "config #2 values (cell atomic)")
    _adjudicate(bad, got1, ref, cells, coords, ri, co, 1, "laplace", None, "config #2 values (cell atomic)")


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
    bad0 = _violations_on_device(mat.values(), ref, owner, counts[1], "config #3 values")
    _adjudicate(bad0, mat.values(), ref, cells, coords, ri, co, bs, "elasticity", p, "config #3 values")
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
    bad1 = _violations_on_device(mat.values(), torch.as_tensor(vals).cuda(), owner, counts[1], "config #3 values after row elimination")
    assert np.isin(bad1, bad0).all()  # eliminated rows are exact; elsewhere the values are the ones adjudicated above
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
            assert _violations_on_device(first, ref, owner, nn, "config #4 values").size == 0
        else:
            assert torch.equal(mat.values(), first)
    assert bsr.stats()["columns_not_found"] == 0
    # K 1 = 0, so the row sums are those of the mass matrix / dt: they add up to area / dt = 1 / dt
    y = bsr.spmv(torch.ones(nn, dtype=torch.float64, device="cuda"))
    assert abs(float(y.sum()) - p["inv_dt"]) <= 1e-9 * p["inv_dt"]
    _assert_symmetric(bsr, nn, 3)
