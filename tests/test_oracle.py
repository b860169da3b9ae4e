"""CPU tests of the oracle itself: invariants and an independent derivation of every element matrix
(B^T D B from the inverse of the P1 basis matrix — shares no code or formula with oracle/oracle.cc)."""
import numpy as np
import pytest
import scipy.sparse as sp

from arcanefem_b200 import mesh as gen
from oracle import oracle as orc
from tests.util import MESHES, op_params


def p1_gradients(X):
    """Gradients of the P1 basis on a simplex with vertices X [nv, dim] (rows of inv([1 X])[1:, :].T) and its measure."""
    nv, dim = X.shape
    M = np.hstack([np.ones((nv, 1)), X])
    G = np.linalg.inv(M)[1:, :].T  # [nv, dim]
    meas = abs(np.linalg.det(M)) / (2.0 if dim == 2 else 6.0)
    return G, meas


def independent_ke(kind, X, bs, params, cell):
    nv, dim = X.shape
    G, meas = p1_gradients(X)
    if kind == "laplace":
        return meas * G @ G.T
    if kind.startswith("heat"):
        lam = params["cell_lambda"][cell] if params.get("cell_lambda") is not None else params["lambda_"]
        mass = meas / ((dim + 1) * (dim + 2)) * (np.ones((nv, nv)) + np.eye(nv))
        return lam * meas * G @ G.T + mass * params["inv_dt"]
    lam, mu = params["lambda_"], params["mu2"] / 2
    nstrain = 3 if dim == 2 else 6
    B = np.zeros((nstrain, nv * dim))
    for a in range(nv):
        if dim == 2:
            B[0, 2 * a] = G[a, 0]
            B[1, 2 * a + 1] = G[a, 1]
            B[2, 2 * a] = G[a, 1]
            B[2, 2 * a + 1] = G[a, 0]
        else:
            B[0, 3 * a] = G[a, 0]
            B[1, 3 * a + 1] = G[a, 1]
            B[2, 3 * a + 2] = G[a, 2]
            B[3, 3 * a], B[3, 3 * a + 1] = G[a, 1], G[a, 0]
            B[4, 3 * a + 1], B[4, 3 * a + 2] = G[a, 2], G[a, 1]
            B[5, 3 * a], B[5, 3 * a + 2] = G[a, 2], G[a, 0]
    D = np.zeros((nstrain, nstrain))
    D[:dim, :dim] = lam
    D[:dim, :dim] += 2 * mu * np.eye(dim)
    D[dim:, dim:] = mu * np.eye(nstrain - dim)
    return meas * B.T @ D @ B


@pytest.mark.parametrize("mesh_name,bs,kind", [("cube5", 1, "laplace"), ("square9_shuffled", 1, "laplace"), ("square9_shuffled", 2, "elasticity"),
                                                ("cube7_shuffled", 3, "elasticity"), ("square20", 1, "heat"), ("square20", 1, "heat_cell"),
                                                ("cube5", 1, "heat"), ("cube7_shuffled", 1, "heat_cell")])
def test_element_matrices_against_independent_derivation(mesh_name, bs, kind):
    dim, cells, coords = MESHES[mesh_name]()
    params = op_params(kind, nb_cell=cells.shape[0])
    rng = np.random.default_rng(0)
    for cell in rng.choice(cells.shape[0], size=min(40, cells.shape[0]), replace=False):
        Ke = orc.element_matrix(cells, coords, int(cell), bs=bs, **params)
        ref = independent_ke(kind, coords[cells[cell]][:, :dim], bs, params, cell)
        assert np.abs(Ke - ref).max() <= 1e-11 * np.abs(ref).max()


@pytest.mark.parametrize("mesh_name", sorted(MESHES))
def test_pattern_invariants(mesh_name):
    dim, cells, coords = MESHES[mesh_name]()
    nn = coords.shape[0]
    ri, co = orc.build_pattern(cells, nn)
    # unique undirected edges from the cells
    nv = cells.shape[1]
    pairs = [(a, b) for a in range(nv) for b in range(a + 1, nv)]
    e = np.concatenate([np.sort(cells[:, p], axis=1) for p in pairs])
    ue = np.unique(e, axis=0)
    assert co.shape[0] == 2 * ue.shape[0] + nn  # femutils/BSRFormat.h:401
    for r in range(nn):
        row = co[ri[r]:ri[r + 1]]
        assert row[0] == r  # diagonal first (BSRFormat.h:712-717)
        assert (np.diff(row[1:]) > 0).all()  # then ascending (sorted keys processed in order)
    A = sp.csr_matrix((np.ones(co.shape[0]), co, ri), shape=(nn, nn))
    assert (A != A.T).nnz == 0


def test_kuhn_cube_counts_match_survey_table():
    for N in (1, 3, 6):
        cells, coords = gen.cube_kuhn(N)
        nc, nn, e, nnzb = gen.kuhn_counts(N)
        ri, co = orc.build_pattern(cells, nn)
        assert (cells.shape[0], coords.shape[0], co.shape[0]) == (nc, nn, nnzb)
    assert gen.kuhn_counts(203) == (50192562, 8489664, 58929479, 126348622)
    assert gen.kuhn_counts(149)[3] == 50086798  # config #3


@pytest.mark.parametrize("mesh_name,bs,kind", [("cube5", 1, "laplace"), ("square9_shuffled", 2, "elasticity"), ("cube7_shuffled", 3, "elasticity")])
def test_assembly_layouts_and_variants_agree(mesh_name, bs, kind):
    dim, cells, coords = MESHES[mesh_name]()
    params = op_params(kind)
    nn = coords.shape[0]
    ri, co = orc.build_pattern(cells, nn)
    v_blk, nf = orc.assemble(cells, coords, ri, co, bs=bs, per_block=True, **params)
    v_row, _ = orc.assemble(cells, coords, ri, co, bs=bs, per_block=False, **params)
    v_nw, _ = orc.assemble(cells, coords, ri, co, bs=bs, per_block=False, nodewise=True, **params)
    assert nf == 0
    rows, nbcol, cols = orc.to_csr(bs, ri, co)
    A_row = sp.csr_matrix((v_row, cols, rows))
    # per-block layout -> scipy BSR
    A_blk = sp.bsr_matrix((v_blk.reshape(-1, bs, bs), co, ri), shape=(nn * bs, nn * bs)).tocsr()
    assert abs(A_row - A_blk).max() == 0
    assert np.abs(v_nw - v_row).max() <= 1e-13 * np.abs(v_row).max()
    # symmetric, and rigid translations are in the kernel (rows sum to zero per component)
    assert abs(A_row - A_row.T).max() <= 1e-12 * abs(A_row).max()
    for i in range(bs):
        t = np.zeros(nn * bs)
        t[i::bs] = 1.0
        assert np.abs(A_row @ t).max() <= 1e-11 * abs(A_row).max()
    # independent COO assembly of the independent element matrices
    acc = sp.lil_matrix((nn * bs, nn * bs))
    for c in range(cells.shape[0]):
        Ke = independent_ke(kind, coords[cells[c]][:, :dim], bs, params, c)
        dofs = (cells[c][:, None] * bs + np.arange(bs)[None, :]).ravel()
        acc[np.ix_(dofs, dofs)] += Ke
    assert abs(A_row - acc.tocsr()).max() <= 1e-10 * abs(A_row).max()


def test_dirichlet_semantics_small_system():
    """Penalty / row / row+column elimination all impose u = g: solve and compare."""
    import scipy.sparse.linalg as spl
    dim, cells, coords = MESHES["square9_shuffled"]()
    nn = coords.shape[0]
    ri, co = orc.build_pattern(cells, nn)
    group = gen.boundary_nodes(coords, 1, 0.0)
    sols = []
    for mode in ("penalty", "weak", "row", "rowcol"):
        vals, _ = orc.assemble(cells, coords, ri, co)
        rows, nbcol, cols = orc.to_csr(1, ri, co)
        rhs = np.zeros(nn)
        orc.rhs_constant_source(cells, coords, 1, 5.5, rhs)
        finfo, fval, einfo, ev = np.zeros(nn, np.uint8), np.zeros(nn), np.zeros(nn, np.uint8), np.zeros(nn)
        if mode in ("penalty", "weak"):
            orc.dirichlet_flag_penalty(1, 0, 0.5, 1e30, group, finfo, fval, rhs, weak=(mode == "weak"))
        else:
            orc.dirichlet_flag_elimination(1, 0, 0.5, 1 if mode == "row" else 2, group, einfo, ev)
        orc.dirichlet_apply(rows, cols, vals, einfo, ev, finfo, fval, rhs)
        A = sp.csr_matrix((vals, cols, rows))
        if mode == "rowcol":
            assert abs(A - A.T).max() <= 1e-12 * abs(A).max()  # row+column keeps symmetry
        u = spl.spsolve(A.tocsc(), rhs)
        assert np.allclose(u[group], 0.5, rtol=1e-12)
        sols.append(u)
    for a, b in zip(sols[:-1], sols[1:]):
        assert np.allclose(a, b, rtol=1e-9, atol=1e-12)


def test_legacy_cpu_baseline_equals_bsr_assembly():
    cells, coords = gen.cube_kuhn(6)
    ri, co = orc.build_pattern(cells, coords.shape[0])
    ref, _ = orc.assemble(cells, coords, ri, co)
    for threads in (1, 4):
        v = np.zeros_like(ref)
        orc.legacy_assemble_csr(cells, coords, ri, co, v, threads=threads)
        assert np.abs(v - ref).max() <= 1e-13 * np.abs(ref).max()


@pytest.mark.parametrize("N", [1, 2, 3, 6, 11])
def test_closed_form_patterns_equal_the_edge_key_algorithm(N):
    """gen.kuhn_pattern / gen.square_pattern (index arithmetic on the lattice) = the oracle's K1-K5 restatement
    (femutils/BSRFormat.h:602-744), bit for bit.  The full-size GPU tests (tests/test_gpu_zz_fullsize.py) use the closed
    forms, where the sequential edge-key sort would take minutes."""
    cells, coords = gen.cube_kuhn(N)
    ri, co = orc.build_pattern(cells, coords.shape[0])
    r2, c2 = gen.kuhn_pattern(N)
    np.testing.assert_array_equal(ri, r2)
    np.testing.assert_array_equal(co, c2)
    assert co.shape[0] == gen.kuhn_counts(N)[3]
    cells, coords = gen.square(N)
    ri, co = orc.build_pattern(cells, coords.shape[0])
    r2, c2 = gen.square_pattern(N)
    np.testing.assert_array_equal(ri, r2)
    np.testing.assert_array_equal(co, c2)


@pytest.mark.parametrize("mesh_name", ["square9_shuffled", "square20", "cube5", "cube7_shuffled"])
def test_heat_boundary_and_transient_terms_against_numpy(mesh_name):
    """Convection face term (modules/heat/FemModule.cc:628-672,813-840) and the (T_old / dt) source (:480-495), restated
    with numpy from the closed forms h |f| / (Nvf (Nvf + 1)) (1 + delta) and field * lumped mass."""
    dim, cells, coords = MESHES[mesh_name]()
    nn = coords.shape[0]
    ri, co = orc.build_pattern(cells, nn)
    faces = gen.boundary_faces(cells, coords, 0, coords[:, 0].max())
    assert faces.shape[0] > 0
    vals = np.zeros(co.shape[0])
    assert orc.add_face_mass(faces, coords, 3.5, ri, co, vals) == 0
    rows = np.repeat(np.arange(nn), np.diff(ri))
    got = sp.csr_matrix((vals, (rows, co)), shape=(nn, nn)).toarray()
    ref = np.zeros((nn, nn))
    nvf = faces.shape[1]
    for f in faces:
        X = coords[f]
        meas = np.hypot(*(X[1] - X[0])[:2]) if nvf == 2 else 0.5 * np.linalg.norm(np.cross(X[1] - X[0], X[2] - X[0]))
        ref[np.ix_(f, f)] += 3.5 * meas / (nvf * (nvf + 1)) * (np.ones((nvf, nvf)) + np.eye(nvf))
    np.testing.assert_allclose(got, ref, rtol=1e-13, atol=1e-15 * np.abs(ref).max())
    # a face edge outside the pattern is reported, not written anywhere
    far = np.array([[0, nn - 1] + ([1] if nvf == 3 else [])], dtype=np.int32)
    before = vals.copy()
    assert orc.add_face_mass(far, coords, 1.0, ri, co, vals) > 0
    assert np.count_nonzero(vals != before) <= nvf * nvf
    # transient source: field * scale * (sum of |cell| / Nv around the node), fixed nodes skipped
    rng = np.random.default_rng(3)
    field = rng.uniform(280.0, 320.0, size=nn)
    skip = (rng.uniform(size=nn) < 0.2).astype(np.uint8)
    rhs = np.zeros(nn)
    orc.rhs_nodal_source(cells, coords, 1, 10.0, field, rhs, node_skip=skip)
    X = coords[cells]
    if dim == 2:
        e1, e2 = X[:, 1] - X[:, 0], X[:, 2] - X[:, 0]
        meas = 0.5 * (e1[:, 0] * e2[:, 1] - e2[:, 0] * e1[:, 1])
    else:
        meas = np.abs(np.einsum("ij,ij->i", X[:, 1] - X[:, 0], np.cross(X[:, 2] - X[:, 0], X[:, 3] - X[:, 0]))) / 6.0
    lumped = np.zeros(nn)
    np.add.at(lumped, cells.ravel(), np.repeat(meas / cells.shape[1], cells.shape[1]))
    np.testing.assert_allclose(rhs, np.where(skip, 0.0, field * 10.0 * lumped), rtol=1e-12, atol=1e-13 * np.abs(rhs).max())
