"""Parity of the CUDA path (through the C ABI) against the CPU oracle.  Needs a B200: run with -m gpu.

Bars (BASELINE.json north_star): sparsity pattern, indices and Dirichlet masks bit-exact; matrix and RHS values
within 1e-12 relative per entry (tests/util.py states the rule, with the row-scale floor for cancelling entries).
"""
import numpy as np
import pytest

from tests import util
from tests.util import MESHES, assert_values_close, assert_vec_close, op_params

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _fem():
    import arcanefem_b200 as fem
    return fem


def _setup(mesh_name, bs, use_csr, variant, node_is_own=None, fan_mode=0, sparsity_build=0):
    """variant: b2f_assembly_variant; 31 / 32 = node gather (3) with the select / lockstep walk forced (b2f_fan_mode)."""
    fem = _fem()
    from arcanefem_b200.bsr import Mesh
    dim, cells, coords = MESHES[mesh_name]()
    m = Mesh.from_numpy(dim, cells, coords, node_is_own)
    bsr = fem.BSRFormat(bs)
    if variant in (31, 32):
        variant, fan_mode = 3, variant - 30
    bsr.initialize(m, use_csr, assembly_variant=variant, fan_mode=fan_mode, sparsity_build=sparsity_build)
    bsr.computeSparsity()
    return bsr, m, dim, cells, coords


def _np(t):
    return t.detach().cpu().numpy()


def _va(variant):
    """initialize() keywords of a test variant id (31 / 32: node gather with the select / lockstep walk forced)."""
    if variant in (31, 32):
        return dict(assembly_variant=3, fan_mode=variant - 30)
    return dict(assembly_variant=variant)


VARIANTS = [1, 2, 3, 4, 31, 32]  # 31 / 32: node gather with the select / lockstep walk forced  # cell atomic, cell atomic warp-aggregated, node gather (fan walk), node gather (slot words)


@pytest.mark.parametrize("sparsity_build", [0, 1, 2])
@pytest.mark.parametrize("mesh_name", sorted(MESHES))
def test_sparsity_pattern_bit_exact(mesh_name, sparsity_build):
    """Both builds (1: the reference's global edge sort; 2: row-wise from the node->cell lists, the default) give the
    oracle's pattern bit for bit."""
    from oracle import oracle as orc
    bsr, m, dim, cells, coords = _setup(mesh_name, 1, True, 3, sparsity_build=sparsity_build)
    assert bsr.stats()["sparsity_by_lists"] == (0 if sparsity_build == 1 else 1)
    ri, co = orc.build_pattern(cells, coords.shape[0])
    mat = bsr.matrix()
    assert mat.nbRow() == coords.shape[0]
    assert mat.nbCol() == co.shape[0]
    assert mat.nbNz() == co.shape[0]
    np.testing.assert_array_equal(_np(mat.rowIndexWithEnd()), ri)
    np.testing.assert_array_equal(_np(mat.columns()), co)
    np.testing.assert_array_equal(_np(mat.nbNzPerRow()), orc.nz_per_row(ri))
    # nnzb = 2E + Nn (femutils/BSRFormat.h:401)
    assert mat.nbCol() == 2 * mat.nbEdge() + mat.nbRow()


CASES = [
    # (mesh, bs, operator)
    ("cube1", 1, "laplace"), ("cube2", 1, "laplace"), ("cube5", 1, "laplace"), ("cube5_flat", 1, "laplace"),
    ("cube7_shuffled", 1, "laplace"), ("cube12", 1, "laplace"),
    ("square1", 1, "laplace"), ("square3", 1, "laplace"), ("square9_shuffled", 1, "laplace"), ("square20", 1, "laplace"),
    ("square3", 2, "elasticity"), ("square9_shuffled", 2, "elasticity"), ("square20", 2, "elasticity"),
    ("cube2", 3, "elasticity"), ("cube7_shuffled", 3, "elasticity"), ("cube12", 3, "elasticity"),
    ("square9_shuffled", 1, "heat"), ("square20", 1, "heat_cell"), ("cube5", 1, "heat"), ("cube7_shuffled", 1, "heat_cell"),
    # negatively oriented cells (the reference takes |det| for tets and the signed area for the heat triangle)
    ("cube4_mixed_orientation", 1, "laplace"), ("cube4_mixed_orientation", 3, "elasticity"), ("cube4_mixed_orientation", 1, "heat_cell"),
    ("square7_mixed_orientation", 1, "laplace"), ("square7_mixed_orientation", 2, "elasticity"), ("square7_mixed_orientation", 1, "heat_cell"),
]


def _operator(kind, params, device_cell_lambda):
    fem = _fem()
    if kind == "laplace":
        return fem.Operator.laplace()
    if kind == "elasticity":
        return fem.Operator.elasticity(21.0e5, 0.28)
    return fem.Operator.heat(params["lambda_"], 0.1, cell_lambda=device_cell_lambda)


@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("use_csr", [False, True])
@pytest.mark.parametrize("mesh_name,bs,kind", CASES)
def test_assembled_values_match_oracle(mesh_name, bs, kind, use_csr, variant):
    from oracle import oracle as orc
    bsr, m, dim, cells, coords = _setup(mesh_name, bs, use_csr, variant)
    params = op_params(kind, nb_cell=cells.shape[0])
    ri, co = orc.build_pattern(cells, coords.shape[0])
    ref, nf = orc.assemble(cells, coords, ri, co, bs=bs, per_block=not use_csr, **params)
    assert nf == 0
    cl = params.get("cell_lambda")
    dcl = torch.as_tensor(cl).cuda() if cl is not None else None
    op = _operator(kind, params, dcl)
    bsr.assembleBilinear(op)
    torch.cuda.synchronize()
    got = _np(bsr.matrix().values())
    assert_values_close(got, ref, ri, bs, what=f"{mesh_name}/{kind}/bs{bs}/csr{use_csr}/v{variant}")
    st = bsr.stats()
    assert st["columns_not_found"] == 0
    # assembling twice overwrites (zero-fill semantics), accumulate=True adds
    bsr.assembleBilinear(op)
    torch.cuda.synchronize()
    assert_values_close(_np(bsr.matrix().values()), ref, ri, bs, what="second assembly")
    bsr.assembleBilinear(op, accumulate=True)
    torch.cuda.synchronize()
    assert_values_close(_np(bsr.matrix().values()), 2.0 * ref, ri, bs, what="accumulate")
    bsr.resetMatrixValues()
    torch.cuda.synchronize()
    assert not _np(bsr.matrix().values()).any()


def test_node_gather_is_deterministic_and_symmetric():
    bsr, m, dim, cells, coords = _setup("cube12", 1, True, 3)
    fem = _fem()
    bsr.assembleBilinear(fem.Operator.laplace())
    a = _np(bsr.matrix().values()).copy()
    bsr.assembleBilinear(fem.Operator.laplace())
    b = _np(bsr.matrix().values())
    np.testing.assert_array_equal(a, b)  # atomic-free: bitwise reproducible
    import scipy.sparse as sp
    mat = bsr.matrix()
    A = sp.csr_matrix((a, _np(mat.columns()), _np(mat.rowIndexWithEnd())))
    assert abs(A - A.T).max() <= 1e-12 * abs(A).max()
    assert np.abs(A @ np.ones(A.shape[0])).max() <= 1e-12 * abs(A).max()


@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("mesh_name,bs,kind", [("cube5", 1, "laplace"), ("square9_shuffled", 2, "elasticity"), ("cube7_shuffled", 3, "elasticity")])
def test_rows_of_ghost_nodes_stay_zero(mesh_name, bs, kind, variant):
    """isOwn(row_node) rule, femutils/BSRFormat.h:815,870,974,1041."""
    from oracle import oracle as orc
    dim, cells, coords = MESHES[mesh_name]()
    rng = np.random.default_rng(3)
    own = (rng.uniform(size=coords.shape[0]) > 0.3).astype(np.uint8)
    bsr, m, dim, cells, coords = _setup(mesh_name, bs, True, variant, node_is_own=own)
    params = op_params(kind)
    ri, co = orc.build_pattern(cells, coords.shape[0])
    ref, _ = orc.assemble(cells, coords, ri, co, bs=bs, per_block=False, node_is_own=own, **params)
    bsr.assembleBilinear(_operator(kind, params, None))
    torch.cuda.synchronize()
    got = _np(bsr.matrix().values())
    assert_values_close(got, ref, ri, bs)
    ghost_rows = np.nonzero(own == 0)[0]
    for r in ghost_rows[:50]:
        assert not got[ri[r] * bs * bs:ri[r + 1] * bs * bs].any()


@pytest.mark.parametrize("mesh_name,bs,kind", [("cube5", 1, "laplace"), ("square9_shuffled", 2, "elasticity"), ("cube7_shuffled", 3, "elasticity")])
def test_to_csr_and_spmv(mesh_name, bs, kind):
    from oracle import oracle as orc
    bsr, m, dim, cells, coords = _setup(mesh_name, bs, True, 3)
    params = op_params(kind)
    ri, co = orc.build_pattern(cells, coords.shape[0])
    ref, _ = orc.assemble(cells, coords, ri, co, bs=bs, per_block=False, **params)
    rows, nbcol, cols = orc.to_csr(bs, ri, co)
    bsr.assembleBilinear(_operator(kind, params, None))
    ls = _fem().DoFLinearSystem()
    bsr.toLinearSystem(ls)
    torch.cuda.synchronize()
    view = ls.getCSRValues()
    np.testing.assert_array_equal(_np(view.rowsWithEnd()), rows)
    np.testing.assert_array_equal(_np(view.rowsNbColumn()), nbcol)
    np.testing.assert_array_equal(_np(view.columns()), cols)
    assert view.nbRow() == coords.shape[0] * bs and view.nbValue() == ref.shape[0]
    assert view.values().data_ptr() == bsr.matrix().values().data_ptr()  # zero copy
    rng = np.random.default_rng(11)
    x = rng.standard_normal(coords.shape[0] * bs)
    y_ref = orc.csr_spmv(rows, cols, ref, x)
    y = _np(bsr.spmv(torch.as_tensor(x).cuda()))
    assert_vec_close(y, y_ref, rtol=1e-11, what="spmv")
    # per-block layout gives the same product
    bsr2, *_ = _setup(mesh_name, bs, False, 3)
    bsr2.assembleBilinear(_operator(kind, params, None))
    y2 = _np(bsr2.spmv(torch.as_tensor(x).cuda()))
    assert_vec_close(y2, y_ref, rtol=1e-11, what="spmv per-block")


@pytest.mark.parametrize("mesh_name", ["cube5", "cube7_shuffled", "square9_shuffled", "square20"])
def test_rhs_source_and_neumann(mesh_name):
    from oracle import oracle as orc
    from arcanefem_b200 import mesh as gen
    bsr, m, dim, cells, coords = _setup(mesh_name, 1, True, 3)
    nn = coords.shape[0]
    L = coords[:, 0].max()
    faces = gen.boundary_faces(cells, coords, 0, L)
    assert faces.shape[0] > 0
    outside = (np.arange(faces.shape[0]) % 3 != 0).astype(np.uint8)
    ref = np.zeros(nn)
    orc.rhs_constant_source(cells, coords, 1, 5.5, ref)
    orc.rhs_neumann(faces, coords, 1, ref, value=2.0)
    orc.rhs_neumann(faces, coords, 1, ref, value=None, value_xyz=(1.5, -0.7, 0.3), face_outside=outside)
    ls = _fem().DoFLinearSystem()
    bsr.toLinearSystem(ls)
    ls.clearValues()
    bsr.applyConstantSourceToRhs(5.5)
    dfaces = torch.as_tensor(faces).cuda()
    bsr.applyNeumannToRhs(dfaces, value=2.0)
    bsr.applyNeumannToRhs(dfaces, value_xyz=(1.5, -0.7, 0.3), face_outside=torch.as_tensor(outside).cuda())
    torch.cuda.synchronize()
    assert_vec_close(_np(ls.rhsVariable()), ref, what="rhs")


@pytest.mark.parametrize("variant", [3, 31, 32, 1])
@pytest.mark.parametrize("mesh_name,bs,kind", [("cube5", 1, "laplace"), ("cube7_shuffled", 1, "laplace"), ("cube7_shuffled", 1, "heat_cell"), ("cube4_mixed_orientation", 3, "elasticity"),
                                                ("square9_shuffled", 1, "laplace"), ("square20", 2, "elasticity"), ("square7_mixed_orientation", 1, "heat_cell")])
def test_fused_source_rhs(mesh_name, bs, kind, variant):
    """b2f_assemble_with_source = assembleBilinear + applyConstantSourceToRhs: same matrix, same rhs as the two separate calls
    (variant 3 sums the cell measures along the fan walk; variant 1 takes the two-kernel fallback)."""
    from oracle import oracle as orc
    bsr, m, dim, cells, coords = _setup(mesh_name, bs, True, variant)
    nn = coords.shape[0]
    params = op_params(kind, nb_cell=cells.shape[0])
    cl = params.get("cell_lambda")
    lam = torch.as_tensor(cl).cuda() if cl is not None else None
    ref = np.zeros(nn * bs)
    orc.rhs_constant_source(cells, coords, bs, 5.5, ref)
    ls = _fem().DoFLinearSystem()
    bsr.toLinearSystem(ls)
    ls.clearValues()
    bsr.assembleBilinear(_operator(kind, params, lam))
    torch.cuda.synchronize()
    v_plain = _np(bsr.matrix().values()).copy()
    bsr.assembleBilinearWithSource(_operator(kind, params, lam), 5.5)
    torch.cuda.synchronize()
    if variant in (3, 31, 32):
        assert np.array_equal(_np(bsr.matrix().values()), v_plain)
    else:
        assert_vec_close(_np(bsr.matrix().values()), v_plain, what="values")
    assert_vec_close(_np(ls.rhsVariable()), ref, what="fused rhs")
    # a second call adds the source again (rhs is +=, like the reference) and accumulate mode works too
    bsr.assembleBilinearWithSource(_operator(kind, params, lam), 5.5, accumulate=True)
    torch.cuda.synchronize()
    assert_vec_close(_np(ls.rhsVariable()), 2 * ref, what="fused rhs twice")
    assert_vec_close(_np(bsr.matrix().values()), 2 * v_plain, what="accumulated values")


def _dirichlet_case(mesh_name, bs, kind, mode):
    """mode: 'penalty' | 'row' | 'rowcol' — returns (got values, got rhs, masks...) vs oracle."""
    from oracle import oracle as orc
    from arcanefem_b200 import mesh as gen
    bsr, m, dim, cells, coords = _setup(mesh_name, bs, True, 3)
    params = op_params(kind)
    nn = coords.shape[0]
    ndof = nn * bs
    ri, co = orc.build_pattern(cells, nn)
    vals, _ = orc.assemble(cells, coords, ri, co, bs=bs, per_block=False, **params)
    rows, nbcol, cols = orc.to_csr(bs, ri, co)
    axis = dim - 1
    group = gen.boundary_nodes(coords, axis, 0.0)  # "horizontal"
    assert group.size > 0
    rhs = np.zeros(ndof)
    orc.rhs_constant_source(cells, coords, bs, 5.5, rhs)
    finfo = np.zeros(ndof, np.uint8)
    fval = np.zeros(ndof)
    einfo = np.zeros(ndof, np.uint8)
    evalue = np.zeros(ndof)
    comps = [(0, 0.5)] if bs == 1 else [(0, 0.5), (bs - 1, -0.25)]  # vectorial: one call per non-"NULL" component
    for dof_index, g in comps:
        if mode in ("penalty", "weak"):
            orc.dirichlet_flag_penalty(bs, dof_index, g, 1.0e30 if mode == "penalty" else 1.0e8, group, finfo, fval, rhs, weak=(mode == "weak"))
        else:
            orc.dirichlet_flag_elimination(bs, dof_index, g, 1 if mode == "row" else 2, group, einfo, evalue)
    orc.dirichlet_apply(rows, cols, vals, einfo, evalue, finfo, fval, rhs)

    ls = _fem().DoFLinearSystem()
    bsr.assembleBilinear(_operator(kind, params, None))
    bsr.toLinearSystem(ls)
    ls.clearValues()
    bsr.applyConstantSourceToRhs(5.5)
    dgroup = torch.as_tensor(group).cuda()
    for dof_index, g in comps:
        if mode == "penalty":
            bsr.applyDirichletViaPenalty(dgroup, g, 1.0e30, dof_index)
        elif mode == "weak":
            bsr.applyDirichletViaWeakPenalty(dgroup, g, 1.0e8, dof_index)
        else:
            bsr.applyDirichletViaRowElimination(dgroup, g, dof_index, mode=1 if mode == "row" else 2)
    ls.applyDirichlet()
    torch.cuda.synchronize()
    np.testing.assert_array_equal(_np(ls.getForcedInfo()), finfo)  # masks bit-exact
    np.testing.assert_array_equal(_np(ls.getEliminationInfo()), einfo)
    np.testing.assert_array_equal(_np(ls.getForcedValue()), fval)
    np.testing.assert_array_equal(_np(ls.getEliminationValue()), evalue)
    assert_values_close(_np(bsr.matrix().values()), vals, ri, bs, what=f"dirichlet {mode} values")
    assert_vec_close(_np(ls.rhsVariable()), rhs, what=f"dirichlet {mode} rhs")
    return _np(bsr.matrix().values()), vals


@pytest.mark.parametrize("mode", ["penalty", "weak", "row", "rowcol"])
@pytest.mark.parametrize("mesh_name,bs,kind", [("cube5", 1, "laplace"), ("cube7_shuffled", 1, "laplace"), ("square9_shuffled", 1, "laplace"),
                                                ("square9_shuffled", 2, "elasticity"), ("cube7_shuffled", 3, "elasticity")])
def test_dirichlet_on_device(mesh_name, bs, kind, mode):
    got, ref = _dirichlet_case(mesh_name, bs, kind, mode)
    if mode == "penalty":
        assert (got == 1.0e30).sum() == (ref == 1.0e30).sum() > 0  # exact stores of P


def test_set_get_value_and_dump(tmp_path):
    fem = _fem()
    bsr, m, dim, cells, coords = _setup("square3", 2, True, 3)
    bsr.assembleBilinear(fem.Operator.elasticity(21.0e5, 0.28))
    mat = bsr.matrix()
    # testlab's host-side penalty: bsr_matrix->setValue(dof, dof, Penalty), modules/testlab/FemModule.cc:677-678
    mat.setValue(5, 5, 1.0e31)
    assert mat.getValue(5, 5) == 1.0e31
    v = mat.getValue(4, 5)
    assert v == mat.getValue(5, 4) or abs(v - mat.getValue(5, 4)) <= 1e-12 * abs(v)
    with pytest.raises(fem.B2FError) as e:
        mat.getValue(0, 2 * (coords.shape[0] - 1))  # far corner is not a neighbour of node 0
    assert e.value.status == -5 and "Value not found" in e.value.message
    out = tmp_path / "m.txt"
    bsr.dumpMatrix(out)
    lines = out.read_text().splitlines()
    assert lines[0] == f"size :{mat.nbNz()}"
    assert len(lines[2].split()) == mat.nbCol() and len(lines[3].split()) == mat.nbNz()


def test_error_conventions():
    fem = _fem()
    from arcanefem_b200.bsr import Mesh
    dim, cells, coords = MESHES["square3"]()
    m = Mesh.from_numpy(dim, cells, coords)
    bsr = fem.BSRFormat(1)
    bsr.initialize(m, True)
    with pytest.raises(fem.B2FError) as e:
        bsr.assembleBilinear(fem.Operator.laplace())  # before computeSparsity
    assert e.value.status == -4
    bad = fem.BSRFormat(4)
    with pytest.raises(fem.B2FError) as e:
        bad.initialize(m, True)
    assert e.value.status == -2
    m1 = Mesh(1, m.cell_node, m.coords)
    with pytest.raises(fem.B2FError) as e:
        fem.BSRFormat(1).initialize(m1, True)
    assert e.value.status == -2 and "Only supports 2D and 3D" in e.value.message
    bsr.computeSparsity()
    with pytest.raises(fem.B2FError) as e:
        bsr.assembleBilinear(fem.Operator.elasticity(1.0, 0.3))  # BS mismatch
    assert e.value.status == -1
    # node id out of range is reported, not dereferenced
    cells_bad = cells.copy()
    cells_bad[0, 0] = coords.shape[0] + 5
    b2 = fem.BSRFormat(1)
    b2.initialize(Mesh.from_numpy(dim, cells_bad, coords), True)
    with pytest.raises(fem.B2FError) as e:
        b2.computeSparsity()
    assert e.value.status == -1
    # BS > 1 CSR view needs the per-row layout
    b3 = fem.BSRFormat(2)
    b3.initialize(m, False)
    b3.computeSparsity()
    with pytest.raises(fem.B2FError) as e:
        b3.matrix().toCsr()
    assert e.value.status == -4
    # the multi-GPU entry points refuse a handle without an exchange plan (b2f_ghost_plan), also the host-buffer one
    import ctypes as C
    from arcanefem_b200 import _lib
    lib = _lib.load_library()
    d = fem.Operator.laplace().desc(False)
    host = torch.zeros(coords.shape[0] * 3, dtype=torch.float64).pin_memory()
    assert lib.b2f_assemble_exchange(bsr._h, C.byref(d), None, None) == -4
    assert lib.b2f_assemble_exchange_host(bsr._h, C.byref(d), None, C.c_void_p(host.data_ptr()), None, None) == -4
    assert lib.b2f_assemble_exchange_host(bsr._h, C.byref(d), None, None, None, None) == -1
    assert lib.b2f_ghost_exchange(bsr._h, None, None) == -4


@pytest.mark.parametrize("variant", VARIANTS)
def test_medium_cube_parity(variant):
    """N = 40 (384 000 tets): oracle runs in about a second."""
    from oracle import oracle as orc
    from arcanefem_b200 import mesh as gen
    from arcanefem_b200.bsr import Mesh
    fem = _fem()
    cells, coords = gen.cube_kuhn(40)
    m = Mesh.from_numpy(3, cells, coords)
    bsr = fem.BSRFormat(1)
    bsr.initialize(m, True, **_va(variant))
    bsr.computeSparsity()
    ri, co = orc.build_pattern(cells, coords.shape[0])
    np.testing.assert_array_equal(_np(bsr.matrix().rowIndexWithEnd()), ri)
    np.testing.assert_array_equal(_np(bsr.matrix().columns()), co)
    nc, nn, e, nnzb = gen.kuhn_counts(40)
    assert bsr.matrix().nbEdge() == e and bsr.matrix().nbCol() == nnzb
    assert bsr.stats()["sparsity_by_lists"] == 1
    ref, _ = orc.assemble(cells, coords, ri, co)
    bsr.assembleBilinear(fem.Operator.laplace())
    torch.cuda.synchronize()
    assert_values_close(_np(bsr.matrix().values()), ref, ri, 1)
    st = bsr.stats()
    assert st["gather_fallback_tiles"] == 0
    assert st["fan_lockstep"] == (1 if variant == 32 else 0)
    if variant == 32:
        # the rows of a structured mesh mostly walk in step: few step bodies per warp step
        assert 1000 <= st["fan_lockstep_bodies_milli"] < 2000, st
    if variant in (3, 31, 32):
        # fan plan: one chain start per row on this mesh (every node link is a disc or a sphere), so
        # steps = node/cell pairs + rows
        assert st["fan_steps"] == 4 * nc + nn


@pytest.mark.parametrize("K", [500, 3000, 5000])
def test_long_rows_take_the_global_fallback_path(K):
    """A fan of triangles around one hub node: one row far longer than the shared-memory tile (K = 5000: also longer
    than the 10-bit column slots of the gather plan, so the search-based kernel runs)."""
    from oracle import oracle as orc
    from arcanefem_b200.bsr import Mesh
    fem = _fem()
    ang = np.linspace(0.0, 2 * np.pi, K, endpoint=False)
    coords = np.zeros((K + 1, 3))
    coords[1:, 0] = np.cos(ang)
    coords[1:, 1] = np.sin(ang)
    cells = np.stack([np.zeros(K, np.int32), 1 + np.arange(K, dtype=np.int32), 1 + (np.arange(K, dtype=np.int32) + 1) % K], axis=1).astype(np.int32)
    ri, co = orc.build_pattern(cells, K + 1)
    ref, _ = orc.assemble(cells, coords, ri, co)
    for variant in VARIANTS:
        bsr = fem.BSRFormat(1)
        bsr.initialize(Mesh.from_numpy(2, cells, coords), True, **_va(variant))
        bsr.computeSparsity()
        np.testing.assert_array_equal(_np(bsr.matrix().columns()), co)
        # the hub row has K neighbours: more than the row-wise build holds, the sort-based build takes over
        assert bsr.stats()["sparsity_by_lists"] == 0
        bsr.assembleBilinear(fem.Operator.laplace())
        torch.cuda.synchronize()
        assert_values_close(_np(bsr.matrix().values()), ref, ri, 1, rtol=1e-11, what=f"fan v{variant}")
        if variant in (3, 31, 32) and K >= 3000:
            assert bsr.stats()["gather_fallback_tiles"] >= 1
        # fused constant source on the same paths (oversized tile: the source term comes from the node's cell list)
        ls = fem.DoFLinearSystem()
        bsr.toLinearSystem(ls)
        ls.clearValues()
        bsr.assembleBilinearWithSource(fem.Operator.laplace(), 5.5)
        torch.cuda.synchronize()
        rhs_ref = np.zeros(K + 1)
        orc.rhs_constant_source(cells, coords, 1, 5.5, rhs_ref)
        assert_vec_close(_np(ls.rhsVariable()), rhs_ref, what=f"fused rhs v{variant}")
        assert_values_close(_np(bsr.matrix().values()), ref, ri, 1, rtol=1e-11, what=f"fan v{variant} with source")


def emulate_fan_plan(cells, ri, co, plan, lockstep=False):
    """Replay the fan-walk plan on the host: returns, per row, the list of (cell, (row, N, O, C) node ids, odd) it emits
    (lockstep form: (row, set f+1, set f+2, C))."""
    nv = cells.shape[1]
    fan_ptr = plan["fan_ptr"]
    steps = plan["fan_steps"].view(np.uint32)
    fcell = plan["fan_cell"].view(np.uint32)
    out = []
    for row in range(ri.shape[0] - 1):
        cols = co[ri[row]:ri[row + 1]]
        n = o = None
        emitted = []
        sets, free = [None, None, None], 3
        for q in range(fan_ptr[row], fan_ptr[row + 1]):
            w = int(steps[q])
            if lockstep:
                # three fixed sets; START: bits 20-21 = free set of the window that ends; N -> set 1, O -> set 2, free 0
                if w & 0x80000000:
                    assert ((w >> 20) & 3) == free, f"row {row}: START names free set {(w >> 20) & 3}, walk has {free}"
                    sets, free = [None, int(cols[w & 1023]), int(cols[(w >> 10) & 1023])], 0
                    continue
                code = (w >> 10) & 7
                assert code < 6 and code >> 1 == free, f"row {row}: step variant {code} with free set {free}"
                x, y = (free + 1) % 3, (free + 2) % 3
                c = int(cols[w & 1023])
                emitted.append((int(fcell[q]), (row, sets[x], sets[y], c), (w >> 13) & 1))
                sets[free] = c
                free = y if code & 1 else x
                sets[free] = None
                continue
            if w & 0x80000000:
                n = int(cols[w & 1023])
                o = int(cols[(w >> 10) & 1023]) if nv == 4 else None
                continue
            c = int(cols[w & 1023])
            assert n is not None
            verts = (row, n, o, c) if nv == 4 else (row, n, c)
            emitted.append((int(fcell[q]), verts, (w >> 11) & 1))
            if nv == 4:
                if not (w >> 10) & 1:  # O leaves: the old N becomes O
                    o = n
            n = c
        out.append(emitted)
    return out


@pytest.mark.parametrize("fan_mode", [1, 2])
@pytest.mark.parametrize("mesh_name", sorted(MESHES))
def test_fan_plan_visits_every_incident_cell_once(mesh_name, fan_mode):
    """The plan is a cover of every node's incident cells by face-adjacent walks; the orientation bit is the parity of
    the emitted vertex order relative to the cell's own.  Both forms of the step words (select walk, lockstep walk)."""
    bsr, m, dim, cells, coords = _setup(mesh_name, 1, True, 3, fan_mode=fan_mode)
    plan = {k: _np(v) for k, v in bsr.plan().items()}
    mat = bsr.matrix()
    ri, co = _np(mat.rowIndexWithEnd()), _np(mat.columns())
    st = bsr.stats()
    assert st["fan_lockstep"] == (1 if fan_mode == 2 and cells.shape[1] == 4 else 0)
    rows = emulate_fan_plan(cells, ri, co, plan, lockstep=bool(st["fan_lockstep"]))
    incident = [[] for _ in range(coords.shape[0])]
    for c, vs in enumerate(cells):
        for v in vs:
            incident[int(v)].append(c)
    for row, emitted in enumerate(rows):
        assert sorted(e[0] for e in emitted) == sorted(incident[row]), f"row {row}"
        for cell, verts, odd in emitted:
            assert sorted(verts) == sorted(int(v) for v in cells[cell]), f"row {row} cell {cell}: {verts} vs {cells[cell]}"
            perm = [list(cells[cell]).index(v) for v in verts]
            inv = sum(1 for i in range(len(perm)) for j in range(i + 1, len(perm)) if perm[i] > perm[j])
            assert (inv & 1) == odd, f"row {row} cell {cell}"


@pytest.mark.parametrize("mesh_name", ["cube5", "cube7_shuffled", "cube4_mixed_orientation", "square9_shuffled", "square20"])
def test_fan_plan_marks_first_visits(mesh_name):
    """k_fan_first_flags: bit 13 of an emit word / bits 20, 21 of a START word are set exactly when the slot they load has
    not been loaded before in its row (the kernel then starts that accumulator from zero and skips clearing the slab), and
    every slot of a walked row is loaded at least once."""
    bsr, m, dim, cells, coords = _setup(mesh_name, 1, True, 3, fan_mode=1)
    plan = {k: _np(v) for k, v in bsr.plan().items()}
    ri = _np(bsr.matrix().rowIndexWithEnd())
    nv = cells.shape[1]
    fan_ptr, steps = plan["fan_ptr"], plan["fan_steps"].view(np.uint32)
    for row in range(ri.shape[0] - 1):
        seen = set()

        def check(slot, flag, what):
            assert bool(flag) == (slot not in seen), f"row {row}: {what} slot {slot}, flag {flag}, seen {sorted(seen)}"
            seen.add(slot)

        for q in range(fan_ptr[row], fan_ptr[row + 1]):
            w = int(steps[q])
            if w & 0x80000000:
                check(w & 1023, (w >> 20) & 1, "START N")
                if nv == 4:
                    check((w >> 10) & 1023, (w >> 21) & 1, "START O")
            else:
                check(w & 1023, (w >> 13) & 1, "emit")
        if fan_ptr[row + 1] > fan_ptr[row]:
            assert seen == set(range(1, int(ri[row + 1] - ri[row]))), f"row {row}: slots {sorted(seen)}"


@pytest.mark.parametrize("mesh_name,bs", [("cube5", 1), ("square9_shuffled", 2), ("cube7_shuffled", 3)])
def test_csr_global_numbering_on_device(mesh_name, bs):
    """Device replacement of the host renumbering loops of femutils/HypreDoFLinearSystem.cc:474-519."""
    bsr, m, dim, cells, coords = _setup(mesh_name, bs, True, 3)
    csr = bsr.matrix().toCsr()
    ndof = coords.shape[0] * bs
    rng = np.random.default_rng(9)
    numbering = (rng.permutation(ndof) + 1000).astype(np.int32)
    rows, cols = bsr.csrGlobalNumbering(torch.as_tensor(numbering).cuda())
    torch.cuda.synchronize()
    np.testing.assert_array_equal(_np(rows), numbering)
    np.testing.assert_array_equal(_np(cols), numbering[_np(csr.columns())])


@pytest.mark.parametrize("with_ghosts", [False, True])
@pytest.mark.parametrize("mesh_name,bs,kind", [("cube5", 1, "laplace"), ("cube7_shuffled", 1, "laplace"), ("square20", 2, "elasticity"), ("cube4_mixed_orientation", 3, "elasticity")])
def test_hypre_ij_hand_off_on_device(mesh_name, bs, kind, with_ghosts):
    """b2f_to_hypre_ij: the own rows of the CSR view with global ids, as HypreDoFLinearSystemImpl::solve() builds them on
    the host (femutils/HypreDoFLinearSystem.cc:474-519, :587-640) — checked against a numpy walk over the own dofs."""
    dim, cells, coords = MESHES[mesh_name]()
    nn = coords.shape[0]
    rng = np.random.default_rng(13)
    own = (rng.uniform(size=nn) > 0.3).astype(np.uint8) if with_ghosts else None
    bsr, m, dim, cells, coords = _setup(mesh_name, bs, True, 3, node_is_own=own)
    params = op_params(kind, nb_cell=cells.shape[0])
    ls = _fem().DoFLinearSystem()
    bsr.toLinearSystem(ls)
    ls.clearValues()
    bsr.assembleBilinearWithSource(_operator(kind, params, None), 5.5)
    torch.cuda.synchronize()
    csr = bsr.matrix().toCsr()
    rp, nbc, cl, va = _np(csr.rows()), _np(csr.rowsNbColumn()), _np(csr.columns()), _np(csr.values())
    rhs = _np(ls.rhsVariable())
    ndof = nn * bs
    numbering = (rng.permutation(ndof) + 77).astype(np.int32)
    for num in (None, numbering):
        ij = bsr.toHypreIJ(torch.as_tensor(num).cuda() if num is not None else None)
        torch.cuda.synchronize()
        ident = num if num is not None else np.arange(ndof, dtype=np.int32)
        e_ncols, e_rows, e_local, e_cols, e_vals, e_rhs = [], [], [], [], [], []
        for r in range(ndof):
            if own is not None and not own[r // bs]:
                continue
            e_ncols.append(nbc[r])
            e_rows.append(ident[r])
            e_local.append(r)
            seg = slice(rp[r], rp[r] + nbc[r])
            e_cols.extend(ident[cl[seg]])
            e_vals.extend(va[seg])
            e_rhs.append(rhs[r])
        np.testing.assert_array_equal(_np(ij["ncols"]), np.array(e_ncols, dtype=np.int32))
        np.testing.assert_array_equal(_np(ij["rows"]), np.array(e_rows, dtype=np.int32))
        np.testing.assert_array_equal(_np(ij["row_local"]), np.array(e_local, dtype=np.int32))
        np.testing.assert_array_equal(_np(ij["cols"]), np.array(e_cols, dtype=np.int32))
        np.testing.assert_array_equal(_np(ij["values"]), np.array(e_vals))
        np.testing.assert_array_equal(_np(ij["rhs"]), np.array(e_rhs))


@pytest.mark.parametrize("case", ["small", "structured", "shuffled", "heat_shuffled_cells"])
def test_host_entry_point_matches_device_path(case):
    """b2f_assemble_host (host coordinates in, host values out).  Small meshes take the one-shot path; from 65536 rows on
    the call is pipelined over row chunks (coordinate prefix up, chunk assembled, chunk values down): same values either
    way, also when the numbering has no locality and every chunk needs all coordinates."""
    from oracle import oracle as orc
    from arcanefem_b200 import mesh as gen
    from arcanefem_b200.bsr import Mesh
    fem = _fem()
    if case == "small":
        cells, coords = gen.cube_kuhn(6)
    elif case == "structured":
        cells, coords = gen.cube_kuhn(41)
    elif case == "shuffled":
        cells, coords = gen.cube_kuhn(41, shuffle_cells=42, shuffle_nodes=7)
    else:
        cells, coords = gen.cube_kuhn(41, shuffle_cells=5)
    nn = coords.shape[0]
    assert (nn >= 65536) == (case != "small")
    if case == "heat_shuffled_cells":
        lam = np.random.default_rng(3).uniform(0.5, 2.0, size=cells.shape[0])
        params = dict(op=orc.OP_HEAT, lambda_=0.0, inv_dt=10.0, cell_lambda=lam)
        dlam = torch.as_tensor(lam).cuda()
        op = fem.Operator.heat(0.0, 0.1, cell_lambda=dlam)
    else:
        params, op = dict(op=orc.OP_LAPLACE), fem.Operator.laplace()
    # the handle is created on a mesh whose device coordinates are garbage: the host entry point must use the host array
    bsr = fem.BSRFormat(1)
    bsr.initialize(Mesh.from_numpy(3, cells, np.full_like(coords, 1.0e9)), True, assembly_variant=3)
    bsr.computeSparsity()
    ri, co = orc.build_pattern(cells, nn)
    ref, nf = orc.assemble(cells, coords, ri, co, bs=1, per_block=False, **params)
    assert nf == 0
    coords_host = torch.as_tensor(coords).pin_memory()
    values_host = torch.full((ri[-1],), np.nan, dtype=torch.float64).pin_memory()
    for rep in range(2):
        values_host.fill_(float("nan"))
        bsr.assembleBilinearHost(op, coords_host, values_host)
        assert_values_close(values_host.numpy(), ref, ri, 1, what=f"{case} host values (call {rep})")
        np.testing.assert_array_equal(_np(bsr.matrix().values()), values_host.numpy())
    assert bsr.stats()["columns_not_found"] == 0
    # no read-back requested: device values only
    bsr.resetMatrixValues()
    bsr.assembleBilinearHost(op, coords_host, None)
    torch.cuda.synchronize()
    assert_values_close(_np(bsr.matrix().values()), ref, ri, 1, what=f"{case} device values")


@pytest.mark.parametrize("sparsity_build", [1, 2])
@pytest.mark.parametrize("variant", [1, 3, 4])
def test_empty_ragged_and_repeated_inputs(variant, sparsity_build):
    """Edge inputs: no cells at all (every row is its diagonal), nodes that no cell references, a single cell, the same
    cell listed twice (contributions add, the pattern does not change), a mesh with no nodes (refused)."""
    from oracle import oracle as orc
    from arcanefem_b200.bsr import Mesh
    fem = _fem()
    rng = np.random.default_rng(21)

    def run(dim, cells, coords, bs=1, op=None, params=None):
        nn = coords.shape[0]
        bsr = fem.BSRFormat(bs)
        bsr.initialize(Mesh.from_numpy(dim, cells, coords), True, assembly_variant=variant, sparsity_build=sparsity_build)
        bsr.computeSparsity()
        ri, co = orc.build_pattern(cells, nn)
        mat = bsr.matrix()
        assert mat.nbRow() == nn and mat.nbCol() == co.shape[0]
        np.testing.assert_array_equal(_np(mat.rowIndexWithEnd()), ri)
        np.testing.assert_array_equal(_np(mat.columns()), co)
        bsr.assembleBilinear(op or fem.Operator.laplace())
        torch.cuda.synchronize()
        ref, nf = orc.assemble(cells, coords, ri, co, bs=bs, per_block=False, **(params or {}))
        assert nf == 0 and bsr.stats()["columns_not_found"] == 0
        assert_values_close(_np(mat.values()), ref, ri, bs)
        return bsr

    # no cells: nnzb = Nn, all values zero
    run(3, np.zeros((0, 4), np.int32), rng.uniform(size=(7, 3)))
    run(2, np.zeros((0, 3), np.int32), rng.uniform(size=(5, 3)))
    # one cell, and nodes nobody references (rows 4..9 hold their diagonal only)
    b = run(3, np.array([[2, 0, 3, 1]], np.int32), rng.uniform(size=(10, 3)))
    assert _np(b.matrix().nbNzPerRow()).tolist() == [4, 4, 4, 4, 1, 1, 1, 1, 1, 1]
    # the same cell twice, next to another one
    c2 = np.array([[0, 1, 2], [0, 1, 2], [1, 3, 2]], np.int32)
    x2 = np.array([[0.0, 0.0, 0.0], [1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [1.1, 0.9, 0.0], [5.0, 5.0, 0.0]])
    run(2, c2, x2)
    run(2, c2, x2, bs=2, op=fem.Operator.elasticity(21.0e5, 0.28), params=op_params("elasticity"))
    c3 = np.array([[0, 1, 2, 3], [0, 1, 2, 3], [1, 2, 3, 4]], np.int32)
    x3 = rng.uniform(size=(6, 3))
    run(3, c3, x3)
    run(3, c3, x3, bs=3, op=fem.Operator.elasticity(21.0e5, 0.28), params=op_params("elasticity"))
    # a mesh with no nodes is refused like the reference does (ArgumentException, femutils/BSRFormat.h:121-125)
    with pytest.raises(fem.B2FError) as err:
        fem.BSRFormat(1).initialize(Mesh.from_numpy(3, np.zeros((0, 4), np.int32), np.zeros((0, 3))), True, assembly_variant=variant, sparsity_build=sparsity_build)
    assert err.value.status == -1 and "should be positive and not null" in err.value.message
