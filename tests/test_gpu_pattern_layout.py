"""The diagonal-last pattern layout (b2f_pattern_layout = B2F_PATTERN_DIAG_LAST), the layout of the reference's connectivity
("atomic-free") sparsity build, femutils/BSRFormat.h:479-578: neighbours first, the node itself last.  The reference lists
the neighbours in the order of Arcane's node-node / node-face connectivity, which cannot be reproduced outside Arcane
("only the per-row neighbour SET is pinned"); here they are listed by ascending id.  Pinned:
  * per-row neighbour set == the set of the sorted-edge build K1-K5 (:602-744), diagonal in the last slot, same row_index;
  * every kernel that addresses a row by slot agrees with the oracle once the row is rotated back: the three assembly
    variants, both value layouts, BS 1 / 2 / 3, Dirichlet (penalty, row, row + column), face mass, SpMV, CSR view, getValue.
"""
import numpy as np
import pytest

from tests.util import MESHES, assert_values_close, assert_vec_close, op_params
from tests.test_gpu_multirank import block_rows

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _np(t):
    return t.detach().cpu().numpy()


def _setup(mesh_name, bs, use_csr, variant, sparsity_build=0):
    import arcanefem_b200 as fem
    from arcanefem_b200.bsr import Mesh
    dim, cells, coords = MESHES[mesh_name]()
    bsr = fem.BSRFormat(bs)
    bsr.initialize(Mesh.from_numpy(dim, cells, coords), use_csr, assembly_variant=variant, sparsity_build=sparsity_build, pattern_layout=fem.bsr.PATTERN_DIAG_LAST)
    bsr.computeSparsity()
    return fem, bsr, dim, cells, coords


def _maps(ri):
    """(to_storage, to_canonical): storage_array == canonical_array[to_storage], canonical_array == storage_array[to_canonical].
    Canonical = diagonal first; storage = diagonal last: canonical slot 0 lives at k - 1, canonical slot x >= 1 at x - 1."""
    k = np.diff(ri)
    start = np.repeat(ri[:-1], k)
    pos = np.arange(ri[-1]) - start
    kk = np.repeat(k, k)
    return start + (pos + 1) % kk, start + (pos + kk - 1) % kk


def _rotation(ri):
    return _maps(ri)[0]


def _to_canonical(ri):
    return _maps(ri)[1]


@pytest.mark.parametrize("sparsity_build", [0, 1])
@pytest.mark.parametrize("mesh_name", ["cube1", "cube5", "cube7_shuffled", "square9_shuffled", "cube4_mixed_orientation"])
def test_neighbour_sets_equal_the_sorted_edge_build(mesh_name, sparsity_build):
    from oracle import oracle as orc
    fem, bsr, dim, cells, coords = _setup(mesh_name, 1, True, 3, sparsity_build)
    ri, co = orc.build_pattern(cells, coords.shape[0])  # K1-K5: diagonal first, ascending
    mat = bsr.matrix()
    np.testing.assert_array_equal(_np(mat.rowIndexWithEnd()), ri)
    got = _np(mat.columns())
    for r in range(coords.shape[0]):
        row = got[ri[r]:ri[r + 1]]
        assert row[-1] == r, "the node itself closes its row (femutils/BSRFormat.h:554-567)"
        assert set(row[:-1].tolist()) == set(co[ri[r] + 1:ri[r + 1]].tolist()), "same neighbour set as the sorted-edge build"
        assert (np.diff(row[:-1]) > 0).all()
    np.testing.assert_array_equal(got, co[_rotation(ri)])
    assert mat.nbCol() == 2 * mat.nbEdge() + mat.nbRow()


CASES = [("cube5", 1, "laplace"), ("cube7_shuffled", 1, "laplace"), ("square9_shuffled", 1, "laplace"), ("square20", 2, "elasticity"), ("cube7_shuffled", 3, "elasticity"),
         ("cube4_mixed_orientation", 3, "elasticity"), ("square20", 1, "heat_cell"), ("cube5", 1, "heat")]


@pytest.mark.parametrize("use_csr", [True, False])
@pytest.mark.parametrize("variant", [1, 2, 3])
@pytest.mark.parametrize("mesh_name,bs,kind", CASES)
def test_values_match_the_oracle_after_rotating_the_rows(mesh_name, bs, kind, variant, use_csr):
    from oracle import oracle as orc
    fem, bsr, dim, cells, coords = _setup(mesh_name, bs, use_csr, variant)
    okw = op_params(kind, cells.shape[0])
    ri, co = orc.build_pattern(cells, coords.shape[0])
    ref, _ = orc.assemble(cells, coords, ri, co, bs=bs, per_block=True, **okw)
    cl = torch.as_tensor(okw["cell_lambda"]).cuda() if okw.get("cell_lambda") is not None else None
    op = {"laplace": fem.Operator.laplace, "elasticity": lambda: fem.Operator.elasticity(21.0e5, 0.28)}.get(kind, lambda: fem.Operator.heat(okw.get("lambda_", 0.0), 0.1, cell_lambda=cl))()
    for _ in range(2):
        bsr.assembleBilinear(op)
    torch.cuda.synchronize()
    st = bsr.stats()
    assert st["columns_not_found"] == 0 and st["gather_fallback_tiles"] == 0
    got = block_rows(_np(bsr.matrix().values()), ri, bs, not use_csr)  # blocks in storage order
    canon = np.empty_like(got)
    canon[:] = got[_to_canonical(ri)]
    assert_values_close(canon.reshape(-1), ref, ri, bs, what=f"{mesh_name} bs={bs} {kind} variant {variant} diag-last")
    # accumulate mode doubles
    bsr.assembleBilinear(op, accumulate=True)
    torch.cuda.synchronize()
    twice = block_rows(_np(bsr.matrix().values()), ri, bs, not use_csr)[_to_canonical(ri)]
    assert_values_close(twice.reshape(-1), 2.0 * ref, ri, bs, what="accumulate")


@pytest.mark.parametrize("use_csr", [True, False])
@pytest.mark.parametrize("mesh_name,bs,kind", [("cube5", 1, "laplace"), ("cube7_shuffled", 3, "elasticity"), ("square20", 2, "elasticity")])
def test_dirichlet_spmv_csr_and_get_value(mesh_name, bs, kind, use_csr):
    """Everything downstream of the assembly on a diagonal-last matrix equals the same calls on a diagonal-first one."""
    import arcanefem_b200 as fem
    from arcanefem_b200 import mesh as gen
    from arcanefem_b200.bsr import Mesh
    dim, cells, coords = MESHES[mesh_name]()
    op = fem.Operator.laplace() if kind == "laplace" else fem.Operator.elasticity(21.0e5, 0.28)
    group = torch.as_tensor(gen.boundary_nodes(coords, dim - 1, coords[:, dim - 1].min())).cuda()
    results = []
    for layout in (0, 1):
        bsr = fem.BSRFormat(bs)
        bsr.initialize(Mesh.from_numpy(dim, cells, coords), use_csr, pattern_layout=layout)
        bsr.computeSparsity()
        ls = fem.DoFLinearSystem()
        bsr.toLinearSystem(ls)  # use_csr = False, BS > 1: the per-block branch (no CSR view, attached all the same)
        bsr.assembleBilinearWithSource(op, 5.5)
        ls.clearValues()
        bsr.applyDirichletViaPenalty(group[::3].contiguous(), 0.5, 1.0e30, dof_index=0)
        bsr.applyDirichletViaRowElimination(group[1::3].contiguous(), 0.25, dof_index=bs - 1)
        bsr.applyDirichletViaRowElimination(group[2::3].contiguous(), -0.75, dof_index=0, mode=fem.bsr.ELIMINATE_ROW_COLUMN)
        ls.applyDirichlet()
        if bs == 1 and dim == 3:
            faces = torch.as_tensor(gen.boundary_faces(cells, coords, 2, coords[:, 2].max())).cuda()
            bsr.addFaceMass(faces, 3.5)
        x = torch.as_tensor(np.random.default_rng(3).uniform(-1, 1, size=coords.shape[0] * bs)).cuda()
        y = bsr.spmv(x)
        torch.cuda.synchronize()
        mat = bsr.matrix()
        ri = _np(mat.rowIndexWithEnd())
        blocks = block_rows(_np(mat.values()), ri, bs, not use_csr)
        d = mat.getValue(bs * 7, bs * 7)
        csr = None
        if use_csr or bs == 1:
            c = mat.toCsr()
            csr = (_np(c.rowsWithEnd()), _np(c.columns()), _np(c.values()))
        results.append((ri, _np(mat.columns()), blocks, _np(y), _np(ls.rhsVariable()), d, csr))
    (ri, c0, b0, y0, r0, d0, csr0), (_, c1, b1, y1, r1, d1, csr1) = results
    rot = _rotation(ri)
    np.testing.assert_array_equal(c1, c0[rot])
    scale = np.abs(b0).max()
    finite = np.abs(b0) < 1e29  # penalty diagonals compared exactly
    assert np.abs(b1 - b0[rot])[finite[rot]].max() <= 1e-12 * scale
    np.testing.assert_array_equal(b1[~finite[rot]], b0[rot][~finite[rot]])
    assert_vec_close(r1, r0, what="rhs")
    assert abs(d1 - d0) <= 1e-12 * max(abs(d0), 1.0)
    big = np.abs(y0) < 1e25
    assert np.abs(y1 - y0)[big].max() <= 1e-11 * np.abs(y0[big]).max()
    if csr0 is not None:  # the CSR views describe the same scalar matrix
        import scipy.sparse as sp
        n = ri.shape[0] - 1
        A0 = sp.csr_matrix((np.where(np.abs(csr0[2]) > 1e29, 0.0, csr0[2]), csr0[1], csr0[0]), shape=(n * bs, n * bs))
        A1 = sp.csr_matrix((np.where(np.abs(csr1[2]) > 1e29, 0.0, csr1[2]), csr1[1], csr1[0]), shape=(n * bs, n * bs))
        assert abs(A0 - A1).max() <= 1e-12 * scale


def test_slot_word_kernel_refuses_the_layout():
    import arcanefem_b200 as fem
    from arcanefem_b200.bsr import Mesh
    from arcanefem_b200._lib import B2FError
    dim, cells, coords = MESHES["cube2"]()
    bsr = fem.BSRFormat(1)
    with pytest.raises(B2FError) as e:
        bsr.initialize(Mesh.from_numpy(dim, cells, coords), True, assembly_variant=fem.ASM_NODE_GATHER_SLOTS, pattern_layout=1)
    assert e.value.status == -2
