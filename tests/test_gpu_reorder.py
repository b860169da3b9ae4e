"""b2f_options.reorder (MORTON / PENCIL): the library renumbers the connectivity internally along a space-filling order of the
nodes, builds its plan there and writes every row where the CALLER's row_index and column order put it.  Everything a caller
sees must be the same as without reordering: pattern bit-exact in the caller's numbering, values / fused right-hand side /
accumulate mode within rounding of the unreordered kernel and within 1e-12 of the oracle, tiles on the global fallback path
included.  The reference has no locality precondition (one thread per cell, femutils/BSRFormat.h:807-836)."""
import numpy as np
import pytest

from tests.util import MESHES, assert_values_close, assert_vec_close, op_params

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _np(t):
    return t.detach().cpu().numpy()


def _build(dim, cells, coords, bs, use_csr, reorder, pattern_layout=0, node_is_own=None):
    import arcanefem_b200 as fem
    from arcanefem_b200.bsr import Mesh
    bsr = fem.BSRFormat(bs)
    bsr.initialize(Mesh.from_numpy(dim, cells, coords, node_is_own), use_csr, reorder=reorder, pattern_layout=pattern_layout)
    bsr.computeSparsity()
    return fem, bsr


CASES = [("cube7_shuffled", 1, "laplace"), ("cube12", 1, "laplace"), ("square9_shuffled", 1, "laplace"), ("square20", 2, "elasticity"), ("cube7_shuffled", 3, "elasticity"),
         ("cube4_mixed_orientation", 1, "heat_cell"), ("square20", 1, "heat_cell"), ("cube5", 1, "heat")]


@pytest.mark.parametrize("mode", [1, 2])
@pytest.mark.parametrize("pattern_layout", [0, 1])
@pytest.mark.parametrize("use_csr", [True, False])
@pytest.mark.parametrize("mesh_name,bs,kind", CASES)
def test_reordered_assembly_matches_the_plain_kernel_and_the_oracle(mesh_name, bs, kind, use_csr, pattern_layout, mode):
    from oracle import oracle as orc
    dim, cells, coords = MESHES[mesh_name]()
    okw = op_params(kind, cells.shape[0])
    own = (np.random.default_rng(9).uniform(size=coords.shape[0]) < 0.8).astype(np.uint8)
    out = []
    for reorder in (0, mode):
        fem, bsr = _build(dim, cells, coords, bs, use_csr, reorder, pattern_layout, node_is_own=own)
        assert bsr.stats()["reordered"] == (1 if reorder else 0)
        cl = torch.as_tensor(okw["cell_lambda"]).cuda() if okw.get("cell_lambda") is not None else None
        op = {"laplace": fem.Operator.laplace, "elasticity": lambda: fem.Operator.elasticity(21.0e5, 0.28)}.get(kind, lambda: fem.Operator.heat(okw.get("lambda_", 0.0), 0.1, cell_lambda=cl))()
        ls = fem.DoFLinearSystem()
        bsr.toLinearSystem(ls)
        ls.clearValues()
        bsr.rhsFill(0.0)
        bsr.assembleBilinearWithSource(op, 5.5)
        torch.cuda.synchronize()
        st = bsr.stats()
        assert st["columns_not_found"] == 0 and st["gather_fallback_tiles"] == 0
        mat = bsr.matrix()
        first = mat.values().clone()
        bsr.assembleBilinear(op, accumulate=True)
        torch.cuda.synchronize()
        out.append((_np(mat.rowIndexWithEnd()), _np(mat.columns()), first, mat.values().clone(), _np(ls.rhsVariable()).copy()))
    (ri0, co0, v0, a0, r0), (ri1, co1, v1, a1, r1) = out
    np.testing.assert_array_equal(ri1, ri0)  # the views stay in the caller's numbering
    np.testing.assert_array_equal(co1, co0)
    # the rows are walked in another order: equal to rounding (row-norm relative, the rule of tests/util.py)
    assert_values_close(_np(v1), _np(v0), ri0, bs, what="reordered vs plain kernel")
    assert_values_close(_np(a1), _np(a0), ri0, bs, what="accumulate mode")
    assert_vec_close(r1, r0, what="fused rhs")
    if pattern_layout == 0 and use_csr:
        ri, co = orc.build_pattern(cells, coords.shape[0])
        ref, _ = orc.assemble(cells, coords, ri, co, bs=bs, per_block=False, node_is_own=own, **okw)
        assert_values_close(_np(v1), ref, ri, bs, what=f"{mesh_name} {kind} reordered")


def test_coordinates_are_re_permuted_at_every_assembly():
    """The caller moves its nodes between two assemblies (same array, new values): the internal copy must follow."""
    from oracle import oracle as orc
    dim, cells, coords = MESHES["cube7_shuffled"]()
    fem, bsr = _build(dim, cells, coords, 1, True, 1)
    op = fem.Operator.laplace()
    bsr.assembleBilinear(op)
    moved = coords * np.array([1.0, 1.3, 0.7])
    bsr._mesh.coords.copy_(torch.as_tensor(moved))
    bsr.assembleBilinear(op)
    torch.cuda.synchronize()
    ri, co = orc.build_pattern(cells, coords.shape[0])
    ref, _ = orc.assemble(cells, moved, ri, co)
    assert_values_close(_np(bsr.matrix().values()), ref, ri, 1, what="moved nodes")


def test_rows_too_long_for_the_walk_plan_switch_the_reordering_off():
    """A hub node with 3000 triangles (more than 1023 blocks in a row): no walk plan, the search-based kernel runs on the caller's
    arrays and the option is reported as not applied."""
    from oracle import oracle as orc
    K = 3000
    ang = np.linspace(0.0, 2 * np.pi, K, endpoint=False)
    coords = np.zeros((K + 1, 3))
    coords[1:, 0] = np.cos(ang)
    coords[1:, 1] = np.sin(ang)
    cells = np.stack([np.zeros(K, np.int32), 1 + np.arange(K, dtype=np.int32), 1 + (np.arange(K, dtype=np.int32) + 1) % K], axis=1).astype(np.int32)
    ri, co = orc.build_pattern(cells, K + 1)
    ref, _ = orc.assemble(cells, coords, ri, co)
    fem, bsr = _build(2, cells, coords, 1, True, 1)
    assert bsr.stats()["reordered"] == 0
    bsr.assembleBilinear(fem.Operator.laplace())
    torch.cuda.synchronize()
    np.testing.assert_array_equal(_np(bsr.matrix().columns()), co)
    assert_values_close(_np(bsr.matrix().values()), ref, ri, 1, rtol=1e-11, what="hub")


@pytest.mark.parametrize("bs,kind,mesh_name", [(1, "laplace", "cube7_shuffled"), (3, "elasticity", "cube7_shuffled"), (2, "elasticity", "square20")])
def test_tiles_on_the_fallback_path_work_in_the_caller_numbering(bs, kind, mesh_name, monkeypatch):
    """B2F_FAN_TILE_CAP_BLOCKS (test hook of the launcher) makes the shared-memory tile so small that ordinary tiles overflow and
    take the search-based global path; in the reorder mode that path maps the tile's internal rows back to the caller's."""
    from oracle import oracle as orc
    dim, cells, coords = MESHES[mesh_name]()
    okw = op_params(kind, cells.shape[0])
    ri, co = orc.build_pattern(cells, coords.shape[0])
    ref, _ = orc.assemble(cells, coords, ri, co, bs=bs, per_block=False, **okw)
    rhs_ref = np.zeros(coords.shape[0] * bs)
    orc.rhs_constant_source(cells, coords, bs, 5.5, rhs_ref)
    monkeypatch.setenv("B2F_FAN_TILE_CAP_BLOCKS", "64")
    for reorder in (0, 1, 2):
        fem, bsr = _build(dim, cells, coords, bs, True, reorder)
        op = fem.Operator.laplace() if kind == "laplace" else fem.Operator.elasticity(21.0e5, 0.28)
        ls = fem.DoFLinearSystem()
        bsr.toLinearSystem(ls)
        ls.clearValues()
        bsr.rhsFill(0.0)
        bsr.assembleBilinearWithSource(op, 5.5)
        torch.cuda.synchronize()
        assert bsr.stats()["gather_fallback_tiles"] >= 1
        assert_values_close(_np(bsr.matrix().values()), ref, ri, bs, what=f"fallback tiles, reorder {reorder}")
        assert_vec_close(_np(ls.rhsVariable()), rhs_ref, what=f"fused rhs on fallback tiles, reorder {reorder}")


def test_host_entry_and_other_variants_ignore_or_survive_the_option():
    """b2f_assemble_host falls back to its one-shot path (row chunks follow the caller's order); the cell-wise variant ignores
    the internal order."""
    from oracle import oracle as orc
    import arcanefem_b200 as fem
    from arcanefem_b200.bsr import Mesh
    dim, cells, coords = MESHES["cube12"]()
    ri, co = orc.build_pattern(cells, coords.shape[0])
    ref, _ = orc.assemble(cells, coords, ri, co)
    fem_, bsr = _build(dim, cells, coords, 1, True, 1)
    ch = torch.as_tensor(coords).pin_memory()
    vh = torch.empty(co.shape[0], dtype=torch.float64).pin_memory()
    bsr.assembleBilinearHost(fem.Operator.laplace(), ch, vh)
    assert_values_close(vh.numpy(), ref, ri, 1, what="host entry, reordered")
    b2 = fem.BSRFormat(1)
    b2.initialize(Mesh.from_numpy(dim, cells, coords), True, assembly_variant=fem.ASM_CELL_ATOMIC, reorder=1)
    b2.computeSparsity()
    assert b2.stats()["reordered"] == 0
    b2.assembleBilinear(fem.Operator.laplace())
    torch.cuda.synchronize()
    assert_values_close(_np(b2.matrix().values()), ref, ri, 1, what="cell atomic with reorder requested")
