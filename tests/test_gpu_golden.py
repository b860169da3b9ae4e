"""The CUDA path (through the C ABI) on the reference's golden cases: device assembly + RHS + Dirichlet, matrix/RHS
compared with the oracle entry by entry (1e-12) and the solved field with the reference's result file (1e-4)."""
import os

import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spl

from tests.test_oracle_golden import GOLDEN, check_against_reference, elasticity_rhs, is_elasticity, load_case, oracle_system
from tests.util import assert_values_close, assert_vec_close

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


@pytest.mark.parametrize("variant", [1, 3])
@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_cuda_path_reproduces_reference_case(path, variant):
    import arcanefem_b200 as fem
    from arcanefem_b200.bsr import Mesh
    case = load_case(path)
    cells, coords = case["cells"], case["coords"]
    if is_elasticity(case):
        return _elasticity_case(case, variant)
    bsr = fem.BSRFormat(1)
    bsr.initialize(Mesh.from_numpy(int(case["dim"]), cells, coords), True, assembly_variant=variant)
    bsr.computeSparsity()
    bsr.assembleBilinear(fem.Operator.laplace())
    ls = fem.DoFLinearSystem()
    bsr.toLinearSystem(ls)
    ls.clearValues()
    bsr.rhsFill(0.0)
    bsr.applyConstantSourceToRhs(float(case["f"]))
    keep = []
    for i in range(int(case["nb_neumann"])):
        faces = torch.as_tensor(case[f"neumann{i}_faces"]).cuda()
        keep.append(faces)
        bsr.applyNeumannToRhs(faces, value_xyz=tuple(case[f"neumann{i}_flux"]))
    for i in range(int(case["nb_dirichlet"])):
        ids = torch.as_tensor(case[f"dirichlet{i}_nodes"]).cuda()
        keep.append(ids)
        bsr.applyDirichletViaPenalty(ids, float(case[f"dirichlet{i}_value"]), float(case["penalty"]))
    ls.applyDirichlet()
    torch.cuda.synchronize()
    csr = ls.getCSRValues()
    rows, cols, vals = csr.rowsWithEnd().cpu().numpy(), csr.columns().cpu().numpy(), csr.values().cpu().numpy()
    rhs = ls.rhsVariable().cpu().numpy()
    ri, co, ref_vals, ref_rhs, _ = oracle_system(case)
    np.testing.assert_array_equal(rows, ri)
    np.testing.assert_array_equal(cols, co)
    assert_values_close(vals, ref_vals, ri, 1, what=case["name"] + " matrix")
    assert_vec_close(rhs, ref_rhs, what=case["name"] + " rhs")
    u = spl.spsolve(sp.csr_matrix((vals, cols, rows)).tocsc(), rhs)
    check_against_reference(case, u)


def _elasticity_case(case, variant):
    """modules/elasticity cases (BS = 2): device matrix, per-component penalty Dirichlet and solve() prelude; the body
    force / traction vector comes from the host, as in the reference (modules/elasticity/FemModule.cc:220-272)."""
    import arcanefem_b200 as fem
    from arcanefem_b200.bsr import Mesh
    cells, coords = case["cells"], case["coords"]
    bsr = fem.BSRFormat(2)
    bsr.initialize(Mesh.from_numpy(int(case["dim"]), cells, coords), True, assembly_variant=variant)
    bsr.computeSparsity()
    bsr.assembleBilinear(fem.Operator.elasticity(float(case["E"]), float(case["nu"])))
    ls = fem.DoFLinearSystem()
    bsr.toLinearSystem(ls)
    ls.clearValues()
    ls.rhsVariable().copy_(torch.as_tensor(elasticity_rhs(case)))
    keep = []
    for i in range(int(case["nb_dirichlet"])):
        ids = torch.as_tensor(case[f"dirichlet{i}_nodes"]).cuda()
        keep.append(ids)
        for comp, value in enumerate(case[f"dirichlet{i}_value2"]):
            if not np.isnan(value):
                bsr.applyDirichletViaPenalty(ids, float(value), float(case["penalty"]), dof_index=comp)
    ls.applyDirichlet()
    torch.cuda.synchronize()
    csr = ls.getCSRValues()
    rows, cols, vals = csr.rowsWithEnd().cpu().numpy(), csr.columns().cpu().numpy(), csr.values().cpu().numpy()
    rhs = ls.rhsVariable().cpu().numpy()
    ri, co, ref_vals, ref_rhs, (ref_rows, ref_cols) = oracle_system(case)
    np.testing.assert_array_equal(rows, ref_rows)
    np.testing.assert_array_equal(cols, ref_cols)
    assert_values_close(vals, ref_vals, ri, 2, what=case["name"] + " matrix")
    assert_vec_close(rhs, ref_rhs, what=case["name"] + " rhs")
    u = spl.spsolve(sp.csr_matrix((vals, cols, rows)).tocsc(), rhs)
    check_against_reference(case, u)
