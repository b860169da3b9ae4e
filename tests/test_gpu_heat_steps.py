"""BASELINE config #4 as the reference's heat module runs it (modules/heat/FemModule.cc:113-131,386-537,775-840): every time
step re-assembles conduction + mass / dt, adds the convection boundary term, rebuilds the right-hand side from the last
temperature, the flux and the convection boundary, applies Dirichlet, and solves.  The reference does all of it on the
CPU through matrixAddValue; here every term is a device kernel behind the C ABI and is compared with the oracle per step
(pattern bit-exact, values and rhs 1e-12, Dirichlet masks bit-exact).  The solve (Hypre/PETSc in the reference) is a host
sparse direct solve of the DEVICE-assembled system, so a wrong device matrix would also drift the temperatures."""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from tests.util import MESHES, assert_values_close, assert_vec_close

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.mark.parametrize("dirichlet", ["penalty", "row"])
@pytest.mark.parametrize("mesh_name", ["square20", "square9_shuffled", "cube5", "cube7_shuffled"])
def test_transient_heat_steps_match_oracle(mesh_name, dirichlet):
    import arcanefem_b200 as fem
    from arcanefem_b200 import mesh as gen
    from arcanefem_b200.bsr import Mesh
    from oracle import oracle as orc
    dim, cells, coords = MESHES[mesh_name]()
    nn = coords.shape[0]
    lam, dt, h_conv, t_ext, flux, t_fixed, t_init = 1.75, 0.1, 12.0, 300.0, 40.0, 350.0, 290.0
    xmax, ymax = coords[:, 0].max(), coords[:, 1].max()
    conv_faces = gen.boundary_faces(cells, coords, 0, xmax)  # convection on x = L
    flux_faces = gen.boundary_faces(cells, coords, 1, ymax)  # constant flux on y = L
    fixed = gen.boundary_nodes(coords, 0, 0.0)  # Dirichlet on x = 0
    skip = np.zeros(nn, np.uint8)
    skip[fixed] = 1

    bsr = fem.BSRFormat(1)
    bsr.initialize(Mesh.from_numpy(dim, cells, coords), True)
    bsr.computeSparsity()
    ri, co = orc.build_pattern(cells, nn)
    np.testing.assert_array_equal(bsr.matrix().rowIndexWithEnd().cpu().numpy(), ri)
    np.testing.assert_array_equal(bsr.matrix().columns().cpu().numpy(), co)
    ls = fem.DoFLinearSystem()
    bsr.toLinearSystem(ls)
    d_conv, d_flux = torch.as_tensor(conv_faces).cuda(), torch.as_tensor(flux_faces).cuda()
    d_fixed, d_skip = torch.as_tensor(fixed).cuda(), torch.as_tensor(skip).cuda()
    rows = np.repeat(np.arange(nn), np.diff(ri))

    T = np.full(nn, t_init)
    T[fixed] = t_fixed
    for step in range(3):
        # ---- oracle: the module's sequence
        vals, nf = orc.assemble(cells, coords, ri, co, bs=1, per_block=False, op=orc.OP_HEAT, lambda_=lam, inv_dt=1.0 / dt)
        assert nf == 0
        assert orc.add_face_mass(conv_faces, coords, h_conv, ri, co, vals) == 0
        rhs = np.zeros(nn)
        orc.rhs_nodal_source(cells, coords, 1, 1.0 / dt, T, rhs, node_skip=skip)
        flux_rhs = np.zeros(nn)
        orc.rhs_neumann(flux_faces, coords, 1, flux_rhs, value=flux)
        orc.rhs_neumann(conv_faces, coords, 1, flux_rhs, value=h_conv * t_ext)
        finfo, fval = np.zeros(nn, np.uint8), np.zeros(nn)
        einfo, evalue = np.zeros(nn, np.uint8), np.zeros(nn)
        rhs += flux_rhs
        if dirichlet == "penalty":
            orc.dirichlet_flag_penalty(1, 0, t_fixed, 1.0e30, fixed, finfo, fval, rhs)
        else:
            orc.dirichlet_flag_elimination(1, 0, t_fixed, 1, fixed, einfo, evalue)
        orc.dirichlet_apply(ri, co, vals, einfo, evalue, finfo, fval, rhs)
        # ---- device, through the C ABI
        ls.clearValues()
        bsr.assembleBilinear(fem.Operator.heat(lam, dt))
        bsr.addFaceMass(d_conv, h_conv)
        bsr.applyNodalSourceToRhs(torch.as_tensor(T).cuda(), 1.0 / dt, node_skip=d_skip)
        bsr.applyNeumannToRhs(d_flux, value=flux)
        bsr.applyNeumannToRhs(d_conv, value=h_conv * t_ext)
        if dirichlet == "penalty":
            bsr.applyDirichletViaPenalty(d_fixed, t_fixed, 1.0e30)
        else:
            bsr.applyDirichletViaRowElimination(d_fixed, t_fixed)
        ls.applyDirichlet()
        torch.cuda.synchronize()
        assert bsr.stats()["columns_not_found"] == 0
        got_vals, got_rhs = bsr.matrix().values().cpu().numpy(), ls.rhsVariable().cpu().numpy()
        np.testing.assert_array_equal(ls.getForcedInfo().cpu().numpy(), finfo)
        np.testing.assert_array_equal(ls.getEliminationInfo().cpu().numpy(), einfo)
        assert_values_close(got_vals, vals, ri, 1, what=f"heat step {step} values")
        free = skip == 0  # the penalty entries (1e30 * T) would swamp a common scale: compared exactly, the others on their own
        assert_vec_close(got_rhs[free], rhs[free], what=f"heat step {step} rhs")
        np.testing.assert_array_equal(got_rhs[~free], rhs[~free])
        # ---- next temperature from the device-assembled system
        T = spla.spsolve(sp.csr_matrix((got_vals, (rows, co)), shape=(nn, nn)).tocsc(), got_rhs)
        assert np.isfinite(T).all()
        np.testing.assert_allclose(T[fixed], t_fixed, rtol=1e-12)
    assert T.min() > 250.0 and T.max() < 400.0  # a physical temperature field after three steps


def test_reference_heat_case_on_device():
    """modules/heat/inputs/conduction.convection.arc on the reference's own mesh (tests/golden/heat_plate_convection.npz): plate at 30,
    left edge held at 10 by the module's "Penalty" (a_ii += P), convection h = 1, Text = 20 on the other three edges, dt = 0.4.
    Every step's matrix and right-hand side from the device against the oracle (tests/test_oracle_golden.py pins that one on
    an independent scheme); the temperatures advance with the device-assembled system."""
    import arcanefem_b200 as fem
    from arcanefem_b200.bsr import Mesh
    from oracle import oracle as orc
    from tests.test_oracle_golden import HEAT, heat_initial_temperature, heat_oracle_step, load_case
    case = load_case([p for p in HEAT if p.endswith("heat_plate_convection.npz")][0])
    cells, coords = case["cells"], case["coords"]
    nn = coords.shape[0]
    lam, dt, P = float(case["lam"]), float(case["dt"]), float(case["penalty"])
    bsr = fem.BSRFormat(1)
    bsr.initialize(Mesh.from_numpy(2, cells, coords), True)
    bsr.computeSparsity()
    ri, co = orc.build_pattern(cells, nn)
    np.testing.assert_array_equal(bsr.matrix().rowIndexWithEnd().cpu().numpy(), ri)
    np.testing.assert_array_equal(bsr.matrix().columns().cpu().numpy(), co)
    ls = fem.DoFLinearSystem()
    bsr.toLinearSystem(ls)
    conv = [(torch.as_tensor(case[f"convection{i}_faces"]).cuda(), float(case[f"convection{i}_h"]), float(case[f"convection{i}_text"]))
            for i in range(int(case["nb_convection"]))]
    dirichlet = [(torch.as_tensor(case[f"dirichlet{i}_nodes"]).cuda(), float(case[f"dirichlet{i}_value"])) for i in range(int(case["nb_dirichlet"]))]
    rows = np.repeat(np.arange(nn), np.diff(ri))
    T = heat_initial_temperature(case)
    d_skip = None
    for step in range(6):
        vals, rhs, fixed, finfo = heat_oracle_step(case, ri, co, T)
        if d_skip is None:
            d_skip = torch.as_tensor(fixed).cuda()
        ls.clearValues()
        bsr.assembleBilinear(fem.Operator.heat(lam, dt))
        bsr.applyNodalSourceToRhs(torch.as_tensor(T).cuda(), 1.0 / dt, node_skip=d_skip)
        for faces, h, text in conv:
            bsr.addFaceMass(faces, h)
            bsr.applyNeumannToRhs(faces, value=h * text)
        for nodes, g in dirichlet:
            bsr.applyDirichletViaWeakPenalty(nodes, g, P)
        ls.applyDirichlet()
        torch.cuda.synchronize()
        assert bsr.stats()["columns_not_found"] == 0
        got_vals, got_rhs = bsr.matrix().values().cpu().numpy(), ls.rhsVariable().cpu().numpy()
        np.testing.assert_array_equal(ls.getForcedInfo().cpu().numpy(), finfo)
        assert_values_close(got_vals, vals, ri, 1, what=f"plate step {step} values")
        free = fixed == 0
        assert_vec_close(got_rhs[free], rhs[free], what=f"plate step {step} rhs")
        np.testing.assert_array_equal(got_rhs[~free], rhs[~free])
        T = spla.spsolve(sp.csr_matrix((got_vals, (rows, co)), shape=(nn, nn)).tocsc(), got_rhs)
    assert np.allclose(T[fixed == 1], 10.0, rtol=1e-12) and T.max() < 30.0
