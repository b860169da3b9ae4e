"""Pins the oracle on the reference's own golden data: every fixture under tests/golden/ is one reference ctest case
(mesh + parameters of the cited .arc + the reference's expected nodal solution, see tests/golden/make_golden.py).

The reference holds no matrix-level vectors for this path; its tests assemble, apply the boundary conditions, solve with
Hypre/PETSc and compare the nodal field with a relative epsilon of 1e-4 (femutils/FemUtils.cc:84-170).  Here the oracle
assembles matrix and RHS, a sparse direct solve (scipy) stands in for the third-party solver, and the same comparison
rule is applied.  A wrong element matrix, pattern, source/Neumann term or Dirichlet rule fails these tests.
"""
import glob
import os

import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spl

from oracle import oracle as orc

_ALL = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "*.npz")))
HEAT = [p for p in _ALL if os.path.basename(p).startswith("heat_")]  # transient cases: no result file in the reference, replayed below
GOLDEN = [p for p in _ALL if p not in HEAT]


def load_case(path):
    z = np.load(path)
    case = {k: z[k] for k in z.files}
    case["name"] = os.path.splitext(os.path.basename(path))[0]
    return case


def is_elasticity(case):
    return "kind" in case and str(case["kind"]) == "elasticity"


def elasticity_rhs(case):
    """Body force and traction of modules/elasticity/FemModule.cc:220-272 (host loops in the reference as well; "NULL"
    components, stored as NaN, carry no load): rhs[dof(n,i)] += f_i * area / 3 per cell node and t_i * length / 2 per
    face node."""
    cells, coords = case["cells"], case["coords"]
    rhs = np.zeros(2 * coords.shape[0])
    f = np.nan_to_num(case["f2"], nan=0.0)
    if not np.isnan(case["f2"]).all():
        p = coords[cells]
        cr = np.cross(p[:, 1] - p[:, 0], p[:, 2] - p[:, 0])
        area = 0.5 * np.sqrt((cr * cr).sum(axis=1)) if cr.ndim == 2 else 0.5 * np.abs(cr)
        for i in range(2):
            np.add.at(rhs, 2 * cells.ravel() + i, np.repeat(f[i] * area / 3.0, 3))
    for k in range(int(case["nb_traction"])):
        faces, t = case[f"traction{k}_faces"], np.nan_to_num(case[f"traction{k}_t"], nan=0.0)
        d = coords[faces[:, 1]] - coords[faces[:, 0]]
        length = np.sqrt((d * d).sum(axis=1))
        for i in range(2):
            np.add.at(rhs, 2 * faces.ravel() + i, np.repeat(t[i] * length / 2.0, 2))
    return rhs


def elasticity_params(case):
    E, nu = float(case["E"]), float(case["nu"])
    return dict(op=orc.OP_ELASTICITY, lambda_=E * nu / ((1 + nu) * (1 - 2 * nu)), mu2=(E / (2 * (1 + nu))) * 2)  # FemModule.cc:133-134


def check_against_reference(case, u):
    """femutils/FemUtils.cc:84-90,146-163: |ref - v| / (|ref| + |v|) compared with epsilon (Arcane isNearlyEqualWithEpsilon),
    values below min_value on both sides are skipped."""
    uid_to_lid = {int(t): i for i, t in enumerate(case["node_uid"])}
    eps, min_value = float(case["eps"]), 1.0e-16
    bad = []
    if is_elasticity(case):  # Real3 rule: every component with the scalar rule (femutils/FemUtils.cc:92-99)
        pairs = [(uid, ref[i], u[2 * uid_to_lid[int(uid)] + i]) for uid, ref in zip(case["expected_uid"], case["expected_vec"]) for i in range(2)]
    else:
        pairs = [(uid, ref, u[uid_to_lid[int(uid)]]) for uid, ref in zip(case["expected_uid"], case["expected_value"])]
    for uid, ref, v in pairs:
        if abs(ref) < min_value and abs(v) < min_value:
            continue
        if abs(ref - v) / (abs(ref) + abs(v)) > eps:
            bad.append((int(uid), float(ref), float(v)))
    assert not bad, f"{case['name']}: {len(bad)} nodes differ from the reference result file, e.g. {bad[:3]}"


def oracle_system(case):
    """Matrix, RHS after the reference's call sequence (modules/poisson/FemModule.cc:100-191): bilinear operator, source,
    Neumann, Dirichlet by penalty, then the solve() prelude (_applyForcedValuesToLhs)."""
    cells, coords = case["cells"], case["coords"]
    nn = coords.shape[0]
    ri, co = orc.build_pattern(cells, nn)
    if is_elasticity(case):
        # modules/elasticity/FemModule.cc:102-116: bilinear operator (BS = 2), toLinearSystem, body force + traction,
        # per-component penalty Dirichlet (ArcaneFemFunctionsGpu.h:514-523), then the solve() prelude
        vals, nf = orc.assemble(cells, coords, ri, co, bs=2, per_block=False, **elasticity_params(case))
        assert nf == 0
        rows, nbcol, cols = orc.to_csr(2, ri, co)
        rhs = elasticity_rhs(case)
        finfo, fval = np.zeros(2 * nn, np.uint8), np.zeros(2 * nn)
        einfo, ev = np.zeros(2 * nn, np.uint8), np.zeros(2 * nn)
        for i in range(int(case["nb_dirichlet"])):
            for comp, value in enumerate(case[f"dirichlet{i}_value2"]):
                if not np.isnan(value):
                    orc.dirichlet_flag_penalty(2, comp, float(value), float(case["penalty"]), case[f"dirichlet{i}_nodes"], finfo, fval, rhs)
        orc.dirichlet_apply(rows, cols, vals, einfo, ev, finfo, fval, rhs)
        return ri, co, vals, rhs, (rows, cols)
    vals, nf = orc.assemble(cells, coords, ri, co, per_block=False)
    assert nf == 0
    rows, nbcol, cols = orc.to_csr(1, ri, co)
    rhs = np.zeros(nn)
    orc.rhs_constant_source(cells, coords, 1, float(case["f"]), rhs)
    for i in range(int(case["nb_neumann"])):
        orc.rhs_neumann(case[f"neumann{i}_faces"], coords, 1, rhs, value_xyz=tuple(case[f"neumann{i}_flux"]))
    finfo, fval = np.zeros(nn, np.uint8), np.zeros(nn)
    einfo, ev = np.zeros(nn, np.uint8), np.zeros(nn)
    for i in range(int(case["nb_dirichlet"])):
        orc.dirichlet_flag_penalty(1, 0, float(case[f"dirichlet{i}_value"]), float(case["penalty"]), case[f"dirichlet{i}_nodes"], finfo, fval, rhs)
    orc.dirichlet_apply(rows, cols, vals, einfo, ev, finfo, fval, rhs)
    return ri, co, vals, rhs, (rows, cols)


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_oracle_reproduces_reference_solution(path):
    case = load_case(path)
    ri, co, vals, rhs, (rows, cols) = oracle_system(case)
    A = sp.csr_matrix((vals, cols, rows))
    u = spl.spsolve(A.tocsc(), rhs)
    check_against_reference(case, u)


def test_golden_fixtures_present():
    assert len(GOLDEN) >= 12


# ---------------------------------------------------------------------------------------------------
# Heat (modules/heat): the reference runs its cases without comparing anything (modules/heat/CMakeLists.txt:43-58), so there
# is no result file to pin on.  The case of inputs/conduction.convection.arc is replayed on the reference's own mesh with
# the oracle (the module's time step: FemModule.cc:113-131,386-537,775-840) beside an independent scipy formulation of the
# same scheme (dense element matrices from the P1 basis, test_oracle.py); both must give the same temperatures.
# ---------------------------------------------------------------------------------------------------
def heat_oracle_step(case, ri, co, T):
    """Matrix and right-hand side of one implicit Euler step, the module's sequence: conduction + mass / dt, convection
    faces, (T_old / dt) source on the free nodes, h * Text on the convection faces, then "Penalty" = a_ii += P, b_i = P * T."""
    cells, coords = case["cells"], case["coords"]
    nn = coords.shape[0]
    vals, nf = orc.assemble(cells, coords, ri, co, bs=1, per_block=False, op=orc.OP_HEAT, lambda_=float(case["lam"]), inv_dt=1.0 / float(case["dt"]))
    assert nf == 0
    rhs = np.zeros(nn)
    fixed = np.zeros(nn, np.uint8)
    for i in range(int(case["nb_dirichlet"])):
        fixed[case[f"dirichlet{i}_nodes"]] = 1
    orc.rhs_nodal_source(cells, coords, 1, 1.0 / float(case["dt"]), T, rhs, node_skip=fixed)
    for i in range(int(case["nb_convection"])):
        faces, h, text = case[f"convection{i}_faces"], float(case[f"convection{i}_h"]), float(case[f"convection{i}_text"])
        assert orc.add_face_mass(faces, coords, h, ri, co, vals) == 0
        orc.rhs_neumann(faces, coords, 1, rhs, value=h * text)
    finfo, fval = np.zeros(nn, np.uint8), np.zeros(nn)
    einfo, ev = np.zeros(nn, np.uint8), np.zeros(nn)
    for i in range(int(case["nb_dirichlet"])):
        orc.dirichlet_flag_penalty(1, 0, float(case[f"dirichlet{i}_value"]), float(case["penalty"]), case[f"dirichlet{i}_nodes"], finfo, fval, rhs, weak=True)
    orc.dirichlet_apply(ri, co, vals, einfo, ev, finfo, fval, rhs)
    return vals, rhs, fixed, finfo


def heat_initial_temperature(case):
    T = np.full(case["coords"].shape[0], float(case["tinit"]))
    for i in range(int(case["nb_dirichlet"])):
        T[case[f"dirichlet{i}_nodes"]] = float(case[f"dirichlet{i}_value"])  # _initTemperature + _applyDirichletBoundaryConditions
    return T


@pytest.mark.parametrize("path", HEAT, ids=[os.path.basename(p)[:-4] for p in HEAT])
def test_oracle_heat_case_against_independent_scheme(path):
    from tests.test_oracle import p1_gradients
    case = load_case(path)
    cells, coords = case["cells"], case["coords"]
    nn = coords.shape[0]
    lam, dt, P = float(case["lam"]), float(case["dt"]), float(case["penalty"])
    # independent: K (conduction), M (consistent mass), C (convection face mass), c (h * Text load) as dense-element scipy sums
    K, M, Cm = sp.lil_matrix((nn, nn)), sp.lil_matrix((nn, nn)), sp.lil_matrix((nn, nn))
    for cn in cells:
        G, area = p1_gradients(coords[cn][:, :2])
        K[np.ix_(cn, cn)] += lam * area * (G @ G.T)
        M[np.ix_(cn, cn)] += area / 12.0 * (np.ones((3, 3)) + np.eye(3))
    load = np.zeros(nn)
    for i in range(int(case["nb_convection"])):
        h, text = float(case[f"convection{i}_h"]), float(case[f"convection{i}_text"])
        for f in case[f"convection{i}_faces"]:
            L = np.hypot(*(coords[f[1]] - coords[f[0]])[:2])
            Cm[np.ix_(f, f)] += h * L / 6.0 * (np.ones((2, 2)) + np.eye(2))
            load[f] += h * text * L / 2.0
    lumped = np.asarray(M.sum(axis=1)).ravel()  # the module's source uses area / 3 per cell node = the lumped mass
    fixed = np.zeros(nn, bool)
    g = np.zeros(nn)
    for i in range(int(case["nb_dirichlet"])):
        fixed[case[f"dirichlet{i}_nodes"]] = True
        g[case[f"dirichlet{i}_nodes"]] = float(case[f"dirichlet{i}_value"])
    A = (K + M / dt + Cm).tocsr() + sp.diags(np.where(fixed, P, 0.0))
    ri, co = orc.build_pattern(cells, nn)
    rows = np.repeat(np.arange(nn), np.diff(ri))
    T = heat_initial_temperature(case)
    T2 = T.copy()
    nsteps = int(round(float(case["tmax"]) / dt))
    assert nsteps == 50
    for step in range(nsteps):
        vals, rhs, _, _ = heat_oracle_step(case, ri, co, T)
        T = spl.spsolve(sp.csr_matrix((vals, (rows, co)), shape=(nn, nn)).tocsc(), rhs)
        b = np.where(fixed, P * g, lumped * T2 / dt + load)
        T2 = spl.spsolve(A.tocsc(), b)
        np.testing.assert_allclose(T, T2, rtol=1e-10)
    # physics of the case: plate at 30 with the left edge held at 10 and air at 20 around it
    assert np.allclose(T[fixed], 10.0, rtol=1e-12)
    assert 10.0 - 1e-9 <= T.min() and T.max() < 30.0
    assert T[coords[:, 0] > 0.9 * coords[:, 0].max()].mean() < 25.0  # the far side has been cooling towards the air temperature
