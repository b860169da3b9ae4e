#!/usr/bin/env python
"""Generates tests/golden/*.npz from the reference's own case data (run in the build container, where
/root/reference exists; the GPU box and the test-suite only read the committed .npz files).

Each fixture = one reference test case (an .arc case file + its mesh + its result file):
  mesh       meshes/msh/<name>.msh read by arcanefem_b200.msh (MSH 4.1 binary)
  parameters copied by hand from the cited .arc file
  expected   the reference's result file (`uid value` per line, femutils/FemUtils.cc:100-170), which the reference's own
             ctest compares the solved field against with a relative epsilon (1e-4 poisson/testlab)
No reference source code is copied: only test DATA (mesh connectivity/coordinates and expected nodal values).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from arcanefem_b200.msh import read_msh  # noqa: E402

REF = "/root/reference"

# name -> (mesh, result file, parameters); citations = the .arc the parameters come from
CASES = {
    # modules/poisson/inputs/circle.2D.bsr.hypre.arc
    "poisson_circle_2D": dict(mesh="circle_cut", result="modules/poisson/check/poisson_test_ref_circle_2D.txt", f=5.5,
                              dirichlet=[("horizontal", 0.5)], penalty=1.0e30, eps=1.0e-4),
    # modules/poisson/inputs/circle.neumann.2D.bsr.hypre.arc
    "poisson_circle_neumann_2D": dict(mesh="circle_cut", result="modules/poisson/check/poisson_test_ref_circle_neumann_2D.txt", f=5.5,
                                      dirichlet=[("horizontal", 0.5)], neumann=[("curved", (-0.35, 1.65, 0.0))], penalty=1.0e30, eps=1.0e-4),
    # modules/poisson/inputs/perforatedSquare.pointDirichlet.2D.bsr.hypre.arc
    "poisson_point_dirichlet_2D": dict(mesh="plancher", result="modules/poisson/check/poisson_test_point_dirichlet_2D.txt", f=0.0,
                                       dirichlet=[("topLeftCorner", 50.0), ("topRightCorner", 20.0), ("botLeftCorner", 20.0), ("botRightCorner", 50.0)],
                                       penalty=1.0e30, eps=1.0e-4),
    # modules/testlab/inputs/Test.sphere.3D.bsr.hypre.arc
    "testlab_sphere_3D": dict(mesh="sphere_cut", result="modules/testlab/tests/poisson_test_ref_sphere_3D.txt", f=5.5,
                              dirichlet=[("horizontal", 0.5)], penalty=1.0e31, eps=1.0e-4),
    # modules/poisson/inputs/sphere.neumann.3D.bsr.hypre.arc
    "poisson_sphere_neumann_3D": dict(mesh="sphere_cut", result="modules/poisson/check/poisson_test_ref_sphere_neumann_3D.txt", f=5.5,
                                      dirichlet=[("horizontal", 0.5)], neumann=[("curved", (0.35, 1.65, 3.75))], penalty=1.0e30, eps=1.0e-4),
    # modules/testlab/inputs/Test.L-shape.2D.bsr.hypre.arc
    "testlab_Lshape_2D": dict(mesh="L-shape", result="modules/testlab/tests/poisson_test_ref_L-shape_2D.txt", f=-5.5,
                              dirichlet=[("boundary", 0.5)], penalty=1.0e30, eps=1.0e-4),
    # modules/testlab/inputs/Test.L-shape.3D.arc
    "testlab_Lshape_3D": dict(mesh="L-shape-3D", result="modules/testlab/tests/poisson_test_ref_L-shape_3D.txt", f=5.5,
                              dirichlet=[("bot", 50.0), ("bc", 10.0)], penalty=1.0e30, eps=1.0e-4),
    # modules/laplace/inputs/L-shape.3D.bsr.arc (Laplace = no source; penalty default of femutils/FemBoundaryConditions.axl:67)
    "laplace_Lshape_3D": dict(mesh="L-shape-3D", result="modules/laplace/check/test_3D_L-shape.txt", f=0.0,
                              dirichlet=[("bot", 50.0), ("bc", 10.0)], penalty=1.0e30, eps=1.0e-4),
    # modules/laplace/inputs/PointDirichlet.bsr.hypreDirect.arc
    "laplace_point_dirichlet_2D": dict(mesh="plancher", result="modules/laplace/check/test3_results.txt", f=0.0,
                                       dirichlet=[("topLeftCorner", 50.0), ("topRightCorner", 20.0), ("botLeftCorner", 20.0), ("botRightCorner", 50.0)],
                                       penalty=1.0e30, eps=1.0e-4),
}


NULL = float("nan")  # a component the .arc leaves "NULL" (not imposed / no load)

# Elasticity cases (modules/elasticity): vector unknown (ux, uy), BS = 2.  Result files hold `uid ux uy uz`; the module
# validates with epsilon = 1e-3 and min value 1e-16 (modules/elasticity/FemModule.cc:547-553).  Penalty default 1e30
# (modules/elasticity/Fem.axl:37).
ELASTICITY_CASES = {
    # modules/elasticity/inputs/bar.2D.traction.bsr.hypre.arc
    "elasticity_traction_bar": dict(mesh="bar", result="modules/elasticity/check/elasticity_traction_bar_test_ref.txt", E=21.0e5, nu=0.28, f=(NULL, NULL),
                                    dirichlet=[("left", (0.0, 0.0))], traction=[("right", (1.0, NULL))], penalty=1.0e30, eps=1.0e-3),
    # modules/elasticity/inputs/bar.2D.traction.bodyforce.bsr.arc
    "elasticity_traction_bodyforce_bar": dict(mesh="bar", result="modules/elasticity/check/elasticity_traction_bodyforce_bar_test_ref.txt", E=21.0e5, nu=0.28,
                                              f=(3.33, -6.66), dirichlet=[("left", (0.0, 0.0))], traction=[("right", (1.33, 2.13))], penalty=1.0e30, eps=1.0e-3),
    # modules/elasticity/inputs/bar.2D.PointDirichlet.bsr.hypre.arc
    "elasticity_point_dirichlet_bar": dict(mesh="bar", result="modules/elasticity/check/elasticity_point-dirichlet_bar_test_ref.txt", E=21.0e5, nu=0.28,
                                           f=(NULL, -1.0), dirichlet=[("left", (0.0, NULL)), ("right", (1.0, NULL)), ("botLeft", (0.0, 0.0)), ("botRight", (NULL, 0.0))],
                                           traction=[], penalty=1.0e30, eps=1.0e-3),
}


def orient_outward(faces, cells, coords):
    """Order the nodes of every boundary face so that the reference's normal formula
    (2D: (y1-y0, x0-x1); 3D: (p1-p0)x(p2-p0), femutils/ArcaneFemFunctionsGpu.h:137-197) points out of the adjacent cell —
    the state Arcane reaches with isSubDomainBoundaryOutside() == true."""
    faces = faces.copy()
    key = {}
    for c, vs in enumerate(cells):
        for i in range(len(vs)):
            sub = tuple(sorted(int(v) for j, v in enumerate(vs) if j != i))
            key[sub] = (c, int(vs[i]))
    for f in range(faces.shape[0]):
        c, opp = key[tuple(sorted(int(v) for v in faces[f]))]
        p = coords[faces[f]]
        if faces.shape[1] == 2:
            n = np.array([p[1, 1] - p[0, 1], p[0, 0] - p[1, 0], 0.0])
        else:
            n = np.cross(p[1] - p[0], p[2] - p[0])
        if np.dot(n, coords[opp] - p[0]) > 0:  # points towards the cell: flip
            faces[f, 0], faces[f, 1] = faces[f, 1], faces[f, 0]
    return faces


def main():
    for name, case in CASES.items():
        m = read_msh(os.path.join(REF, "meshes", "msh", case["mesh"] + ".msh"))
        cells, coords = m.cells(), m.coords
        exp = np.loadtxt(os.path.join(REF, case["result"]))
        out = dict(dim=np.int32(m.dim), cells=cells, coords=coords, node_uid=m.node_tags, expected_uid=exp[:, 0].astype(np.int64), expected_value=exp[:, 1],
                   f=np.float64(case["f"]), penalty=np.float64(case["penalty"]), eps=np.float64(case["eps"]))
        for i, (group, value) in enumerate(case["dirichlet"]):
            out[f"dirichlet{i}_nodes"] = m.group_nodes(group)
            out[f"dirichlet{i}_value"] = np.float64(value)
        out["nb_dirichlet"] = np.int32(len(case["dirichlet"]))
        for i, (group, q) in enumerate(case.get("neumann", [])):
            out[f"neumann{i}_faces"] = orient_outward(m.group_elements(group), cells, coords)
            out[f"neumann{i}_flux"] = np.array(q, dtype=np.float64)
        out["nb_neumann"] = np.int32(len(case.get("neumann", [])))
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **out)
        print(name, cells.shape, coords.shape, exp.shape, os.path.getsize(path), "bytes")


def main_elasticity():
    for name, case in ELASTICITY_CASES.items():
        m = read_msh(os.path.join(REF, "meshes", "msh", case["mesh"] + ".msh"))
        cells, coords = m.cells(), m.coords
        exp = np.loadtxt(os.path.join(REF, case["result"]))
        out = dict(kind=np.array("elasticity"), dim=np.int32(m.dim), cells=cells, coords=coords, node_uid=m.node_tags, expected_uid=exp[:, 0].astype(np.int64),
                   expected_vec=exp[:, 1:3], E=np.float64(case["E"]), nu=np.float64(case["nu"]), f2=np.array(case["f"], dtype=np.float64),
                   penalty=np.float64(case["penalty"]), eps=np.float64(case["eps"]))
        for i, (group, value) in enumerate(case["dirichlet"]):
            out[f"dirichlet{i}_nodes"] = m.group_nodes(group)
            out[f"dirichlet{i}_value2"] = np.array(value, dtype=np.float64)
        out["nb_dirichlet"] = np.int32(len(case["dirichlet"]))
        for i, (group, t) in enumerate(case["traction"]):
            out[f"traction{i}_faces"] = m.group_elements(group)
            out[f"traction{i}_t"] = np.array(t, dtype=np.float64)
        out["nb_traction"] = np.int32(len(case["traction"]))
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **out)
        print(name, cells.shape, coords.shape, exp.shape, os.path.getsize(path), "bytes")


# Heat cases (modules/heat): transient, re-assembled every step.  The module ships NO result files (its ctest only runs the
# cases, modules/heat/CMakeLists.txt:43-58), so these fixtures carry the mesh, the groups and the .arc parameters only.
# Its "Penalty" method adds P to the diagonal (modules/heat/FemModule.cc:386-406), i.e. the WeakPenalty of the other modules;
# penalty default 1e30 (modules/heat/Fem.axl:55).
HEAT_CASES = {
    # modules/heat/inputs/conduction.convection.arc
    "heat_plate_convection": dict(mesh="plate", lam=1.75, dt=0.4, tmax=20.0, tinit=30.0, dirichlet=[("left", 10.0)],
                                  convection=[("right", 1.0, 20.0), ("top", 1.0, 20.0), ("bottom", 1.0, 20.0)], penalty=1.0e30),
}


def main_heat():
    for name, case in HEAT_CASES.items():
        m = read_msh(os.path.join(REF, "meshes", "msh", case["mesh"] + ".msh"))
        cells, coords = m.cells(), m.coords
        out = dict(kind=np.array("heat"), dim=np.int32(m.dim), cells=cells, coords=coords, node_uid=m.node_tags, lam=np.float64(case["lam"]),
                   dt=np.float64(case["dt"]), tmax=np.float64(case["tmax"]), tinit=np.float64(case["tinit"]), penalty=np.float64(case["penalty"]))
        for i, (group, value) in enumerate(case["dirichlet"]):
            out[f"dirichlet{i}_nodes"] = m.group_nodes(group)
            out[f"dirichlet{i}_value"] = np.float64(value)
        out["nb_dirichlet"] = np.int32(len(case["dirichlet"]))
        for i, (group, h, text) in enumerate(case["convection"]):
            out[f"convection{i}_faces"] = m.group_elements(group)
            out[f"convection{i}_h"] = np.float64(h)
            out[f"convection{i}_text"] = np.float64(text)
        out["nb_convection"] = np.int32(len(case["convection"]))
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **out)
        print(name, cells.shape, coords.shape, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
    main_elasticity()
    main_heat()
