"""Python host mirror of femutils' BSR assembly API over the C ABI (include/b2f.h).

Names and call order follow the reference so tests read like its callers
(modules/poisson/FemModule.cc:24-117, modules/elasticity/FemModule.cc:74-195):

    bsr = BSRFormat(nb_dof)                                  # BSRFormat<NB_DOF>(tm, queue, dofs), BSRFormat.h:358
    bsr.initialize(mesh, does_linear_system_use_csr, use_atomic_free)   # :389
    bsr.computeSparsity()                                    # :775
    bsr.assembleBilinear(Operator.laplace())                 # :1105 (element lambda -> operator descriptor)
    bsr.toLinearSystem(linear_system)                        # :414
    ... BC helpers (ArcaneFemFunctionsGpu.h:434-768) ...
    linear_system.applyDirichlet()                           # prelude of HypreDoFLinearSystemImpl::solve(), :387-391

torch is used only to own device memory and to expose the library's device arrays as tensors (zero copy).
"""
import ctypes as C
from dataclasses import dataclass
from typing import Optional

import numpy as np

from . import _lib
from ._lib import check

ASM_DEFAULT, ASM_CELL_ATOMIC, ASM_CELL_ATOMIC_WARPAGG, ASM_NODE_GATHER, ASM_NODE_GATHER_SLOTS = 0, 1, 2, 3, 4
PATTERN_DIAG_FIRST, PATTERN_DIAG_LAST = 0, 1
OP_LAPLACE, OP_ELASTICITY, OP_HEAT = 0, 1, 2
ELIMINATE_NONE, ELIMINATE_ROW, ELIMINATE_ROW_COLUMN = 0, 1, 2


def _torch():
    import torch
    return torch


def _stream():
    return C.c_void_p(_torch().cuda.current_stream().cuda_stream)


class _DevArray:
    """Zero-copy view of a library-owned device array through __cuda_array_interface__."""

    def __init__(self, ptr, n, typestr, owner):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr, "data": (int(ptr), False), "version": 2}
        self._owner = owner


def _as_tensor(ptr, n, typestr, owner):
    torch = _torch()
    if n == 0 or not ptr:
        dt = {"<f8": torch.float64, "<i4": torch.int32, "|u1": torch.uint8}[typestr]
        return torch.empty(0, dtype=dt, device="cuda")
    t = torch.as_tensor(_DevArray(ptr, n, typestr, owner), device="cuda")
    t._b2f_owner = owner  # keep the handle alive while the view exists
    return t


@dataclass
class Mesh:
    """What the reference kernels read from Arcane: cellNode() connectivity, VariableNodeReal3, isOwn."""
    dim: int
    cell_node: "object"  # torch.int32 cuda [nb_cell, dim+1]
    coords: "object"  # torch.float64 cuda [nb_node, 3]
    node_is_own: Optional["object"] = None  # torch.uint8 cuda [nb_node] or None
    cell_is_own: Optional["object"] = None  # torch.uint8 cuda [nb_cell] or None (0 = ghost cell: pattern only)

    @property
    def nb_cell(self):
        return int(self.cell_node.shape[0])

    @property
    def nb_node(self):
        return int(self.coords.shape[0])

    @staticmethod
    def from_numpy(dim, cell_node, coords, node_is_own=None, device="cuda", cell_is_own=None):
        torch = _torch()
        cn = torch.as_tensor(np.ascontiguousarray(cell_node, dtype=np.int32)).to(device)
        xy = torch.as_tensor(np.ascontiguousarray(coords, dtype=np.float64)).to(device)
        own = None if node_is_own is None else torch.as_tensor(np.ascontiguousarray(node_is_own, dtype=np.uint8)).to(device)
        cown = None if cell_is_own is None else torch.as_tensor(np.ascontiguousarray(cell_is_own, dtype=np.uint8)).to(device)
        return Mesh(dim, cn, xy, own, cown)


class Operator:
    """Descriptor replacing the reference's element lambda (BSRFormat.h:786-787)."""

    def __init__(self, kind, lambda_=0.0, mu2=0.0, inv_dt=0.0, cell_lambda=None):
        self.kind, self.lambda_, self.mu2, self.inv_dt, self.cell_lambda = kind, float(lambda_), float(mu2), float(inv_dt), cell_lambda

    @staticmethod
    def laplace():
        """modules/poisson/FemModule.h:139-147 (Tria3) / :177-186 (Tetra4)."""
        return Operator(OP_LAPLACE)

    @staticmethod
    def elasticity(E, nu):
        """modules/elasticity/FemModule.h:112-152 with mu2/lambda of FemModule.cc:133-134."""
        mu2 = (E / (2 * (1 + nu))) * 2
        lam = E * nu / ((1 + nu) * (1 - 2 * nu))
        return Operator(OP_ELASTICITY, lambda_=lam, mu2=mu2)

    @staticmethod
    def heat(lambda_, dt, cell_lambda=None):
        """modules/heat/FemModule.cc:678-770."""
        return Operator(OP_HEAT, lambda_=lambda_, inv_dt=1.0 / dt, cell_lambda=cell_lambda)

    def desc(self, accumulate=False):
        d = _lib.OperatorDesc()
        d.kind = self.kind
        d.accumulate = 1 if accumulate else 0
        d.lambda_ = self.lambda_
        d.mu2 = self.mu2
        d.inv_dt = self.inv_dt
        d.cell_lambda = C.c_void_p(self.cell_lambda.data_ptr()) if self.cell_lambda is not None else None
        return d


class CSRFormatView:
    """femutils/DoFLinearSystem.h:42-76."""

    def __init__(self, rows, rows_nb_column, columns, values):
        self._rows, self._rows_nb_column, self._columns, self._values = rows, rows_nb_column, columns, values

    def rows(self):
        return self._rows[:-1]  # the reference view has nbRow entries, no closing entry

    def rowsWithEnd(self):
        return self._rows

    def rowsNbColumn(self):
        return self._rows_nb_column

    def columns(self):
        return self._columns

    def values(self):
        return self._values

    def nbRow(self):
        return int(self._rows.shape[0]) - 1

    def nbColumn(self):
        return int(self._columns.shape[0])

    def nbValue(self):
        return int(self._values.shape[0])


class BSRMatrix:
    """femutils/BSRFormat.h:104-336 (views on the handle's device arrays)."""

    def __init__(self, fmt):
        self._fmt = fmt

    def _view(self):
        v = _lib.BsrView()
        check(self._fmt._lib.b2f_get_bsr(self._fmt._h, C.byref(v)))
        return v

    def orderValuePerBlock(self):
        return self._view().layout == 0

    def nbNz(self):
        return int(self._view().nb_nz)

    def nbCol(self):
        return int(self._view().nb_col)

    def nbRow(self):
        return int(self._view().nb_row)

    def nbEdge(self):
        return int(self._view().nb_edge)

    def values(self):
        v = self._view()
        return _as_tensor(v.values, v.nb_nz, "<f8", self._fmt)

    def columns(self):
        v = self._view()
        return _as_tensor(v.columns, v.nb_col, "<i4", self._fmt)

    def rowIndex(self):
        """nb_row entries like the reference (BSRFormat.h:138); rowIndexWithEnd() adds the closing entry."""
        return self.rowIndexWithEnd()[:-1]

    def rowIndexWithEnd(self):
        v = self._view()
        return _as_tensor(v.row_index, v.nb_row + 1, "<i4", self._fmt)

    def nbNzPerRow(self):
        v = self._view()
        return _as_tensor(v.nb_nz_per_row, v.nb_row, "<i4", self._fmt)

    def getValue(self, row, col):
        out = C.c_double()
        check(self._fmt._lib.b2f_get_value(self._fmt._h, int(row), int(col), C.byref(out)))
        return out.value

    def setValue(self, row, col, value):
        check(self._fmt._lib.b2f_set_value(self._fmt._h, int(row), int(col), float(value)))

    def toCsr(self):
        v = _lib.CsrView()
        check(self._fmt._lib.b2f_to_csr(self._fmt._h, C.byref(v), _stream()))
        f = self._fmt
        return CSRFormatView(_as_tensor(v.rows, v.nb_row + 1, "<i4", f), _as_tensor(v.rows_nb_column, v.nb_row, "<i4", f),
                             _as_tensor(v.columns, v.nb_value, "<i4", f), _as_tensor(v.values, v.nb_value, "<f8", f))

    def dump(self, filename):
        check(self._fmt._lib.b2f_dump(self._fmt._h, str(filename).encode()))


class DoFLinearSystem:
    """Device-resident part of femutils/DoFLinearSystem.h:126-286 that the assembly path touches (no solver)."""

    def __init__(self):
        self._fmt = None
        self._csr = None

    def _sys(self):
        if self._fmt is None:
            raise RuntimeError("DoFLinearSystem is not attached: call BSRFormat.toLinearSystem(linear_system) first")
        v = _lib.SystemView()
        check(self._fmt._lib.b2f_get_system(self._fmt._h, C.byref(v)))
        return v

    def hasSetCSRValues(self):
        return True

    def setCSRValues(self, csr_view):
        self._csr = csr_view

    def getCSRValues(self):
        return self._csr

    def rhsVariable(self):
        v = self._sys()
        return _as_tensor(v.rhs, v.nb_dof, "<f8", self._fmt)

    def getForcedInfo(self):
        v = self._sys()
        return _as_tensor(v.forced_info, v.nb_dof, "|u1", self._fmt)

    def getForcedValue(self):
        v = self._sys()
        return _as_tensor(v.forced_value, v.nb_dof, "<f8", self._fmt)

    def getEliminationInfo(self):
        v = self._sys()
        return _as_tensor(v.elimination_info, v.nb_dof, "|u1", self._fmt)

    def getEliminationValue(self):
        v = self._sys()
        return _as_tensor(v.elimination_value, v.nb_dof, "<f8", self._fmt)

    def clearValues(self):
        self._sys()
        check(self._fmt._lib.b2f_clear_values(self._fmt._h, _stream()))

    def applyDirichlet(self):
        """_applyRowElimination() + _applyForcedValuesToLhs() (+ row-column elimination)."""
        self._sys()
        check(self._fmt._lib.b2f_dirichlet_apply(self._fmt._h, _stream()))


class BSRFormat:
    """femutils/BSRFormat.h:353-1140."""

    def __init__(self, nb_dof):
        self._lib = _lib.load_library()
        self._nb_dof = int(nb_dof)
        self._h = C.c_void_p()
        self._mesh = None
        self._use_csr = True
        self._keep = []

    def __del__(self):
        try:
            if getattr(self, "_h", None) and self._h.value:
                self._lib.b2f_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def initialize(self, mesh, does_linear_system_use_csr, use_atomic_free=False, assembly_variant=None, fan_mode=0, sparsity_build=0, pattern_layout=0, reorder=0):
        if mesh is None:
            raise ValueError("BSRFormat(initialize): null mesh")  # ARCANE_CHECK_POINTER, BSRFormat.h:391
        if self._h.value:
            self._lib.b2f_destroy(self._h)
            self._h = C.c_void_p()
        mv = _lib.MeshView()
        mv.dim = mesh.dim
        mv.nodes_per_cell = int(mesh.cell_node.shape[1]) if mesh.cell_node.dim() == 2 else mesh.dim + 1
        mv.nb_cell = mesh.nb_cell
        mv.nb_node = mesh.nb_node
        mv.cell_node = C.c_void_p(mesh.cell_node.data_ptr())
        mv.coords = C.c_void_p(mesh.coords.data_ptr())
        mv.node_is_own = C.c_void_p(mesh.node_is_own.data_ptr()) if mesh.node_is_own is not None else None
        mv.cell_is_own = C.c_void_p(mesh.cell_is_own.data_ptr()) if mesh.cell_is_own is not None else None
        opt = _lib.Options()
        opt.block_size = self._nb_dof
        opt.use_csr = 1 if does_linear_system_use_csr else 0
        if assembly_variant is None:
            assembly_variant = ASM_NODE_GATHER if use_atomic_free else ASM_DEFAULT
        opt.assembly_variant = int(assembly_variant)
        opt.fan_mode = int(fan_mode)  # b2f_fan_mode: 0 auto, 1 select walk, 2 lockstep walk
        opt.sparsity_build = int(sparsity_build)  # b2f_sparsity_build: 0 auto, 1 global sort, 2 row-wise from the node->cell lists
        opt.reorder = int(reorder)  # b2f_reorder: 1 = the library sorts the nodes along a Morton curve internally (views stay in the caller's numbering)
        opt.pattern_layout = int(pattern_layout)  # b2f_pattern_layout: 0 diagonal first (sorted-edge build), 1 diagonal last (connectivity build, BSRFormat.h:479-578)
        check(self._lib.b2f_create(C.byref(mv), C.byref(opt), C.byref(self._h)))
        self._mesh = mesh
        self._use_csr = bool(does_linear_system_use_csr)

    def computeSparsity(self):
        check(self._lib.b2f_compute_sparsity(self._h, _stream()))

    def assembleBilinear(self, op, accumulate=False):
        d = op.desc(accumulate)
        check(self._lib.b2f_assemble(self._h, C.byref(d), _stream()))

    def assembleBilinearWithSource(self, op, qdot, accumulate=False):
        """assembleBilinear + applyConstantSourceToRhs(qdot) in one pass over the mesh (b2f_assemble_with_source)."""
        d = op.desc(accumulate)
        check(self._lib.b2f_assemble_with_source(self._h, C.byref(d), float(qdot), _stream()))

    def assembleBilinearHost(self, op, coords_host, values_host=None):
        """Host-buffer entry point b2f_assemble_host (pinned tensors recommended)."""
        d = op.desc(False)
        vp = C.c_void_p(values_host.data_ptr()) if values_host is not None else None
        check(self._lib.b2f_assemble_host(self._h, C.byref(d), C.c_void_p(coords_host.data_ptr()), vp, _stream()))

    def setCoords(self, coords):
        self._keep.append(coords)
        check(self._lib.b2f_set_coords(self._h, C.c_void_p(coords.data_ptr())))

    def toLinearSystem(self, linear_system):
        """femutils/BSRFormat.h:414-430: a CSR consumer gets the CSR view; any other linear system (the per-block branch,
        BSRMatrix::toLinearSystem :261-280) is attached to the same device-resident values."""
        linear_system._fmt = self  # attached first: Dirichlet / rhs entries work for either layout
        if self._use_csr or self._nb_dof == 1:
            if not linear_system.hasSetCSRValues():
                raise ValueError("BSRFormat(toLinearSystem): Linear system was set to use CSR but is incompatible")
            linear_system.setCSRValues(self.matrix().toCsr())

    def matrix(self):
        return BSRMatrix(self)

    def resetMatrixValues(self):
        check(self._lib.b2f_reset_values(self._h, _stream()))

    def dumpMatrix(self, filename):
        self.matrix().dump(filename)

    def stats(self):
        s = _lib.Stats()
        check(self._lib.b2f_get_stats(self._h, C.byref(s)))
        return {k: int(getattr(s, k)) for k, _ in _lib.Stats._fields_}

    def csrGlobalNumbering(self, dof_number):
        """(rows_global, cols_global) of the CSR view for the device int32 numbering array `dof_number` [nb_dof]."""
        torch = _torch()
        v = _lib.CsrView()
        check(self._lib.b2f_to_csr(self._h, C.byref(v), _stream()))
        rows = torch.empty(int(v.nb_row), dtype=torch.int32, device="cuda")
        cols = torch.empty(int(v.nb_value), dtype=torch.int32, device="cuda")
        check(self._lib.b2f_csr_global_numbering(self._h, C.c_void_p(dof_number.data_ptr()), C.c_void_p(rows.data_ptr()), C.c_void_p(cols.data_ptr()), _stream()))
        return rows, cols

    def toHypreIJ(self, dof_number=None):
        """Own rows of the system as the device arrays HYPRE_IJMatrixSetValues / HYPRE_IJVectorSetValues take
        (b2f_to_hypre_ij): dict(ncols, rows, row_local, cols, values, rhs) of tensor views."""
        v = _lib.IjView()
        num = C.c_void_p(dof_number.data_ptr()) if dof_number is not None else None
        check(self._lib.b2f_to_hypre_ij(self._h, num, C.byref(v), _stream()))
        nr, nv = int(v.nb_row), int(v.nb_value)
        out = {"ncols": _as_tensor(v.ncols, nr, "<i4", self), "rows": _as_tensor(v.rows, nr, "<i4", self), "row_local": _as_tensor(v.row_local, nr, "<i4", self),
               "cols": _as_tensor(v.cols, nv, "<i4", self), "values": _as_tensor(v.values, nv, "<f8", self)}
        if v.rhs:
            out["rhs"] = _as_tensor(v.rhs, nr, "<f8", self)
        return out

    def plan(self):
        """Gather plan arrays as tensors (views; "<u4" words are exposed as int32 bit patterns)."""
        v = _lib.PlanView()
        check(self._lib.b2f_get_plan(self._h, C.byref(v)))
        out = {"nc_ptr": _as_tensor(v.nc_ptr, v.nb_node + 1, "<i4", self), "nc_list": _as_tensor(v.nc_list, v.nb_pair, "<i4", self)}
        if v.fan_ptr:
            out["fan_ptr"] = _as_tensor(v.fan_ptr, v.nb_node + 1, "<i4", self)
            out["fan_steps"] = _as_tensor(v.fan_steps, v.nb_fan_step, "<i4", self)
            out["fan_cell"] = _as_tensor(v.fan_cell, v.nb_fan_step, "<i4", self)
        return out

    def spmv(self, x, y=None):
        torch = _torch()
        if y is None:
            y = torch.empty_like(x)
        check(self._lib.b2f_spmv(self._h, C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()), _stream()))
        return y

    # --- RHS / boundary-condition helpers (FemUtils::Gpu::BoundaryConditions*, ArcaneFemFunctionsGpu.h:434-768)
    def applyConstantSourceToRhs(self, qdot):
        check(self._lib.b2f_rhs_constant_source(self._h, float(qdot), _stream()))

    def applyNodalSourceToRhs(self, node_field, scale, node_skip=None):
        """modules/heat/FemModule.cc:480-495: rhs[n] += (node_field[n] * scale) * |cell| / Nv (T_old / dt of the transient heat module)."""
        check(self._lib.b2f_rhs_nodal_source(self._h, C.c_void_p(node_field.data_ptr()), float(scale),
                                             C.c_void_p(node_skip.data_ptr()) if node_skip is not None else None, _stream()))

    def addFaceMass(self, face_node, coefficient):
        """Convection boundary term of the heat module (modules/heat/FemModule.cc:628-672,813-840): K += h * face mass matrix."""
        nf = int(face_node.shape[0])
        check(self._lib.b2f_add_face_mass(self._h, nf, C.c_void_p(face_node.data_ptr()) if nf else None, float(coefficient), _stream()))

    def applyNeumannToRhs(self, face_node, value=None, value_xyz=None, face_outside=None):
        bc = _lib.Neumann()
        bc.nb_face = int(face_node.shape[0])
        bc.face_node = C.c_void_p(face_node.data_ptr()) if bc.nb_face else None
        bc.face_outside = C.c_void_p(face_outside.data_ptr()) if face_outside is not None else None
        bc.has_value = 1 if value is not None else 0
        bc.value = float(value) if value is not None else 0.0
        vx, vy, vz = (value_xyz if value_xyz is not None else (0.0, 0.0, 0.0))
        bc.value_x, bc.value_y, bc.value_z = float(vx), float(vy), float(vz)
        check(self._lib.b2f_rhs_neumann(self._h, C.byref(bc), _stream()))

    def applyDirichletViaPenalty(self, node_ids, value, penalty, dof_index=0):
        check(self._lib.b2f_dirichlet_flag_penalty(self._h, C.c_void_p(node_ids.data_ptr()), int(node_ids.numel()), int(dof_index), float(value),
                                                   float(penalty), _stream()))

    def applyDirichletViaWeakPenalty(self, node_ids, value, penalty, dof_index=0):
        check(self._lib.b2f_dirichlet_flag_weak_penalty(self._h, C.c_void_p(node_ids.data_ptr()), int(node_ids.numel()), int(dof_index), float(value),
                                                        float(penalty), _stream()))

    def applyDirichletViaRowElimination(self, node_ids, value, dof_index=0, mode=ELIMINATE_ROW):
        check(self._lib.b2f_dirichlet_flag_elimination(self._h, C.c_void_p(node_ids.data_ptr()), int(node_ids.numel()), int(dof_index), float(value),
                                                       int(mode), _stream()))

    def rhsFill(self, value):
        check(self._lib.b2f_rhs_fill(self._h, float(value), _stream()))
