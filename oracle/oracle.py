"""ctypes loader of the CPU oracle (oracle/oracle.cc).  TEST INFRASTRUCTURE ONLY.

May be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, never by
the product package (arcanefem_b200).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
OP_LAPLACE, OP_ELASTICITY, OP_HEAT = 0, 1, 2

_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")


def build():
    subprocess.check_call(["make", "-C", _HERE, "-s"])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        _LIB = C.CDLL(path)
        _LIB.orc_build_pattern.restype = C.c_int64
        _LIB.orc_legacy_assemble_csr.restype = C.c_double
        _LIB.orc_assemble_cellwise.restype = C.c_int64
        _LIB.orc_assemble_nodewise.restype = C.c_int64
        _LIB.orc_add_face_mass.restype = C.c_int64
        _LIB.orc_num_threads.restype = C.c_int
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def num_threads():
    return int(lib().orc_num_threads())


def build_pattern(cell_node, nb_node):
    """K1-K5 executed sequentially -> (row_index[nb_node+1], columns[nnzb])."""
    cell_node = np.ascontiguousarray(cell_node, dtype=np.int32)
    nc, nv = cell_node.shape
    row_index = np.zeros(nb_node + 1, dtype=np.int32)
    cols_ptr = C.POINTER(C.c_int32)()
    nnzb = lib().orc_build_pattern(C.c_int(nv), C.c_int64(nc), C.c_int64(nb_node), _p(cell_node), _p(row_index), C.byref(cols_ptr))
    if nnzb < 0:
        raise MemoryError("orc_build_pattern failed")
    columns = np.ctypeslib.as_array(cols_ptr, shape=(max(nnzb, 1),))[:nnzb].copy()
    lib().orc_free(cols_ptr)
    return row_index, columns


def nz_per_row(row_index):
    return np.diff(row_index).astype(np.int32)


def element_matrix(cell_node, coords, cell, bs=1, op=OP_LAPLACE, lambda_=0.0, mu2=0.0, inv_dt=0.0, cell_lambda=None):
    cell_node = np.ascontiguousarray(cell_node, dtype=np.int32)
    coords = np.ascontiguousarray(coords, dtype=np.float64)
    nv = cell_node.shape[1]
    n = nv * bs
    Ke = np.zeros((n, n), dtype=np.float64)
    lib().orc_element_matrix(C.c_int(nv), C.c_int(bs), C.c_int(op), C.c_double(lambda_), C.c_double(mu2), C.c_double(inv_dt), _p(cell_lambda),
                             C.c_int64(cell), _p(cell_node), _p(coords), _p(Ke))
    return Ke


def assemble(cell_node, coords, row_index, columns, bs=1, per_block=True, op=OP_LAPLACE, lambda_=0.0, mu2=0.0, inv_dt=0.0, cell_lambda=None,
             node_is_own=None, nodewise=False, values=None):
    cell_node = np.ascontiguousarray(cell_node, dtype=np.int32)
    coords = np.ascontiguousarray(coords, dtype=np.float64)
    nc, nv = cell_node.shape
    nn = coords.shape[0]
    if values is None:
        values = np.zeros(columns.shape[0] * bs * bs, dtype=np.float64)
    fn = lib().orc_assemble_nodewise if nodewise else lib().orc_assemble_cellwise
    nf = fn(C.c_int(nv), C.c_int(bs), C.c_int(1 if per_block else 0), C.c_int(op), C.c_double(lambda_), C.c_double(mu2), C.c_double(inv_dt), _p(cell_lambda),
            C.c_int64(nc), C.c_int64(nn), _p(cell_node), _p(coords), _p(node_is_own), _p(row_index), _p(columns), _p(values))
    return values, int(nf)


def to_csr(bs, row_index, columns):
    nn = row_index.shape[0] - 1
    nnz = columns.shape[0] * bs * bs
    rows = np.zeros(nn * bs + 1, dtype=np.int32)
    nbcol = np.zeros(nn * bs, dtype=np.int32)
    cols = np.zeros(nnz, dtype=np.int32)
    lib().orc_to_csr(C.c_int(bs), C.c_int64(nn), _p(row_index), _p(columns), _p(rows), _p(nbcol), _p(cols))
    return rows, nbcol, cols


def csr_spmv(rows, columns, values, x):
    y = np.zeros(rows.shape[0] - 1, dtype=np.float64)
    lib().orc_csr_spmv(C.c_int64(rows.shape[0] - 1), _p(rows), _p(columns), _p(values), _p(np.ascontiguousarray(x, dtype=np.float64)), _p(y))
    return y


def rhs_constant_source(cell_node, coords, bs, qdot, rhs, node_is_own=None):
    cell_node = np.ascontiguousarray(cell_node, dtype=np.int32)
    nc, nv = cell_node.shape
    lib().orc_rhs_constant_source(C.c_int(nv), C.c_int(bs), C.c_double(qdot), C.c_int64(nc), _p(cell_node), _p(coords), _p(node_is_own), _p(rhs))
    return rhs


def rhs_nodal_source(cell_node, coords, bs, scale, node_field, rhs, node_skip=None, node_is_own=None):
    """modules/heat/FemModule.cc:480-495: rhs[n] += (field[n] * scale) * |cell| / Nv (signed 2D area like the module)."""
    cell_node = np.ascontiguousarray(cell_node, dtype=np.int32)
    nc, nv = cell_node.shape
    lib().orc_rhs_nodal_source(C.c_int(nv), C.c_int(bs), C.c_double(scale), _p(np.ascontiguousarray(node_field, dtype=np.float64)), _p(node_skip), C.c_int64(nc),
                               _p(cell_node), _p(coords), _p(node_is_own), _p(rhs))


def rhs_neumann(face_node, coords, bs, rhs, value=None, value_xyz=(0.0, 0.0, 0.0), face_outside=None, node_is_own=None):
    face_node = np.ascontiguousarray(face_node, dtype=np.int32)
    nf, nvf = face_node.shape
    lib().orc_rhs_neumann(C.c_int(nvf), C.c_int(bs), C.c_int(1 if value is not None else 0), C.c_double(value if value is not None else 0.0),
                          C.c_double(value_xyz[0]), C.c_double(value_xyz[1]), C.c_double(value_xyz[2]), C.c_int64(nf), _p(face_node), _p(face_outside),
                          _p(coords), _p(node_is_own), _p(rhs))
    return rhs


def add_face_mass(face_node, coords, h, row_index, columns, values, node_is_own=None):
    """Convection boundary term (modules/heat/FemModule.cc:628-672,813-840), BS = 1; returns the not-found count."""
    face_node = np.ascontiguousarray(face_node, dtype=np.int32)
    nf, nvf = face_node.shape
    return int(lib().orc_add_face_mass(C.c_int(nvf), C.c_double(h), C.c_int64(nf), _p(face_node), _p(coords), _p(node_is_own), _p(row_index), _p(columns),
                                       _p(values)))


def dirichlet_flag_penalty(bs, dof_index, value, penalty, node_ids, forced_info, forced_value, rhs, node_is_own=None, weak=False):
    node_ids = np.ascontiguousarray(node_ids, dtype=np.int32)
    lib().orc_dirichlet_flag_penalty(C.c_int(bs), C.c_int(dof_index), C.c_double(value), C.c_double(penalty), C.c_int(1 if weak else 0), C.c_int64(node_ids.shape[0]), _p(node_ids),
                                     _p(node_is_own), _p(forced_info), _p(forced_value), _p(rhs))


def dirichlet_flag_elimination(bs, dof_index, value, mode, node_ids, elimination_info, elimination_value, node_is_own=None):
    node_ids = np.ascontiguousarray(node_ids, dtype=np.int32)
    lib().orc_dirichlet_flag_elimination(C.c_int(bs), C.c_int(dof_index), C.c_double(value), C.c_int(mode), C.c_int64(node_ids.shape[0]), _p(node_ids),
                                         _p(node_is_own), _p(elimination_info), _p(elimination_value))


def dirichlet_apply(rows, columns, values, elimination_info, elimination_value, forced_info, forced_value, rhs):
    lib().orc_dirichlet_apply(C.c_int64(rows.shape[0] - 1), _p(rows), _p(columns), _p(values), _p(elimination_info), _p(elimination_value), _p(forced_info),
                              _p(forced_value), _p(rhs))


def legacy_assemble_csr(cell_node, coords, row_index, columns, values, threads=1, op=OP_LAPLACE, lambda_=0.0, mu2=0.0, inv_dt=0.0):
    """Timed CPU baseline (reference legacy loop); returns seconds."""
    cell_node = np.ascontiguousarray(cell_node, dtype=np.int32)
    nc, nv = cell_node.shape
    return float(lib().orc_legacy_assemble_csr(C.c_int(nv), C.c_int(op), C.c_double(lambda_), C.c_double(mu2), C.c_double(inv_dt), C.c_int64(nc), _p(cell_node),
                                               _p(coords), _p(row_index), _p(columns), _p(values), C.c_int(threads)))
