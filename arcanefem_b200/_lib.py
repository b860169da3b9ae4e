"""ctypes binding of include/b2f.h (one prototype per exported symbol)."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))


def lib_path():
    """In-tree library; B2F_LIBRARY selects another build of the same library (kernel tuning experiments)."""
    return os.environ.get("B2F_LIBRARY") or os.path.join(_HERE, "lib", "libb200fem.so")


class B2FError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"b2f status {status}: {message}")
        self.status = status
        self.message = message


class MeshView(C.Structure):
    _fields_ = [("dim", C.c_int32), ("nodes_per_cell", C.c_int32), ("nb_cell", C.c_int64), ("nb_node", C.c_int64),
                ("cell_node", C.c_void_p), ("coords", C.c_void_p), ("node_is_own", C.c_void_p), ("cell_is_own", C.c_void_p)]


class Options(C.Structure):
    _fields_ = [("block_size", C.c_int32), ("use_csr", C.c_int32), ("assembly_variant", C.c_int32), ("fan_mode", C.c_int32),
                ("sparsity_build", C.c_int32), ("pattern_layout", C.c_int32), ("reorder", C.c_int32), ("reserved", C.c_int32)]


class OperatorDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("accumulate", C.c_int32), ("lambda_", C.c_double), ("mu2", C.c_double),
                ("inv_dt", C.c_double), ("cell_lambda", C.c_void_p)]


class BsrView(C.Structure):
    _fields_ = [("block_size", C.c_int32), ("layout", C.c_int32), ("nb_row", C.c_int64), ("nb_col", C.c_int64),
                ("nb_nz", C.c_int64), ("nb_edge", C.c_int64), ("row_index", C.c_void_p), ("columns", C.c_void_p),
                ("nb_nz_per_row", C.c_void_p), ("values", C.c_void_p)]


class CsrView(C.Structure):
    _fields_ = [("nb_row", C.c_int64), ("nb_value", C.c_int64), ("rows", C.c_void_p), ("rows_nb_column", C.c_void_p),
                ("columns", C.c_void_p), ("values", C.c_void_p)]


class SystemView(C.Structure):
    _fields_ = [("nb_dof", C.c_int64), ("rhs", C.c_void_p), ("forced_info", C.c_void_p), ("forced_value", C.c_void_p),
                ("elimination_info", C.c_void_p), ("elimination_value", C.c_void_p)]


class Neumann(C.Structure):
    _fields_ = [("nb_face", C.c_int64), ("face_node", C.c_void_p), ("face_outside", C.c_void_p), ("has_value", C.c_int32),
                ("reserved", C.c_int32), ("value", C.c_double), ("value_x", C.c_double), ("value_y", C.c_double),
                ("value_z", C.c_double)]


class PlanView(C.Structure):
    _fields_ = [("nb_node", C.c_int64), ("nb_pair", C.c_int64), ("nb_fan_step", C.c_int64), ("nc_ptr", C.c_void_p), ("nc_list", C.c_void_p),
                ("fan_ptr", C.c_void_p), ("fan_steps", C.c_void_p), ("fan_cell", C.c_void_p)]


class IjView(C.Structure):
    _fields_ = [("nb_row", C.c_int64), ("nb_value", C.c_int64), ("ncols", C.c_void_p), ("rows", C.c_void_p), ("row_local", C.c_void_p),
                ("cols", C.c_void_p), ("values", C.c_void_p), ("rhs", C.c_void_p)]


class IjBuffers(C.Structure):
    """b2f_ij_buffers: caller-owned outputs of b2f_ij_own_rows_no_alloc."""
    _fields_ = [("capacity_rows", C.c_int64), ("capacity_values", C.c_int64), ("ncols", C.c_void_p), ("rows", C.c_void_p), ("row_local", C.c_void_p),
                ("cols", C.c_void_p), ("values", C.c_void_p), ("rhs", C.c_void_p), ("totals_out", C.c_void_p)]


class Stats(C.Structure):
    _fields_ = [("nb_edge", C.c_int64), ("nb_col", C.c_int64), ("nb_nz", C.c_int64), ("columns_not_found", C.c_int64),
                ("workspace_bytes", C.c_int64), ("resident_bytes", C.c_int64), ("gather_fallback_tiles", C.c_int64), ("fan_steps", C.c_int64),
                ("sparsity_by_lists", C.c_int64), ("fan_lockstep", C.c_int64), ("fan_lockstep_bodies_milli", C.c_int64), ("reordered", C.c_int64)]


_H = C.c_void_p  # b2f_handle
_S = C.c_void_p  # cudaStream_t
_P = C.POINTER

# symbol -> (restype, argtypes); every function declared in include/b2f.h
PROTOTYPES = {
    "b2f_version": (C.c_int, []),
    "b2f_last_error_string": (C.c_char_p, []),
    "b2f_status_string": (C.c_char_p, [C.c_int]),
    "b2f_create": (C.c_int, [_P(MeshView), _P(Options), _P(_H)]),
    "b2f_destroy": (C.c_int, [_H]),
    "b2f_set_coords": (C.c_int, [_H, C.c_void_p]),
    "b2f_compute_sparsity": (C.c_int, [_H, _S]),
    "b2f_release_workspace": (C.c_int, [_H]),
    "b2f_assemble": (C.c_int, [_H, _P(OperatorDesc), _S]),
    "b2f_assemble_with_source": (C.c_int, [_H, _P(OperatorDesc), C.c_double, _S]),
    "b2f_user_assemble_begin": (C.c_int, [_H, C.c_int32, C.c_int32, C.c_void_p, C.c_size_t, _P(C.c_int32), _P(C.c_int32), _P(C.c_int64), _S]),
    "b2f_reset_values": (C.c_int, [_H, _S]),
    "b2f_get_bsr": (C.c_int, [_H, _P(BsrView)]),
    "b2f_to_csr": (C.c_int, [_H, _P(CsrView), _S]),
    "b2f_to_hypre_ij": (C.c_int, [_H, C.c_void_p, _P(IjView), _S]),
    "b2f_csr_global_numbering": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_void_p, _S]),
    "b2f_get_system": (C.c_int, [_H, _P(SystemView)]),
    "b2f_get_stats": (C.c_int, [_H, _P(Stats)]),
    "b2f_get_plan": (C.c_int, [_H, _P(PlanView)]),
    "b2f_set_value": (C.c_int, [_H, C.c_int64, C.c_int64, C.c_double]),
    "b2f_get_value": (C.c_int, [_H, C.c_int64, C.c_int64, _P(C.c_double)]),
    "b2f_dump": (C.c_int, [_H, C.c_char_p]),
    "b2f_clear_values": (C.c_int, [_H, _S]),
    "b2f_rhs_fill": (C.c_int, [_H, C.c_double, _S]),
    "b2f_rhs_constant_source": (C.c_int, [_H, C.c_double, _S]),
    "b2f_rhs_nodal_source": (C.c_int, [_H, C.c_void_p, C.c_double, C.c_void_p, _S]),
    "b2f_rhs_neumann": (C.c_int, [_H, _P(Neumann), _S]),
    "b2f_add_face_mass": (C.c_int, [_H, C.c_int64, C.c_void_p, C.c_double, _S]),
    "b2f_dirichlet_flag_penalty": (C.c_int, [_H, C.c_void_p, C.c_int64, C.c_int32, C.c_double, C.c_double, _S]),
    "b2f_dirichlet_flag_weak_penalty": (C.c_int, [_H, C.c_void_p, C.c_int64, C.c_int32, C.c_double, C.c_double, _S]),
    "b2f_dirichlet_flag_elimination": (C.c_int, [_H, C.c_void_p, C.c_int64, C.c_int32, C.c_double, C.c_int32, _S]),
    "b2f_dirichlet_apply": (C.c_int, [_H, _S]),
    "b2f_spmv": (C.c_int, [_H, C.c_void_p, C.c_void_p, _S]),
    "b2f_assemble_host": (C.c_int, [_H, _P(OperatorDesc), C.c_void_p, C.c_void_p, _S]),
    "b2f_ghost_plan": (C.c_int, [_H, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, _S]),
    "b2f_ghost_exchange": (C.c_int, [_H, C.c_void_p, _S]),
    "b2f_assemble_exchange": (C.c_int, [_H, _P(OperatorDesc), C.c_void_p, _S]),
    "b2f_assemble_exchange_host": (C.c_int, [_H, _P(OperatorDesc), C.c_void_p, C.c_void_p, C.c_void_p, _S]),
    "b2f_nccl_unique_id": (C.c_int, [C.c_void_p]),
    "b2f_nccl_comm_init": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, _P(C.c_void_p)]),
    "b2f_nccl_comm_destroy": (C.c_int, [C.c_void_p]),
    "b2f_ij_own_rows_no_alloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32,
                                           _P(IjBuffers), _S]),
}

_lib = None


def load_library():
    """Load libb200fem.so; raise (never fall back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise ImportError(f"{path} is missing: build it with `make` (or __graft_entry__.build()); "
                          "arcanefem_b200 has no CPU fallback")
    lib = C.CDLL(path, mode=C.RTLD_GLOBAL)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status):
    if status != 0:
        msg = load_library().b2f_last_error_string()
        raise B2FError(status, msg.decode("utf-8", "replace") if msg else "")
