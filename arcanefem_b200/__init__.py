"""arcanefem_b200 — B200-native (sm_100a) finite-element bilinear-form assembly.

Host-side Python mirror of the femutils assembly API of ArcaneFEM (reference: toutane/arcanefem,
femutils/BSRFormat.h) over the C ABI of ``libb200fem.so`` (include/b2f.h).  The product is the CUDA library;
this package only binds it for tests and benchmarks.  There is no CPU fallback: importing works anywhere, but
every compute call needs the built library and a CUDA device and fails loudly otherwise.
"""
from ._lib import B2FError, lib_path, load_library  # noqa: F401
from .bsr import (  # noqa: F401
    BSRFormat,
    BSRMatrix,
    CSRFormatView,
    DoFLinearSystem,
    Operator,
    ASM_CELL_ATOMIC,
    ASM_CELL_ATOMIC_WARPAGG,
    ASM_DEFAULT,
    ASM_NODE_GATHER,
    ASM_NODE_GATHER_SLOTS,
)
from . import mesh  # noqa: F401

__all__ = [
    "B2FError", "lib_path", "load_library", "BSRFormat", "BSRMatrix", "CSRFormatView", "DoFLinearSystem", "Operator",
    "ASM_DEFAULT", "ASM_CELL_ATOMIC", "ASM_CELL_ATOMIC_WARPAGG", "ASM_NODE_GATHER", "ASM_NODE_GATHER_SLOTS", "mesh",
]
