"""The C-ABI library loads and exports every symbol include/b2f.h declares (no compute calls: no GPU needed)."""
import ctypes
import os
import re

from arcanefem_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b2f.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b2f_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    names = declared_symbols()
    assert len(names) >= 25
    lib = ctypes.CDLL(_lib.lib_path())
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_python_binding_covers_the_header():
    assert sorted(_lib.PROTOTYPES) == declared_symbols()


def test_version_and_status_strings():
    lib = _lib.load_library()
    assert lib.b2f_version() >= 1
    assert lib.b2f_status_string(0) == b"B2F_OK"
    assert lib.b2f_status_string(-5) == b"B2F_ERR_NOT_FOUND"


def test_null_arguments_are_rejected_without_a_device():
    lib = _lib.load_library()
    assert lib.b2f_create(None, None, None) == -1
    assert b"null" in lib.b2f_last_error_string()
    assert lib.b2f_compute_sparsity(None, None) == -1
    assert lib.b2f_destroy(None) == 0


def test_product_never_touches_the_oracle():
    """The shipped package and the CUDA sources must not import, link or mention the oracle."""
    bad = []
    for base in ("arcanefem_b200", "include"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".cc")):
                    s = open(os.path.join(dp, f), errors="replace").read()
                    if re.search(r"liboracle|from oracle|import oracle|oracle/", s):
                        bad.append(os.path.join(dp, f))
    assert not bad, bad
