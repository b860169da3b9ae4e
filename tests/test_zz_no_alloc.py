"""b2f_ij_own_rows_no_alloc (include/b2f.h): the own-row extraction with its scan arrays in a fixed `static __device__`
array of the library instead of cudaMalloc'ed workspace.  No counterpart in the reference,
so the check is against the existing entry point b2f_to_hypre_ij: bit-identical arrays, on calls made
back to back on different systems (every call shares the one array).  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

from tests.util import MESHES

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

MAX_ROWS = 1 << 20  # B2F_NO_ALLOC_MAX_ROWS


def _np(t):
    return t.detach().cpu().numpy()


def _system(mesh_name, bs, kind, with_ghosts, seed):
    """Assembled system + right-hand side on the device, through the existing entry points."""
    import arcanefem_b200 as fem
    from arcanefem_b200.bsr import Mesh, Operator
    dim, cells, coords = MESHES[mesh_name]()
    nn = coords.shape[0]
    rng = np.random.default_rng(seed)
    own = (rng.uniform(size=nn) > 0.3).astype(np.uint8) if with_ghosts else None
    m = Mesh.from_numpy(dim, cells, coords, own)
    bsr = fem.BSRFormat(bs)
    bsr.initialize(m, True)
    bsr.computeSparsity()
    ls = fem.DoFLinearSystem()
    bsr.toLinearSystem(ls)
    ls.clearValues()
    op = Operator.laplace() if kind == "laplace" else Operator.elasticity(21.0e5, 0.28)
    bsr.assembleBilinearWithSource(op, 5.5)
    numbering = torch.as_tensor((rng.permutation(nn * bs) + 77).astype(np.int32)).cuda()
    own_d = torch.as_tensor(own).cuda() if own is not None else None
    return dict(bsr=bsr, mesh=m, ls=ls, own=own_d, numbering=numbering, nb_row=nn * bs, bs=bs)


def _buffers(S, slack_rows=0, slack_values=0, cap_rows=None, cap_values=None):
    """Caller-owned outputs for system S, filled with a sentinel."""
    nrow, nval = S["nb_row"], int(_inputs(S)[0].values().numel())
    cr = nrow + slack_rows if cap_rows is None else cap_rows
    cv = nval + slack_values if cap_values is None else cap_values
    i32 = lambda n: torch.full((n,), -7, dtype=torch.int32, device="cuda")
    f64 = lambda n: torch.full((n,), -7.0, dtype=torch.float64, device="cuda")
    return dict(ncols=i32(cr), rows=i32(cr), row_local=i32(cr), cols=i32(cv), values=f64(cv), rhs=f64(cr),
                totals=torch.full((2,), -7, dtype=torch.int64, device="cuda"))


def _inputs(S):
    """Device views of the CSR arrays and the right-hand side of S (existing entry points; cached on S)."""
    if "csr" not in S:
        S["csr"], S["rhs"] = S["bsr"].matrix().toCsr(), S["ls"].rhsVariable()
    return S["csr"], S["rhs"]


def _call(lib, S, numbering, out):
    """One call of the new entry point on system S, on torch's current stream."""
    from arcanefem_b200 import _lib
    csr, rhs = _inputs(S)
    b = _lib.IjBuffers(out["ncols"].numel(), out["cols"].numel(), out["ncols"].data_ptr(), out["rows"].data_ptr(), out["row_local"].data_ptr(),
                       out["cols"].data_ptr(), out["values"].data_ptr(), out["rhs"].data_ptr(), out["totals"].data_ptr())
    p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.b2f_ij_own_rows_no_alloc(p(csr.rowsWithEnd()), p(csr.rowsNbColumn()), p(csr.columns()), p(csr.values()), p(rhs), p(S["own"]), p(numbering),
                                      S["nb_row"], S["bs"], C.byref(b), stream)
    assert rc == 0, lib.b2f_last_error_string()
    return out


def _no_alloc(lib, S, numbering, **kw):
    return _call(lib, S, numbering, _buffers(S, **kw))


def _compare(S, numbering, out):
    ij = S["bsr"].toHypreIJ(numbering)
    torch.cuda.synchronize()
    nr, nv = int(ij["ncols"].numel()), int(ij["values"].numel())
    assert _np(out["totals"]).tolist() == [nr, nv]
    for k in ("ncols", "rows", "row_local", "rhs"):
        np.testing.assert_array_equal(_np(out[k])[:nr], _np(ij[k]), err_msg=k)  # integer ids and copied fp64: bit-exact
        assert (_np(out[k])[nr:] == -7).all(), f"{k}: written beyond the own rows"
    for k in ("cols", "values"):
        np.testing.assert_array_equal(_np(out[k])[:nv], _np(ij[k]), err_msg=k)
        assert (_np(out[k])[nv:] == -7).all(), f"{k}: written beyond the own entries"
    return nr, nv


def test_no_alloc_matches_to_hypre_ij_on_calls_in_a_row():
    from arcanefem_b200 import _lib
    lib = _lib.load_library()
    # three systems of different size and block size; the larger ones span several 1024-row tiles of the scan
    A = _system("cube12", 1, "laplace", True, 3)  # 2197 rows, ghosts
    B = _system("cube4_mixed_orientation", 3, "elasticity", False, 4)  # 375 rows, BS = 3, no ghosts
    Cc = _system("square20", 2, "elasticity", True, 5)  # 882 rows, BS = 2, ghosts
    for S, num in ((A, A["numbering"]), (B, None), (A, None), (Cc, Cc["numbering"]), (B, B["numbering"]), (A, A["numbering"])):
        out = _no_alloc(lib, S, num, slack_rows=5, slack_values=9)
        nr, nv = _compare(S, num, out)
        assert nr > 0 and nv > 0


def test_no_alloc_runs_under_stream_capture():
    """A stream capture in the global mode refuses cudaMalloc / cudaFree / synchronisation from the capturing thread: the
    call must go through, and the replayed graph must give the same arrays."""
    from arcanefem_b200 import _lib
    lib = _lib.load_library()
    S = _system("cube12", 1, "laplace", True, 11)
    out = _buffers(S)
    _call(lib, S, S["numbering"], out)  # set-up (symbol address of this device) outside the capture
    torch.cuda.synchronize()
    for t in out.values():
        t.fill_(-7)
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    with torch.cuda.graph(g, stream=side, capture_error_mode="global"):
        _call(lib, S, S["numbering"], out)
    torch.cuda.synchronize()
    assert (_np(out["totals"]) == -7).all()  # captured, not run
    for _ in range(2):
        g.replay()
        torch.cuda.synchronize()
        _compare(S, S["numbering"], out)
        for t in out.values():
            t.fill_(-7)


def test_no_alloc_back_to_back_without_synchronising():
    """Two calls queued on one stream with no host synchronisation between them: stream order is all the shared array needs."""
    from arcanefem_b200 import _lib
    lib = _lib.load_library()
    A = _system("cube12", 1, "laplace", True, 7)
    B = _system("cube7_shuffled", 1, "laplace", True, 8)
    torch.cuda.synchronize()
    out_a = _no_alloc(lib, A, A["numbering"])
    out_b = _no_alloc(lib, B, B["numbering"])
    out_a2 = _no_alloc(lib, A, A["numbering"])
    _compare(A, A["numbering"], out_a)
    _compare(B, B["numbering"], out_b)
    _compare(A, A["numbering"], out_a2)


def test_no_alloc_limits():
    from arcanefem_b200 import _lib
    lib = _lib.load_library()
    S = _system("cube5", 1, "laplace", True, 9)
    # beyond the fixed array: refused with a clear message, nothing launched
    b = _lib.IjBuffers()
    one = torch.zeros(8, dtype=torch.int32, device="cuda")
    p = C.c_void_p(one.data_ptr())
    rc = lib.b2f_ij_own_rows_no_alloc(p, None, p, p, None, None, None, MAX_ROWS + 1, 1, C.byref(b), None)
    assert rc == -1 and b"B2F_NO_ALLOC_MAX_ROWS" in lib.b2f_last_error_string()
    assert lib.b2f_ij_own_rows_no_alloc(p, None, p, p, None, None, None, 8, 1, None, None) == -1
    assert lib.b2f_ij_own_rows_no_alloc(None, None, p, p, None, None, None, 8, 1, C.byref(b), None) == -1
    # nb_row = 0: set-up only, totals cleared
    tot = torch.full((2,), 5, dtype=torch.int64, device="cuda")
    b0 = _lib.IjBuffers(0, 0, None, None, None, None, None, None, tot.data_ptr())
    assert lib.b2f_ij_own_rows_no_alloc(None, None, None, None, None, None, None, 0, 1, C.byref(b0), None) == 0
    torch.cuda.synchronize()
    assert _np(tot).tolist() == [0, 0]
    # capacities smaller than the result: the totals still report the full sizes, nothing is written past the capacities
    full = _no_alloc(lib, S, None)
    torch.cuda.synchronize()
    nr, nv = _np(full["totals"]).tolist()
    cut = _no_alloc(lib, S, None, cap_rows=nr // 2, cap_values=nv // 3)
    torch.cuda.synchronize()
    assert _np(cut["totals"]).tolist() == [nr, nv]
    np.testing.assert_array_equal(_np(cut["rows"]), _np(full["rows"])[: nr // 2])
    np.testing.assert_array_equal(_np(cut["values"]), _np(full["values"])[: nv // 3])


def test_no_alloc_at_the_maximum_size():
    """A synthetic CSR matrix with exactly B2F_NO_ALLOC_MAX_ROWS rows (all 1024 tiles of the scan), row lengths 0..7,
    against a numpy walk."""
    from arcanefem_b200 import _lib
    lib = _lib.load_library()
    rng = np.random.default_rng(21)
    n = MAX_ROWS
    ncols = rng.integers(0, 8, size=n).astype(np.int32)
    rp = np.zeros(n + 1, dtype=np.int32)
    np.cumsum(ncols, out=rp[1:])
    nnz = int(rp[-1])
    cols = rng.integers(0, n, size=nnz).astype(np.int32)
    vals = rng.standard_normal(nnz)
    rhs = rng.standard_normal(n)
    own = (rng.uniform(size=n) > 0.25).astype(np.uint8)
    d = {k: torch.as_tensor(v).cuda() for k, v in dict(rp=rp, cols=cols, vals=vals, rhs=rhs, own=own).items()}
    keep = own.astype(bool)
    nr = int(keep.sum())
    mask = np.repeat(keep, ncols)
    nv = int(mask.sum())
    out = dict(ncols=torch.empty(nr, dtype=torch.int32, device="cuda"), rows=torch.empty(nr, dtype=torch.int32, device="cuda"),
               row_local=torch.empty(nr, dtype=torch.int32, device="cuda"), cols=torch.empty(nv, dtype=torch.int32, device="cuda"),
               values=torch.empty(nv, dtype=torch.float64, device="cuda"), rhs=torch.empty(nr, dtype=torch.float64, device="cuda"),
               totals=torch.empty(2, dtype=torch.int64, device="cuda"))
    b = _lib.IjBuffers(nr, nv, out["ncols"].data_ptr(), out["rows"].data_ptr(), out["row_local"].data_ptr(), out["cols"].data_ptr(),
                       out["values"].data_ptr(), out["rhs"].data_ptr(), out["totals"].data_ptr())
    p = lambda t: C.c_void_p(t.data_ptr())
    for _ in range(2):
        rc = lib.b2f_ij_own_rows_no_alloc(p(d["rp"]), None, p(d["cols"]), p(d["vals"]), p(d["rhs"]), p(d["own"]), None, n, 1, C.byref(b),
                                          C.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == 0, lib.b2f_last_error_string()
        torch.cuda.synchronize()
        assert _np(out["totals"]).tolist() == [nr, nv]
        np.testing.assert_array_equal(_np(out["ncols"]), ncols[keep])
        np.testing.assert_array_equal(_np(out["rows"]), np.nonzero(keep)[0].astype(np.int32))
        np.testing.assert_array_equal(_np(out["row_local"]), np.nonzero(keep)[0].astype(np.int32))
        np.testing.assert_array_equal(_np(out["cols"]), cols[mask])
        np.testing.assert_array_equal(_np(out["values"]), vals[mask])
        np.testing.assert_array_equal(_np(out["rhs"]), rhs[keep])
        for t in out.values():
            t.zero_()
