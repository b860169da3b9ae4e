"""Multi-GPU host side: Arcane-style slab decomposition of the synthetic Kuhn cube and the NCCL ghost-row exchange.

One process per GPU (torch.distributed supplies rank/world and carries the NCCL unique id to the other ranks; the exchange
itself runs inside libb200fem.so on its own ncclComm_t, include/b2f.h b2f_ghost_plan / b2f_ghost_exchange).

Decomposition (what Arcane hands to every rank, femutils/BSRFormat.h:807-815, femutils/FemDoFsOnNodes.cc:79-128): the local
mesh holds the rank's own cells plus the layer of ghost cells touching its own nodes; a node is owned by the lowest rank
whose own cells touch it.  For z-slabs of `n_hex` hex layers per rank this gives: own nodes = planes r*n+1 .. (r+1)*n
(plus plane 0 on rank 0), ghost nodes = the bottom plane r*n (owner r-1) and the plane (r+1)*n+1 of the ghost-cell layer
on top (owner r+1).
"""
import ctypes as C

import numpy as np

from . import _lib
from . import mesh as gen
from ._lib import check


def plane_coords(n_hex, nz_total, k, L=1.0, perturb=0.25, seed=1234):
    """Coordinates of the (n_hex+1)^2 nodes of global z-plane k (same values on every rank that holds the plane)."""
    n1 = n_hex + 1
    h = L / n_hex
    ax = np.arange(n1, dtype=np.float64) * h
    y, x = np.meshgrid(ax, ax, indexing="ij")
    c = np.stack([x.ravel(), y.ravel(), np.full(n1 * n1, k * h)], axis=1)
    if perturb and 0 < k < nz_total:
        rng = np.random.default_rng(seed + 7919 * k)
        jit = rng.uniform(-perturb * h, perturb * h, size=c.shape)
        idx = np.arange(n1)
        interior = ((idx[:, None] > 0) & (idx[:, None] < n_hex) & (idx[None, :] > 0) & (idx[None, :] < n_hex)).ravel()
        c[interior] += jit[interior]
    return c


def global_mesh(n_hex, world, L=1.0, perturb=0.25, seed=1234, nz_total=None):
    """The undecomposed mesh (cells, coords) of the same cube, global lexicographic node ids (tests only: small sizes)."""
    n1 = n_hex + 1
    if nz_total is None:
        nz_total = n_hex * world
    coords = np.concatenate([plane_coords(n_hex, nz_total, k, L, perturb, seed) for k in range(nz_total + 1)], axis=0)
    hi = np.arange(n_hex, dtype=np.int64)
    kk, jj, ii = np.meshgrid(np.arange(nz_total, dtype=np.int64), hi, hi, indexing="ij")
    base = (ii + n1 * (jj + n1 * kk)).ravel()
    corner = np.array([dx + n1 * (dy + n1 * dz) for dz in (0, 1) for dy in (0, 1) for dx in (0, 1)], dtype=np.int64)
    cells = (base[:, None, None] + corner[gen.kuhn_table()][None, :, :]).reshape(-1, 4).astype(np.int32)
    return np.ascontiguousarray(cells), np.ascontiguousarray(coords)


def slab_bounds(nz_total, world):
    """Hex layers of every rank for a z-slab split of nz_total layers: world + 1 boundaries, the first nz_total % world
    ranks take one layer more (406 layers: 2 x 203; 102 102 101 101; 6 x 51 + 2 x 50)."""
    base, extra = divmod(nz_total, world)
    b = [0]
    for r in range(world):
        b.append(b[-1] + base + (1 if r < extra else 0))
    return b


def slab_partition(n_hex, rank, world, L=1.0, perturb=0.25, seed=1234, nz_total=None, max_layers=None):
    """Local mesh of `rank` for the global box of n_hex x n_hex x nz_total hexes split into z-slabs (slab_bounds).
    nz_total = None: n_hex layers per rank (weak scaling: the cube grows with the ranks); nz_total = n_hex: one fixed cube
    cut into `world` slabs (strong scaling, BASELINE config #5).  max_layers keeps only the first layers of the slab (the
    bounded CPU sample of bench.py; the top plane's rows are then incomplete).

    Returns dict(cells int32 [nc,4], coords [nn,3], cell_is_own uint8 [nc], node_owner int32 [nn], node_gid int64 [nn],
    nb_own_cell, z0 = first local plane, layers = (first, last+1) own hex layers)."""
    n1 = n_hex + 1
    plane = n1 * n1
    if nz_total is None:
        nz_total = n_hex * world
    bounds = slab_bounds(nz_total, world)
    lay0, lay1 = bounds[rank], bounds[rank + 1]  # own hex layers [lay0, lay1)
    ghost_top = 1 if rank < world - 1 else 0
    if max_layers is not None and lay1 - lay0 > max_layers:
        lay1, ghost_top = lay0 + max_layers, 0
    z0, z1 = lay0, lay1 + ghost_top  # local planes z0 .. z1 inclusive
    h_xy = L / n_hex  # cubic hexes: the box is L x L x (nz_total * h)
    coords = np.concatenate([plane_coords(n_hex, nz_total, k, L, perturb, seed) for k in range(z0, z1 + 1)], axis=0)
    del h_xy
    # cells of the local layers, local lexicographic node ids (built layer by layer: int32 throughout, 200 M cells fit)
    nlay = z1 - z0
    hi = np.arange(n_hex, dtype=np.int32)
    jj, ii = np.meshgrid(hi, hi, indexing="ij")
    base0 = (ii + n1 * jj).ravel().astype(np.int32)
    corner = np.array([dx + n1 * (dy + n1 * dz) for dz in (0, 1) for dy in (0, 1) for dx in (0, 1)], dtype=np.int64)
    off = corner[gen.kuhn_table()].astype(np.int32)
    per_layer = (base0[:, None, None] + off[None, :, :]).reshape(-1, 4)
    ncl = per_layer.shape[0]
    cells = np.empty((nlay * ncl, 4), dtype=np.int32)
    for k in range(nlay):
        np.add(per_layer, np.int32(k * plane), out=cells[k * ncl:(k + 1) * ncl])
    cell_is_own = np.ones(nlay * ncl, dtype=np.uint8)
    if ghost_top:
        cell_is_own[(nlay - 1) * ncl:] = 0
    node_plane = np.repeat(np.arange(z0, z1 + 1, dtype=np.int64), plane)
    # owner = lowest rank whose own cells touch the node: plane k belongs to the slab below when it is a slab boundary
    node_owner = (np.searchsorted(np.asarray(bounds[1:], dtype=np.int64), node_plane, side="left")).clip(0, world - 1).astype(np.int32)
    node_gid = (np.arange(coords.shape[0], dtype=np.int64) + z0 * plane)
    return dict(cells=cells, coords=np.ascontiguousarray(coords), cell_is_own=cell_is_own, node_owner=node_owner, node_gid=node_gid,
                nb_own_cell=int(cell_is_own.sum()), z0=z0, plane=plane, layers=(lay0, lay1), bounds=bounds)


def partition_by_cell_rank(cells, coords, cell_rank, rank, world):
    """Arcane-style local mesh of `rank` for ANY assignment of the cells of a global mesh to ranks (tests: block
    decompositions where a node is shared by three or more ranks).  Own cells = cell_rank == rank; a node is owned by the
    lowest rank among the cells that touch it; the local mesh holds the own cells plus every cell touching a node this
    rank owns (what b2f_ghost_plan requires of the ghost layer).  Local node ids ascend with the global id."""
    cells = np.asarray(cells)
    cell_rank = np.asarray(cell_rank)
    nn = coords.shape[0]
    owner = np.full(nn, world, dtype=np.int64)
    for a in range(cells.shape[1]):
        np.minimum.at(owner, cells[:, a], cell_rank)
    own_cell = cell_rank == rank
    touches_owned = (owner[cells] == rank).any(axis=1)
    keep = own_cell | touches_owned
    lcells_g = cells[keep]
    gid = np.unique(lcells_g)
    lid = np.full(nn, -1, dtype=np.int64)
    lid[gid] = np.arange(gid.shape[0])
    return dict(cells=np.ascontiguousarray(lid[lcells_g].astype(np.int32)), coords=np.ascontiguousarray(coords[gid]),
                cell_is_own=own_cell[keep].astype(np.uint8), node_owner=owner[gid].astype(np.int32), node_gid=gid.astype(np.int64),
                nb_own_cell=int(own_cell.sum()))


def block_cell_rank(n_hex, py, pz, nz_total=None):
    """Rank of every cell of global_mesh(n_hex, ., nz_total=...) for a py x pz block split in y and z (rank = by + py * bz):
    the nodes on the lines where four blocks meet are shared by four ranks."""
    if nz_total is None:
        nz_total = n_hex
    hi = np.arange(n_hex, dtype=np.int64)
    kk, jj, _ = np.meshgrid(np.arange(nz_total, dtype=np.int64), hi, hi, indexing="ij")
    by = (jj * py) // n_hex
    bz = (kk * pz) // nz_total
    return np.repeat((by + py * bz).ravel(), 6)


def exchange_reference(part, ri, co, vals, rank, world, dist):
    """Host mirror of b2f_ghost_exchange on numpy arrays (BS = 1), used by the CPU tests with the gloo backend: every ghost
    row with a non-zero partial sum goes to its owner as (global row, global col, value) triplets; the owner adds them and
    the ghost row is zeroed."""
    import torch
    gid, owner = part["node_gid"], part["node_owner"]
    out = vals.copy()
    send = {p: [] for p in range(world) if p != rank}
    for r in np.nonzero(owner != rank)[0]:
        seg = slice(ri[r], ri[r + 1])
        trip = np.stack([np.full(ri[r + 1] - ri[r], gid[r], dtype=np.float64), gid[co[seg]].astype(np.float64), out[seg]], axis=1)
        send[int(owner[r])].append(trip)
        out[seg] = 0.0
    lid_of = {int(g): i for i, g in enumerate(gid)}
    for p in range(world):
        if p == rank:
            continue
        mine = np.concatenate(send[p], axis=0) if send[p] else np.zeros((0, 3))
        n_mine = torch.tensor([mine.shape[0]], dtype=torch.int64)
        n_theirs = torch.zeros(1, dtype=torch.int64)
        # lower rank sends first: deadlock-free pairwise exchange
        if rank < p:
            dist.send(n_mine, p)
            dist.recv(n_theirs, p)
        else:
            dist.recv(n_theirs, p)
            dist.send(n_mine, p)
        theirs = torch.zeros((int(n_theirs.item()), 3), dtype=torch.float64)
        tm = torch.as_tensor(mine.copy())
        if rank < p:
            if tm.numel():
                dist.send(tm, p)
            if theirs.numel():
                dist.recv(theirs, p)
        else:
            if theirs.numel():
                dist.recv(theirs, p)
            if tm.numel():
                dist.send(tm, p)
        for g_row, g_col, v in theirs.numpy():
            r, c = lid_of[int(g_row)], lid_of[int(g_col)]
            assert owner[r] == rank, "ghost row sent to a rank that does not own it"
            row = co[ri[r]:ri[r + 1]]
            x = int(np.nonzero(row == c)[0][0])
            out[ri[r] + x] += v
    return out


class GhostExchange:
    """NCCL ghost-row exchange bound to one BSRFormat (one rank = one GPU)."""

    def __init__(self, bsr, rank, world, node_owner, node_gid, comm=None):
        """comm: an ncclComm_t (ctypes c_void_p) made by make_comm() to share one communicator between several exchanges;
        None creates one that close() destroys."""
        import torch
        self._lib = _lib.load_library()
        self._bsr = bsr
        self._owns_comm = comm is None
        self._comm = comm if comm is not None else self.make_comm(rank, world)
        owner = np.ascontiguousarray(node_owner, dtype=np.int32)
        gid = np.ascontiguousarray(node_gid, dtype=np.int64)
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        check(self._lib.b2f_ghost_plan(bsr._h, self._comm, int(rank), int(world), owner.ctypes.data_as(C.c_void_p), gid.ctypes.data_as(C.c_void_p), stream))

    @staticmethod
    def make_comm(rank, world):
        """ncclComm_t of libb200fem.so's own NCCL binding; the unique id travels over torch.distributed."""
        import torch
        import torch.distributed as dist
        lib = _lib.load_library()
        comm = C.c_void_p()
        ident = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            buf = (C.c_ubyte * 128)()
            check(lib.b2f_nccl_unique_id(buf))
            ident = torch.tensor(list(buf), dtype=torch.uint8)
        ident = ident.cuda()
        dist.broadcast(ident, 0)
        raw = bytes(ident.cpu().tolist())
        check(lib.b2f_nccl_comm_init(C.c_char_p(raw), int(rank), int(world), C.byref(comm)))
        return comm

    def run(self):
        import torch
        check(self._lib.b2f_ghost_exchange(self._bsr._h, self._comm, C.c_void_p(torch.cuda.current_stream().cuda_stream)))

    def assemble_and_run(self, op):
        """One multi-GPU assembly step with the exchange hidden behind the assembly of the rows that do not travel
        (b2f_assemble_exchange): same values as bsr.assembleBilinear(op) followed by run()."""
        import torch
        d = op.desc(False)
        check(self._lib.b2f_assemble_exchange(self._bsr._h, C.byref(d), self._comm, C.c_void_p(torch.cuda.current_stream().cuda_stream)))

    def assemble_and_run_host(self, op, coords_host, values_host=None):
        """Host-buffer entry point of the multi-GPU step (b2f_assemble_exchange_host): pinned coordinates in, values out,
        pipelined over row chunks with the exchange inside.  Synchronises before returning."""
        import torch
        d = op.desc(False)
        vp = C.c_void_p(values_host.data_ptr()) if values_host is not None else None
        check(self._lib.b2f_assemble_exchange_host(self._bsr._h, C.byref(d), self._comm, C.c_void_p(coords_host.data_ptr()), vp,
                                                   C.c_void_p(torch.cuda.current_stream().cuda_stream)))

    def close(self):
        if self._comm and self._owns_comm:
            self._lib.b2f_nccl_comm_destroy(self._comm)
        self._comm = C.c_void_p()
