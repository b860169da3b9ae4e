"""Host logic of the multi-GPU path on CPU: 2 ranks over gloo.  Each rank takes its slab (own cells + ghost-cell layer),
assembles ONLY its own cells with the oracle, exchanges ghost rows (host mirror of b2f_ghost_exchange) and must end up
with exactly the owner rows of the undecomposed assembly — the reference scheme's result (femutils/BSRFormat.h:807-815)."""
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_hex, q, decomposition="slab", nz_total=None):
    try:
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        from arcanefem_b200 import multigpu as mg
        from oracle import oracle as orc
        if decomposition == "slab":
            part = mg.slab_partition(n_hex, rank, world, nz_total=nz_total)
            gcells, gcoords = mg.global_mesh(n_hex, world, nz_total=nz_total)
        else:  # 2 x 2 blocks in y and z: the centre line is shared by four ranks
            gcells, gcoords = mg.global_mesh(n_hex, 1, nz_total=n_hex)
            part = mg.partition_by_cell_rank(gcells, gcoords, mg.block_cell_rank(n_hex, 2, world // 2), rank, world)
        cells, coords = part["cells"], part["coords"]
        nn = coords.shape[0]
        ri, co = orc.build_pattern(cells, nn)  # pattern from own + ghost cells
        own_cells = np.ascontiguousarray(cells[part["cell_is_own"] != 0])
        vals, nf = orc.assemble(own_cells, coords, ri, co)  # values from own cells only, all local rows
        assert nf == 0
        vals = mg.exchange_reference(part, ri, co, vals, rank, world, dist)
        # undecomposed assembly
        gri, gco = orc.build_pattern(gcells, gcoords.shape[0])
        gvals, _ = orc.assemble(gcells, gcoords, gri, gco)
        gid = part["node_gid"]
        np.testing.assert_array_equal(gcoords[gid], coords)
        worst = 0.0
        for r in range(nn):
            seg = slice(ri[r], ri[r + 1])
            if part["node_owner"][r] != rank:
                assert not vals[seg].any(), "ghost rows must end up zero"
                continue
            g = gid[r]
            gseg = slice(gri[g], gri[g + 1])
            np.testing.assert_array_equal(gid[co[seg]], gco[gseg])  # same columns in the same (canonical) order
            scale = np.abs(gvals[gseg]).max()
            worst = max(worst, np.abs(vals[seg] - gvals[gseg]).max() / scale)
        assert worst <= 1e-12, worst
        # every global node is owned by exactly one rank
        owned = torch.zeros(gcoords.shape[0], dtype=torch.int64)
        owned[torch.as_tensor(gid[part["node_owner"] == rank])] = 1
        dist.all_reduce(owned)
        assert int(owned.min()) == 1 and int(owned.max()) == 1
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_slab_decomposition_and_ghost_exchange_reproduce_global_assembly(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, 4, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in res:
        assert msg == "ok", f"rank {rank}: {msg}"


def _spawn(world, n_hex, **kw):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_hex, q), kwargs=kw) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in res:
        assert msg == "ok", f"rank {rank}: {msg}"


def test_one_cube_cut_into_uneven_slabs():
    """BASELINE configs[4] in small: one 7^3 cube, 3 slabs of 3/2/2 layers."""
    _spawn(3, 7, nz_total=7)


def test_block_decomposition_with_a_node_shared_by_four_ranks():
    _spawn(4, 4, decomposition="block")
