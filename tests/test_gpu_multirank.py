"""N ranks on N GPUs (N = 2, 4, 8; skipped when the box has fewer): Arcane-style decomposition, device assembly of the own
cells, NCCL ghost-row exchange inside libb200fem.so; the owner rows of every rank must equal the UNDECOMPOSED oracle
assembly (the reference scheme's result, femutils/BSRFormat.h:807-815) - pattern bit-exact after local -> global mapping
(femutils/HypreDoFLinearSystem.cc:264-303), values 1e-12.  Cases: z-slabs (2 neighbours) with >= 20^3 hexes per rank so the
overlapped path runs, a 2 x 2 block split whose centre line is shared by FOUR ranks (several peers add into one owner row:
fixed summation order), BS = 1 Laplace and BS = 3 elasticity in both value layouts."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def block_rows(vals, ri, bs, per_block):
    """values -> [nnzb, bs, bs] whatever the layout (femutils/BSRFormat.h:145-171)."""
    nnzb = int(ri[-1])
    if per_block or bs == 1:
        return np.asarray(vals).reshape(nnzb, bs, bs)
    out = np.empty((nnzb, bs, bs))
    for r in range(ri.shape[0] - 1):
        rb, k = int(ri[r]), int(ri[r + 1] - ri[r])
        out[rb:rb + k] = np.asarray(vals[rb * bs * bs:(rb + k) * bs * bs]).reshape(bs, k, bs).transpose(1, 0, 2)
    return out


def _worker(rank, world, port, case, q):
    try:
        import torch.distributed as dist
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
        import arcanefem_b200 as fem
        from arcanefem_b200 import multigpu as mg
        from arcanefem_b200.bsr import Mesh
        from oracle import oracle as orc
        from tests.util import op_params
        n_hex, variant, bs, use_csr, decomposition = case["n_hex"], case["variant"], case["bs"], case["use_csr"], case["decomposition"]
        if decomposition == "slab":
            nz_total = case.get("nz_total")
            part = mg.slab_partition(n_hex, rank, world, nz_total=nz_total)
            gcells, gcoords = mg.global_mesh(n_hex, world, nz_total=nz_total)
        else:  # 2 x 2 blocks in y, z: four ranks meet on the centre line
            gcells, gcoords = mg.global_mesh(n_hex, 1, nz_total=n_hex)
            part = mg.partition_by_cell_rank(gcells, gcoords, mg.block_cell_rank(n_hex, 2, world // 2), rank, world)
        cells, coords = part["cells"], part["coords"]
        kind = "laplace" if bs == 1 else "elasticity"
        okw = op_params(kind)
        op = fem.Operator.laplace() if bs == 1 else fem.Operator.elasticity(21.0e5, 0.28)
        bsr = fem.BSRFormat(bs)
        mesh_obj = Mesh.from_numpy(3, cells, coords, cell_is_own=part["cell_is_own"])
        bsr.initialize(mesh_obj, use_csr, assembly_variant=variant)
        bsr.computeSparsity()
        assert bsr.stats()["sparsity_by_lists"] == 1, "the exchange mode builds its pattern from the node->cell lists too"
        ex = mg.GhostExchange(bsr, rank, world, part["node_owner"], part["node_gid"])
        for _ in range(2):  # twice: the exchange must be repeatable (ghost rows re-assembled, then zeroed again)
            bsr.assembleBilinear(op)
            ex.run()
        torch.cuda.synchronize()
        mat = bsr.matrix()
        sequential = mat.values().clone()
        for _ in range(2):  # the overlapped step (ghost rows first, exchange behind the other rows) gives the same bits
            ex.assemble_and_run(op)
        torch.cuda.synchronize()
        if variant == 3:  # the gather kernel and the exchange add in a fixed order: same bits, also with >= 3 ranks on a node
            assert torch.equal(mat.values(), sequential), "b2f_assemble_exchange differs from b2f_assemble + b2f_ghost_exchange"
            bsr.assembleBilinear(op)
            ex.run()
            torch.cuda.synchronize()
            assert torch.equal(mat.values(), sequential), "the exchange is not bitwise repeatable"
        else:
            torch.testing.assert_close(mat.values(), sequential, rtol=1e-12, atol=1e-14)
        if case.get("host_entry"):
            # b2f_assemble_exchange_host: pinned coordinates in, values out, row chunks with the exchange inside (pipelined from
            # 65536 local nodes on): the host copy and the device matrix must equal the device path bit for bit
            coords_host = torch.as_tensor(coords).pin_memory()
            values_host = torch.full((mat.nbNz(),), float("nan"), dtype=torch.float64).pin_memory()
            bsr.setCoords(torch.zeros_like(mesh_obj.coords))  # the entry point must take its coordinates from the host array
            for _ in range(2):
                ex.assemble_and_run_host(op, coords_host, values_host)
            bsr.setCoords(mesh_obj.coords)
            if variant == 3:
                assert torch.equal(values_host, sequential.cpu()), "b2f_assemble_exchange_host: host values differ from the device path"
                assert torch.equal(mat.values(), sequential), "b2f_assemble_exchange_host: device values differ from the device path"
            else:
                torch.testing.assert_close(values_host, sequential.cpu(), rtol=1e-12, atol=1e-14)
            ex.assemble_and_run_host(op, coords_host, None)  # no read-back
            assert torch.equal(mat.values(), sequential) or variant != 3
        ri, co, vals = mat.rowIndexWithEnd().cpu().numpy(), mat.columns().cpu().numpy(), mat.values().cpu().numpy()
        gri, gco = orc.build_pattern(gcells, gcoords.shape[0])
        gvals, _ = orc.assemble(gcells, gcoords, gri, gco, bs=bs, per_block=True, **okw)
        gblk = block_rows(gvals, gri, bs, True)
        blk = block_rows(vals, ri, bs, not use_csr)
        gid = part["node_gid"]
        worst, shared = 0.0, 0
        for r in range(coords.shape[0]):
            seg = slice(ri[r], ri[r + 1])
            if part["node_owner"][r] != rank:
                assert not blk[seg].any(), "ghost rows must end up zero"
                continue
            gseg = slice(gri[gid[r]], gri[gid[r] + 1])
            np.testing.assert_array_equal(gid[co[seg]], gco[gseg])  # same columns, same canonical order, in global numbering
            worst = max(worst, np.abs(blk[seg] - gblk[gseg]).max() / np.abs(gblk[gseg]).max())
        assert worst <= 1e-12, worst
        ex.close()
        q.put((rank, "ok"))
        dist.destroy_process_group()
    except Exception:  # pragma: no cover
        import traceback
        q.put((rank, traceback.format_exc()))


def _run(world, case):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs (gpurun --gpus {world})")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, case, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in res:
        assert msg == "ok", f"rank {rank}: {msg}"


@pytest.mark.parametrize("n_hex", [6, 20])  # 20: the sent rows are a small part of the numbering, the overlapped path runs
@pytest.mark.parametrize("variant", [1, 3])
def test_two_gpu_ghost_exchange_matches_global_assembly(variant, n_hex):
    _run(2, dict(n_hex=n_hex, variant=variant, bs=1, use_csr=True, decomposition="slab"))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_slabs_of_20_layers_per_rank_match_the_undecomposed_oracle(world):
    """>= 20^3 hexes per rank (weak slabs: 20 x 20 x 20*world), default kernel, overlapped exchange."""
    _run(world, dict(n_hex=20, variant=3, bs=1, use_csr=True, decomposition="slab"))


@pytest.mark.parametrize("world", [4, 8])
def test_one_cube_cut_into_uneven_slabs(world):
    """BASELINE configs[4] in small: ONE 22^3 cube cut into `world` slabs of uneven height (22 = 6+6+5+5 / 3 x 6 + 5 x 2...)."""
    _run(world, dict(n_hex=22, nz_total=22, variant=3, bs=1, use_csr=True, decomposition="slab"))


@pytest.mark.parametrize("use_csr", [True, False])
@pytest.mark.parametrize("bs", [1, 3])
@pytest.mark.parametrize("world", [2, 4])
def test_block_sizes_and_layouts_through_the_exchange(world, bs, use_csr):
    """k_send_index / k_recv_index for BS = 3 and both value layouts; world = 4: 2 x 2 blocks, the centre line is shared by
    four ranks, so one owner entry receives from up to three peers."""
    if bs == 1 and not use_csr:
        pytest.skip("BS = 1: both layouts coincide")
    case = dict(n_hex=8, variant=3, bs=bs, use_csr=use_csr, decomposition="slab" if world == 2 else "block")
    _run(world, case)


@pytest.mark.parametrize("world,n_hex,decomposition", [(2, 6, "slab"), (2, 44, "slab"), (4, 64, "block")])
def test_host_entry_point_with_the_exchange_inside(world, n_hex, decomposition):
    """b2f_assemble_exchange_host.  n_hex = 6: the copy / step / copy fallback.  44: 45^3 = 91 k nodes per rank, the chunk
    pipeline; slabs keep the travelling rows in the first chunk and the receiving rows in the last.  64 in 2 x 2 blocks: 70 k
    nodes per rank with ghost and receiving rows spread over the whole numbering (most chunks wait for the exchange)."""
    _run(world, dict(n_hex=n_hex, variant=3, bs=1, use_csr=True, decomposition=decomposition, host_entry=True))
