#!/usr/bin/env python
"""bench.py — BSR assembly throughput (GNNZ/s) + fraction of the HBM roofline (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload W] [--scaling auto|weak|strong]

Workloads (BASELINE.json configs):
  N = 1   poisson3d (default, configs[1]: Poisson 3D P1 tets, Kuhn cube 203^3, 50.2 M tets, BSR 1x1), elasticity3d (configs[2],
          149^3, BS = 3, + Dirichlet row elimination on z = 0), heat3d / heat2d (configs[3], mass + stiffness re-assembly), poisson2d
  N > 1   poisson3d only.  --scaling auto (default) = configs[4] as written: ONE 406^3 cube (401.5 M tets, 1.007e9 nnz) cut
          into N z-slabs (102/102/101/101, 6 x 51 + 2 x 50 layers), "scaling": "strong"; the weak-scaling run (203^3 hexes
          per GPU, the cube grows with N) rides along in "weak_scaling".  --scaling weak / strong runs one mode only.
A step = one values assembly (zero-fill + element matrices + scatter-add [+ ghost-row exchange]) on an existing pattern
(SURVEY.md 8d).  `value`: device-resident inputs, CUDA events on the launching stream, max over ranks.  `e2e`: the host-buffer
entry (pinned coordinates H2D + assembly + values D2H every step).  `parity_sample`: >= 10^4 own rows per rank recomputed on
the host in extended precision from the same coordinates (numpy, independent of oracle/) and compared entry by entry.
`--impl reference`: the CPU restatement of the reference's legacy assembly loop (the reference needs the Arcane framework and
cannot be built here) on all host cores, on a bounded sample of the same workload.
"""
import argparse
import hashlib
import json
import os
import re
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "bsr_assembly_gnnz_per_s"
UNIT = "GNNZ/s"

# name -> (dim, block size, hexes / quads per side at N = 1, operator, BASELINE.json config)
WORKLOADS = {
    "poisson3d": (3, 1, 203, "laplace", "configs[1]"),
    "elasticity3d": (3, 3, 149, "elasticity", "configs[2]"),
    "heat3d": (3, 1, 203, "heat", "configs[3] (3D)"),
    "heat2d": (2, 1, 5000, "heat", "configs[3] (2D)"),
    "poisson2d": (2, 1, 5000, "laplace", "2D twin of configs[1]"),
}
KERNEL = {"poisson3d": "k_node_fan<LaplaceTet>", "elasticity3d": "k_node_fan<ElasticityTet>", "heat3d": "k_node_fan<HeatTet>", "heat2d": "k_node_fan<HeatTri>",
          "poisson2d": "k_node_fan<LaplaceTri>"}
NCU_FILES = {"poisson3d": ["profiles/r02_fan_ncu.txt", "profiles/r01_fan_ncu.txt"], "elasticity3d": ["profiles/r02_fan_elasticity_ncu.txt"],
             "heat3d": ["profiles/r02_fan_heat3d_ncu.txt"], "heat2d": ["profiles/r02_fan_heat2d_ncu.txt"], "poisson2d": ["profiles/r02_fan_poisson2d_ncu.txt"]}


def algorithmic_bytes(nv, nc, nn, nnzb, bs, per_cell_bytes=0):
    """B_asm of SURVEY.md 8(d): connectivity + coordinates + row_index + columns + values (+ per-cell data), each touched once."""
    return 4 * nv * nc + 24 * nn + 4 * (nn + 1) + 4 * nnzb + 8 * bs * bs * nnzb + per_cell_bytes * nc


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum of the workload's kernel from the committed ncu --set full summary
    (per launch), with the file's git blob hash; (None, None) when no capture of the kernel is committed."""
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for rel in NCU_FILES.get(workload, []):
        path = os.path.join(ROOT, rel)
        if not os.path.exists(path):
            continue
        raw = open(path, "rb").read()
        total, found = 0.0, 0
        for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            m = re.search(re.escape(key) + r"\s*=\s*([0-9.eE+-]+)\s*([KMG]?byte)", raw.decode("utf-8", "replace"))
            if m:
                total += float(m.group(1)) * unit[m.group(2)]
                found += 1
        if found == 2:
            blob = hashlib.sha1(b"blob %d\0" % len(raw) + raw).hexdigest()
            return total, f"{rel} (git blob {blob[:12]}): dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full capture of {KERNEL[workload]}"
    return None, None


def bind_to_gpu_numa_node(local_rank):
    """Pins this rank (and the pinned host buffers it allocates afterwards: first touch) to the CPUs of the NUMA node its GPU
    hangs off.  Round 1: all eight ranks sat on node 0 and pulled 1.2 GB each per step through one socket (e2e 94 ms at N = 8).
    Returns a dict for the bench line; never fails the run."""
    info = {"bound": False}
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        bus = bus.lower()
        if len(bus.split(":")[0]) == 8:  # nvml prints an 8-digit domain, sysfs a 4-digit one
            bus = bus[4:]
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read().strip())
        info.update({"pci": bus, "numa_node": node})
        if node < 0:
            return info
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        allowed = cpus & os.sched_getaffinity(0)
        if allowed:
            os.sched_setaffinity(0, allowed)
            info.update({"bound": True, "cpus": len(allowed)})
    except Exception as e:
        info["error"] = f"{type(e).__name__}: {e}"
    return info


def host_threads():
    """Host cores this process may use.  Not omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


class ClockSampler:
    """Samples SM clocks / clock-event reasons while the timed region runs: NVML in-process, one poll per millisecond
    (the timed region is about 10 ms), nvidia-smi at 100 ms as the fallback.  start() returns once sampling runs, so it
    is called BEFORE the barrier that opens the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index=0):
        self.index = index
        self.rows = []
        self.samples = []  # (sm_mhz, reason bit mask)
        self.proc = None
        self.thread = None
        self.nvml = None
        self.handle = None
        self.sm_max = None
        self.running = False
        self.active = False  # only samples taken while `active` count

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
            self.running = True
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _poll(self):
        n = self.nvml
        while self.running:
            if self.active:
                try:
                    self.samples.append((float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)), int(n.nvmlDeviceGetCurrentClocksEventReasons(self.handle))))
                except Exception:
                    pass
            time.sleep(0.001)

    def _read(self):
        for line in self.proc.stdout:
            if self.active:
                self.rows.append(line.strip())

    def stop(self):
        self.active = False
        if self.nvml is not None:
            self.running = False
            self.thread.join(timeout=1)
            n = self.nvml
            bits = {"hw_slowdown": n.nvmlClocksEventReasonHwSlowdown, "hw_thermal_slowdown": n.nvmlClocksEventReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": n.nvmlClocksEventReasonSwThermalSlowdown, "sw_power_cap": n.nvmlClocksEventReasonSwPowerCap}
            reasons = sorted(k for k, b in bits.items() if any(m & b for _, m in self.samples))
            sm = [x for x, _ in self.samples]
            return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": self.sm_max, "reasons": reasons, "samples": len(sm), "source": "nvml, 1 ms polls inside the timed region"}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for name, val in zip(self.NAMES, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi -lms 100"}


def path_breakdown(torch, fem, gen, bsr, op, coords, sparsity_ms, assembly_ms, reps=5):
    """CUDA-event time of every stage of the hot path on the bench workload (ms, mean of `reps` after one warm-up):
    right-hand side, Dirichlet (penalty on the z = 0 plane: flagging + the solve() prelude), CSR view, own-row extraction for
    HYPRE_IJMatrixSetValues, SpMV check, and the matrix + source fused in one walk."""
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed(fn):
        fn()
        torch.cuda.synchronize()
        ev0.record()
        for _ in range(reps):
            fn()
        ev1.record()
        torch.cuda.synchronize()
        return round(ev0.elapsed_time(ev1) / reps, 4)

    ls = fem.DoFLinearSystem()
    bsr.toLinearSystem(ls)
    ls.clearValues()
    group = torch.as_tensor(gen.boundary_nodes(coords, 2, 0.0)).cuda()
    x = torch.ones(coords.shape[0], dtype=torch.float64, device="cuda")
    y = torch.empty_like(x)
    out = {"sparsity": round(sparsity_ms, 3), "assembly": round(assembly_ms, 4)}
    out["assembly_with_fused_source"] = timed(lambda: bsr.assembleBilinearWithSource(op, 5.5))
    out["rhs_constant_source"] = timed(lambda: bsr.applyConstantSourceToRhs(5.5))
    out["to_csr"] = timed(lambda: bsr.matrix().toCsr())
    out["spmv"] = timed(lambda: bsr.spmv(x, y))

    def dirichlet():
        bsr.applyDirichletViaPenalty(group, 0.5, 1.0e30)
        ls.applyDirichlet()

    out["dirichlet_penalty_flag_and_apply"] = timed(dirichlet)
    out["hypre_ij_own_rows"] = timed(lambda: bsr.toHypreIJ())
    out["dirichlet_nodes"] = int(group.numel())
    out["total_once"] = round(out["sparsity"] + out["assembly_with_fused_source"] + out["dirichlet_penalty_flag_and_apply"] + out["to_csr"] + out["hypre_ij_own_rows"], 3)
    return out


def workload_text(workload, mode, n, world, nz_total):
    """The one description of the job both arms print (config.workload)."""
    dim, bs, _, opname, cfg = WORKLOADS[workload]
    if workload != "poisson3d" or world == 1:
        shape = f"{n}^3 hexes" if dim == 3 else f"{n}^2 quads"
        what = {"laplace": "Poisson", "elasticity": "linear elasticity", "heat": "heat (mass + stiffness)"}[opname]
        return f"BASELINE {cfg}: {what} {dim}D P1 {'tets' if dim == 3 else 'triangles'}, structured {'Kuhn cube' if dim == 3 else 'square'} {shape}, BSR {bs}x{bs} values assembly, 1 GPU"
    if mode == "strong":
        return (f"BASELINE configs[4]: Poisson 3D P1 tets, ONE structured Kuhn cube {n}^3 hexes ({6 * n ** 3} cells) cut into {world} z-slabs of "
                f"{'/'.join(str(b - a) for a, b in zip(slab_bounds_of(nz_total, world)[:-1], slab_bounds_of(nz_total, world)[1:]))} hex layers, BSR 1x1 values assembly + ghost-row exchange")
    return f"Poisson 3D P1 tets, structured Kuhn box {n} x {n} x {nz_total} hexes = {n}^3 per GPU on {world} GPUs (weak scaling of BASELINE configs[1]), BSR 1x1 values assembly + ghost-row exchange"


def slab_bounds_of(nz_total, world):
    from arcanefem_b200 import multigpu as mg
    return mg.slab_bounds(nz_total, world)


def laplace_tet_rows_reference(cells, coords, ri, co, rows, extended=True):
    """Rows `rows` of the P1 Laplace matrix from the cells that touch them, in extended precision (numpy longdouble), from the
    reference's formula Ke[a][b] = V grad_a . grad_b with grad = cofactor / (6V) (femutils/ArcaneFemFunctionsGpu.h:110-122,280-392,
    modules/poisson/FemModule.h:177-186).  Independent of oracle/ and of the CUDA formulation.  Returns (ref [n, kmax], mask)."""
    rows = np.asarray(rows, dtype=np.int64)
    real = np.longdouble if extended else np.float64
    nn = coords.shape[0]
    pos = np.full(nn, -1, dtype=np.int32)
    pos[rows] = np.arange(rows.shape[0], dtype=np.int32)
    touch = (pos[cells] >= 0).any(axis=1)
    sel = cells[touch].astype(np.int64)
    X = coords[sel].astype(real)  # [m, 4, 3]
    e1, e2, e3 = X[:, 1] - X[:, 0], X[:, 2] - X[:, 0], X[:, 3] - X[:, 0]
    g = np.empty((sel.shape[0], 4, 3), dtype=real)
    g[:, 1], g[:, 2], g[:, 3] = np.cross(e2, e3), np.cross(e3, e1), np.cross(e1, e2)
    g[:, 0] = -(g[:, 1] + g[:, 2] + g[:, 3])
    det = (e1 * g[:, 1]).sum(axis=1)
    K = np.einsum("mad,mbd->mab", g, g) / (6 * np.abs(det))[:, None, None]
    k = (ri[rows + 1] - ri[rows]).astype(np.int64)
    kmax = int(k.max())
    cols = np.full((rows.shape[0], kmax), -1, dtype=np.int64)
    mask = np.arange(kmax)[None, :] < k[:, None]
    cols[mask] = co[(ri[rows].astype(np.int64)[:, None] + np.arange(kmax)[None, :])[mask]]
    ref = np.zeros((rows.shape[0], kmax), dtype=real)
    missing = 0
    for a in range(4):
        p = pos[sel[:, a]]
        m = p >= 0
        for b in range(4):
            hit = cols[p[m]] == sel[m, b][:, None]
            missing += int((~hit.any(axis=1)).sum())
            slot = hit.argmax(axis=1)
            np.add.at(ref, (p[m], slot), K[m, a, b])
    return ref, mask, missing


def parity_sample(cells, coords, mat, own_rows, receiving_rows, nrows=10000, seed=99):
    """Compares a fixed sample of this rank's own rows (half of them rows that receive ghost contributions, when there are
    any) with the host recomputation.  max_rel = max |d| / max(|ref|, 1e-3 row max) (row-norm floor only for entries that
    cancel to ~0); max_rel_pure = max |d| / |ref| over the entries with |ref| > 1e-6 row max."""
    rng = np.random.default_rng(seed)
    pick = []
    if receiving_rows is not None and receiving_rows.shape[0] > 0:
        pick.append(rng.choice(receiving_rows, size=min(nrows // 2, receiving_rows.shape[0]), replace=False))
    rest = nrows - sum(p.shape[0] for p in pick)
    pick.append(rng.choice(own_rows, size=min(rest, own_rows.shape[0]), replace=False))
    rows = np.unique(np.concatenate(pick))
    ri = mat.rowIndexWithEnd().cpu().numpy()
    co = mat.columns().cpu().numpy()
    ref, mask, missing = laplace_tet_rows_reference(cells, coords, ri, co, rows)
    import torch
    idx = torch.as_tensor((ri[rows][:, None] + np.arange(ref.shape[1])[None, :])[mask]).cuda()
    got = np.zeros(ref.shape, dtype=np.longdouble)
    got[mask] = mat.values()[idx].cpu().numpy()
    d = np.abs(got - ref)
    rowmax = np.abs(ref).max(axis=1, keepdims=True)
    rel = (d / np.maximum(np.abs(ref), 1e-3 * rowmax))[mask]
    big = mask & (np.abs(ref) > 1e-6 * rowmax)
    pure = (d[big] / np.abs(ref[big]))
    # the same formula in plain fp64 on the host: what ANY fp64 evaluation loses on the sample's worst-conditioned cells (a jittered
    # mesh holds a few nearly flat tetrahedra, where the determinant alone costs kappa * eps)
    ref64, _, _ = laplace_tet_rows_reference(cells, coords, ri, co, rows, extended=False)
    rel64 = (np.abs(ref64.astype(np.longdouble) - ref) / np.maximum(np.abs(ref), 1e-3 * rowmax))[mask]
    # device against that plain fp64 host evaluation (the reference computes in fp64 too): the parity figure proper
    relgh = (np.abs(got - ref64.astype(np.longdouble)) / np.maximum(np.abs(ref), 1e-3 * rowmax))[mask]
    # max_rel / over_1e-12: the device against the plain fp64 host evaluation - the reference computes in fp64 too, so this is the parity
    # figure.  The extended-precision comparison rides along: what the device (and any fp64 evaluation) loses on the sample's
    # worst-conditioned cells.
    return {"rows": int(rows.shape[0]), "entries": int(mask.sum()), "max_rel": float(relgh.max()), "over_1e-12": int((relgh > 1e-12).sum()),
            "max_rel_pure": float(pure.max()),
            "vs_extended_precision_max_rel": float(rel.max()), "vs_extended_precision_over_1e-12": int((rel > 1e-12).sum()),
            "host_fp64_vs_extended_precision_max_rel": float(rel64.max()), "host_fp64_vs_extended_precision_over_1e-12": int((rel64 > 1e-12).sum()),
            "columns_missing": missing, "receiving_rows": int(pick[0].shape[0]) if len(pick) > 1 else 0,
            "rule": "|d| / max(|ref|, 1e-3 * row max) over the sampled rows; ref = host recomputation with the reference formula in fp64 (max_rel) and in "
                    "numpy longdouble (vs_extended_precision_*); max_rel_pure = |d| / |ref| against the longdouble values where |ref| > 1e-6 row max"}


def build_workload(args, workload, mode, rank, world):
    """Mesh + operator of one run.  Returns a dict; the mesh arrays are numpy (host) - they stand in for Arcane's views."""
    import torch
    import arcanefem_b200 as fem
    from arcanefem_b200 import mesh as gen
    from arcanefem_b200 import multigpu as mg
    dim, bs, n_default, opname, _ = WORKLOADS[workload]
    keep = []
    if workload == "poisson3d":
        if world == 1:
            n, nz_total = (args.n or n_default), None
        elif mode == "strong":
            n = args.n_global
            nz_total = n
        else:
            n, nz_total = (args.n or n_default), None
        part = mg.slab_partition(n, rank, world, nz_total=nz_total)
        cells, coords = part["cells"], part["coords"]
        nz = nz_total if nz_total is not None else n * world
    else:
        n = args.n or n_default
        cells, coords = gen.cube_kuhn(n) if dim == 3 else gen.square(n)
        part, nz = None, n
    if opname == "laplace":
        op = fem.Operator.laplace()
    elif opname == "elasticity":
        op = fem.Operator.elasticity(21.0e5, 0.28)
    else:
        lam = torch.as_tensor(np.random.default_rng(5).uniform(0.5, 2.0, size=cells.shape[0])).cuda()
        keep.append(lam)
        op = fem.Operator.heat(1.75, 0.1, cell_lambda=lam)
    return dict(dim=dim, bs=bs, n=n, nz_total=nz, cells=cells, coords=coords, part=part, op=op, opname=opname, keep=keep)


def run_mode(args, workload, mode, torch, dist, comm, rank, world, local_rank, full=True):
    """One measured configuration.  full = False (the ride-along mode at N > 1): device-timed value, parity sample and
    sparsity time only."""
    import arcanefem_b200 as fem
    from arcanefem_b200.bsr import Mesh
    from arcanefem_b200 import mesh as gen
    from arcanefem_b200 import multigpu as mg
    t0 = time.time()
    W = build_workload(args, workload, mode, rank, world)
    gen_s = time.time() - t0
    cells, coords, part, op, dim, bs = W["cells"], W["coords"], W["part"], W["op"], W["dim"], W["bs"]
    nc, nn, nv = cells.shape[0], coords.shape[0], cells.shape[1]
    mesh = Mesh.from_numpy(dim, cells, coords, cell_is_own=part["cell_is_own"] if world > 1 else None)
    bsr = fem.BSRFormat(bs)
    bsr.initialize(mesh, True, assembly_variant=args.variant, fan_mode=args.fan_mode, sparsity_build=args.sparsity_build)
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sp = []
    for _ in range(4):  # first build: module load / local-memory set-up inside the timed region; reported: median of the three warm ones
        ev0.record()
        bsr.computeSparsity()
        ev1.record()
        torch.cuda.synchronize()
        sp.append(ev0.elapsed_time(ev1))
    mat = bsr.matrix()
    nnzb, nnz = mat.nbCol(), mat.nbNz()
    exchange = mg.GhostExchange(bsr, rank, world, part["node_owner"], part["node_gid"], comm=comm) if world > 1 else None
    ls = group = None
    if workload == "elasticity3d":
        ls = fem.DoFLinearSystem()
        bsr.toLinearSystem(ls)
        group = torch.as_tensor(gen.boundary_nodes(coords, dim - 1, 0.0)).cuda()

    def step():
        if exchange is None:
            bsr.assembleBilinear(op)
            if ls is not None:  # configs[2]: Dirichlet row elimination on z = 0, every component
                ls.clearValues()
                for comp in range(bs):
                    bsr.applyDirichletViaRowElimination(group, 0.0, dof_index=comp)
                ls.applyDirichlet()
        elif args.no_overlap:
            bsr.assembleBilinear(op)
            exchange.run()
        else:
            exchange.assemble_and_run(op)  # C ABI b2f_assemble_exchange: ghost rows first, exchange behind the other rows

    if exchange is not None:
        for _ in range(10):  # untimed: NCCL channel / proxy set-up for the neighbour pattern
            exchange.run()
    warm = max(args.warmup, 3)
    for _ in range(warm):
        step()
    torch.cuda.synchronize()
    # the sampler is started before the barrier: starting it takes milliseconds on rank 0 alone, and a rank that enters
    # the timed loop late makes its neighbour's first exchange wait inside the neighbour's timed region
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.active = True
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    torch.cuda.synchronize()
    sampler.active = False
    ms_total = ev0.elapsed_time(ev1)
    if dist is not None:
        t = torch.tensor([ms_total], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
        dist.barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms_step = ms_total / args.steps
    stats = bsr.stats()  # error counters of the timed workload itself

    # nonzeros the whole job produced = rows owned by some rank (ghost copies are not counted twice)
    own_nnz, own_cells_job = nnz, nc
    own_rows_np = np.arange(nn, dtype=np.int64)
    receiving = None
    if world > 1:
        own_mask = part["node_owner"] == rank
        own_rows_np = np.nonzero(own_mask)[0]
        t = torch.stack([mat.nbNzPerRow()[torch.as_tensor(own_mask).cuda()].sum().to(torch.float64), torch.tensor(float(part["nb_own_cell"]), device="cuda", dtype=torch.float64)])
        dist.all_reduce(t)
        own_nnz, own_cells_job = int(t[0].item()), int(t[1].item())
        if rank < world - 1:  # the top own plane is touched by the cells of the rank above: its rows receive partial sums
            plane = part["plane"]
            top = part["bounds"][rank + 1] - part["z0"]
            receiving = np.arange(top * plane, (top + 1) * plane, dtype=np.int64)
    out = dict(W=W, ms_step=ms_step, own_nnz=own_nnz, own_cells_job=own_cells_job, nnz=nnz, nnzb=nnzb, nc=nc, nn=nn, nv=nv, sparsity_ms=sorted(sp[1:])[1], sparsity_cold_ms=sp[0], gen_s=gen_s,
               clocks=clocks, stats=stats, value=own_nnz / (ms_step * 1e-3) / 1e9, overlapped=exchange is not None and not args.no_overlap)

    # --- parity of the matrix the timed loop left behind: a fixed sample of own rows against the host recomputation
    out["parity"] = None
    if workload == "poisson3d" and not args.no_parity:
        try:
            ps = parity_sample(cells, coords, mat, own_rows_np, receiving)
            if dist is not None:  # worst rank
                keys = ["max_rel", "max_rel_pure", "over_1e-12", "columns_missing", "vs_extended_precision_max_rel", "vs_extended_precision_over_1e-12",
                        "host_fp64_vs_extended_precision_max_rel", "host_fp64_vs_extended_precision_over_1e-12"]
                t = torch.tensor([float(ps[q]) for q in keys], device="cuda", dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                r = torch.tensor([float(ps["rows"]), float(ps["entries"]), float(ps["receiving_rows"])], device="cuda", dtype=torch.float64)
                dist.all_reduce(r)
                for i, q in enumerate(keys):
                    ps[q] = float(t[i]) if q.endswith("max_rel") or q == "max_rel_pure" else int(t[i])
                ps.update({"rows": int(r[0]), "entries": int(r[1]), "receiving_rows": int(r[2]), "ranks": world})
            out["parity"] = ps
        except Exception as e:
            out["parity"] = {"error": f"{type(e).__name__}: {e}"}
            if dist is not None:  # keep the collectives of the other ranks matched
                raise

    # where a multi-GPU step goes: assembly alone and exchange alone, same protocol (reported, not part of `value`)
    out["breakdown"] = None
    if exchange is not None and full:
        parts = {}
        for name, fn in (("assemble", lambda: bsr.assembleBilinear(op)), ("exchange", exchange.run)):
            torch.cuda.synchronize()
            dist.barrier()
            ev0.record()
            for _ in range(args.steps):
                fn()
            ev1.record()
            torch.cuda.synchronize()
            t = torch.tensor([ev0.elapsed_time(ev1) / args.steps], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            parts[name + "_ms"] = round(float(t.item()), 4)
        out["breakdown"] = parts

    out["e2e"] = out["e2e_resident"] = out["path"] = out["cpu"] = None
    if full:
        # --- end to end through the host-buffer entry point (rank-local; max over ranks)
        coords_host = torch.as_tensor(coords).pin_memory()
        values_host = torch.empty(nnz, dtype=torch.float64).pin_memory()
        e2e_steps = max(2, min(args.steps, 5))
        coords_dev = mesh.coords

        def e2e_step(read_back=True):
            if exchange is None and ls is None:
                bsr.assembleBilinearHost(op, coords_host, values_host if read_back else None)  # C ABI b2f_assemble_host; synchronises before returning
            elif exchange is not None and ls is None and not args.no_overlap:
                exchange.assemble_and_run_host(op, coords_host, values_host if read_back else None)  # C ABI b2f_assemble_exchange_host
            else:
                coords_dev.copy_(coords_host, non_blocking=True)
                step()
                if read_back:
                    values_host.copy_(mat.values(), non_blocking=True)
                torch.cuda.synchronize()

        def timed_host(fn):
            fn()
            torch.cuda.synchronize()
            if dist is not None:
                dist.barrier()
            t0 = time.perf_counter()
            for _ in range(e2e_steps):
                r = fn()
            torch.cuda.synchronize()
            sec = (time.perf_counter() - t0) / e2e_steps
            if dist is not None:
                t = torch.tensor([sec], device="cuda", dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                sec = float(t.item())
            return sec, r

        e2e_s, _ = timed_host(e2e_step)
        checksum = float(values_host.sum().item())
        out["checksum"] = checksum
        out["e2e"] = {"value": round(own_nnz / e2e_s / 1e9, 3), "unit": UNIT, "h2d_bytes_per_step": int(coords_host.numel() * 8), "d2h_bytes_per_step": int(values_host.numel() * 8),
                      "ms_per_step": round(e2e_s * 1e3, 3),
                      "api": "b2f_assemble_host (pinned host coordinates in, values out, pipelined over row chunks)" if exchange is None and ls is None else
                             "b2f_assemble_exchange_host (pinned host coordinates in, values out, row chunks with the ghost-row exchange inside), per rank, max over ranks"
                             if exchange is not None and ls is None and not args.no_overlap else
                             "pinned coordinates H2D + the timed step + values D2H, per rank, max over ranks"}
        # the reference's own data flow keeps the matrix on the device for the Hypre solve (BASELINE.json north_star): same call
        # with the read-back skipped, coordinates in from pinned host memory, one 8-byte checksum of the values out (reported
        # beside `e2e`, never instead of it)
        try:
            def resident_step():
                e2e_step(read_back=False)
                return float(mat.values().sum().item())

            rs_s, rs = timed_host(resident_step)
            out["e2e_resident"] = {"value": round(own_nnz / rs_s / 1e9, 3), "unit": UNIT, "ms_per_step": round(rs_s * 1e3, 3), "h2d_bytes_per_step": int(coords_host.numel() * 8),
                                   "d2h_bytes_per_step": 8, "checksum_matches": bool(abs(rs - checksum) <= 1e-9 * max(1.0, float(values_host.abs().sum().item()))),
                                   "api": "same call with the values read-back skipped + 8-byte checksum read-back: the matrix stays device resident for the solver"}
        except Exception as e:
            out["e2e_resident"] = {"error": f"{type(e).__name__}: {e}"}
            if dist is not None:
                raise
        del coords_host, values_host
        # --- the rest of the path, each stage on its own (t_sparsity, t_dirichlet, t_tocsr, t_total)
        if world == 1 and workload == "poisson3d":
            try:
                out["path"] = path_breakdown(torch, fem, gen, bsr, op, coords, sorted(sp[1:])[1], ms_step)
            except Exception as e:  # a reporting extra must never cost the bench line
                out["path"] = {"error": f"{type(e).__name__}: {e}"}
        if rank == 0 and world == 1:
            out["cpu"] = cpu_baseline(workload, W, mat.rowIndexWithEnd().cpu().numpy(), mat.columns().cpu().numpy(), args.cpu_cells)
    if exchange is not None:
        exchange.close()
    del bsr, mesh, mat
    torch.cuda.empty_cache()
    return out


def run_b200(args):
    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dist = comm = None
    workload = args.workload
    all_cpus = host_threads()
    affinity = bind_to_gpu_numa_node(local_rank) if world > 1 and not args.no_numa_bind else {"bound": False}
    affinity["host_cores_before_binding"] = all_cpus
    if world > 1:
        import torch.distributed as dist
        from arcanefem_b200 import multigpu as mg
        if workload != "poisson3d":
            raise SystemExit("multi-GPU runs are defined for --workload poisson3d (BASELINE configs[4])")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        comm = mg.GhostExchange.make_comm(rank, world)
    mode = "weak" if world == 1 else ("strong" if args.scaling in ("auto", "strong") else "weak")
    R = run_mode(args, workload, mode, torch, dist, comm, rank, world, local_rank, full=True)
    other = None
    if world > 1 and args.scaling == "auto":
        other = run_mode(args, workload, "weak", torch, dist, comm, rank, world, local_rank, full=False)
    if rank == 0:
        line = make_line(args, workload, mode, world, R, other)
        line["host_affinity_rank0"] = affinity
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def mode_summary(R, world):
    return {"value": round(R["value"], 3), "unit": UNIT, "ms_per_step": round(R["ms_step"], 4), "gnnz_per_s_per_gpu": round(R["value"] / world, 3), "nnz_job": R["own_nnz"],
            "cells_job": R["own_cells_job"], "cells_on_rank0": R["nc"], "sparsity_build_ms": round(R["sparsity_ms"], 3), "parity_sample": R["parity"],
            "columns_not_found": R["stats"]["columns_not_found"], "gather_fallback_tiles": R["stats"]["gather_fallback_tiles"]}


def make_line(args, workload, mode, world, R, other):
    W = R["W"]
    dim, bs, _, opname, _ = WORKLOADS[workload]
    peak, peak_src = measured_peak()
    per_cell = 8 if opname == "heat" else 0
    b_alg = algorithmic_bytes(R["nv"], R["nc"], R["nn"], R["nnzb"], bs, per_cell)  # rank 0's launch (the largest slab)
    achieved = b_alg / (R["ms_step"] * 1e-3) / 1e9
    variant_names = {0: "node_gather_fan", 1: "cell_atomic", 2: "cell_atomic_warpagg", 3: "node_gather_fan", 4: "node_gather_slots"}
    default_kernel = args.variant in (0, 3)
    kernel = KERNEL[workload] if default_kernel else {1: "memset + k_cell_atomic", 2: "memset + k_cell_atomic<warpagg>", 4: "k_node_gather_slots"}.get(args.variant, "?")
    traffic, traffic_src = ncu_traffic(workload) if (default_kernel and world == 1 and not args.n) else (None, None)
    # our kernels per step: the assembly kernel (+ pack and add kernels of the ghost exchange; NCCL's own kernels not counted;
    # overlapped step: the travelling rows, the rows before them and the rows after them are separate launches of the kernel)
    if world > 1:
        launches = (3 if R["overlapped"] else 1) + 2
    else:
        launches = 6 if workload == "elasticity3d" else 1  # + 3 flag kernels + k_any_rowcol + k_dirichlet_rows
    stats = R["stats"]
    line = {
        "metric": METRIC, "value": round(R["value"], 3), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": round(R["ms_step"], 4), "higher_is_better": True, "scaling": mode, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_text(workload, mode, W["n"], world, W["nz_total"]), "name": workload, "n_hex": W["n"], "nz_hex_total": W["nz_total"],
                   "variant": variant_names.get(args.variant, str(args.variant)), "cells_on_rank0": R["nc"], "nnz_on_rank0": R["nnz"], "nnz_job": R["own_nnz"], "cells_job": R["own_cells_job"],
                   "jitter": 0.25, "cache": "inputs larger than L2 (GBs touched per step vs 126 MB L2)",
                   "parallelism": "1 rank per GPU, z-slab decomposition with a ghost-cell layer (Arcane style), own cells assembled once, NCCL ghost-row exchange" if world > 1 else "single GPU",
                   "sparsity_build": "row-wise from node->cell lists" if stats["sparsity_by_lists"] else "global edge sort", "sparsity_build_ms": round(R["sparsity_ms"], 3),
                   "sparsity_build_first_call_ms": round(R["sparsity_cold_ms"], 3), "mesh_gen_s": round(R["gen_s"], 2), "values_checksum": R.get("checksum"),
                   "columns_not_found": stats["columns_not_found"], "gather_fallback_tiles": stats["gather_fallback_tiles"],
                   "step_breakdown": R["breakdown"], "exchange_overlapped": R["overlapped"], "path_breakdown_ms": R["path"],
                   "step": "values assembly + Dirichlet row elimination on z = 0 (3 components)" if workload == "elasticity3d" else "values assembly" + (" + ghost-row exchange" if world > 1 else "")},
        "parity_sample": R["parity"],
        "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4), "traffic": traffic,
                     "peak_source": peak_src, "algorithmic_bytes_per_launch": b_alg, "frac_of_nominal_8TBs": round(achieved / 8000.0, 4), "kernel": kernel, "traffic_source": traffic_src,
                     "note": None if world == 1 else "rank 0's slab; the step time includes the exchange"},
        "cpu_baseline": R["cpu"],
        "e2e": R["e2e"],
        "e2e_matrix_resident": R["e2e_resident"],
        "gpu_launches": launches * args.steps,
        "clocks": R["clocks"],
    }
    if world > 1:
        line["per_gpu"] = {"gnnz_per_s": round(R["value"] / world, 3), "cells": R["own_cells_job"] // world}
        line["cpu_baseline"] = {"value": None, "note": "timed by the reference arm (bench.py --impl reference) and at N = 1; not repeated on every rank"}
        if other is not None:
            line["weak_scaling"] = dict(mode_summary(other, world), workload=workload_text(workload, "weak", other["W"]["n"], world, other["W"]["nz_total"]), scaling="weak")
    return line


def cpu_seconds(orc, workload, cells, coords, ri, co, threads):
    """One pass of the CPU port over `cells` (seconds).  Laplace / heat: the reference's legacy loop (modules/poisson/FemModule.cc:300-323
    + CSR add femutils/CsrFormatMatrix.h:58-87), OpenMP over cells + atomics (Arcane -A,T=n analogue).  Elasticity BS = 3 has no
    legacy twin in the reference: the oracle's sequential cell-wise BSR assembly (femutils/BSRFormat.h:786-837 semantics)."""
    dim, bs, _, opname, _ = WORKLOADS[workload]
    if opname == "elasticity":
        E, nu = 21.0e5, 0.28
        t0 = time.perf_counter()
        orc.assemble(cells, coords, ri, co, bs=bs, per_block=False, op=orc.OP_ELASTICITY, lambda_=E * nu / ((1 + nu) * (1 - 2 * nu)), mu2=E / (1 + nu))
        return time.perf_counter() - t0, 1
    vals = np.zeros(co.shape[0])
    kw = dict(op=orc.OP_HEAT, lambda_=1.75, inv_dt=10.0) if opname == "heat" else {}
    return orc.legacy_assemble_csr(cells, coords, ri, co, vals, threads=threads, **kw), threads


def best_cpu_threads(orc, workload, cells, coords, ri, co, max_threads):
    """The threaded loop uses fp64 atomics like Arcane's thread runner; on some hosts fewer threads are faster: probe."""
    if WORKLOADS[workload][3] == "elasticity":
        return 1
    probe = cells[:min(cells.shape[0], 400000)]
    best_t, best_rate = 1, 0.0
    for t in sorted({1, max(1, max_threads // 2), max_threads}):
        s, _ = cpu_seconds(orc, workload, probe, coords, ri, co, t)
        if probe.shape[0] / s > best_rate:
            best_t, best_rate = t, probe.shape[0] / s
    return best_t


def cpu_baseline(workload, W, ri, co, max_cells):
    """Oracle port of the reference's CPU assembly on the host cores, bounded sample of the same mesh (variant B =
    all threads is `value`; variant A = one thread and variant C = the sparsity build ride along)."""
    from oracle import oracle as orc
    cells, coords = W["cells"], W["coords"]
    bs = W["bs"]
    nc = cells.shape[0]
    sample = min(nc, max_cells if WORKLOADS[workload][3] != "elasticity" else min(max_cells, 2_000_000))
    threads = best_cpu_threads(orc, workload, cells[:sample], coords, ri, co, host_threads())
    secs, threads = cpu_seconds(orc, workload, cells[:sample], coords, ri, co, threads)
    nnz = co.shape[0] * bs * bs
    out = {"value": round(nnz * (sample / nc) / secs / 1e9, 4), "unit": UNIT, "cores": threads, "kind": "port", "host_cores": host_threads(),
           "sample": f"first {sample} of {nc} cells of the same mesh, one pass, {secs:.2f} s; nnz credited pro rata"}
    try:
        if workload == "poisson3d":
            one = min(sample, 5_000_000)
            s1, _ = cpu_seconds(orc, workload, cells[:one], coords, ri, co, 1)
            out["single_thread"] = {"value": round(nnz * (one / nc) / s1 / 1e9, 4), "unit": UNIT, "sample": f"first {one} cells, {s1:.2f} s"}
            from arcanefem_b200 import mesh as gen
            c_sp, x_sp = gen.cube_kuhn(64)
            t0 = time.perf_counter()
            _, co_sp = orc.build_pattern(c_sp, x_sp.shape[0])
            s_sp = time.perf_counter() - t0
            out["sparsity_build"] = {"value": round(co_sp.shape[0] / s_sp / 1e9, 4), "unit": "GNNZ/s of pattern", "cores": 1,
                                     "sample": f"Kuhn cube 64^3 ({c_sp.shape[0]} cells, {co_sp.shape[0]} blocks), {s_sp:.2f} s"}
    except Exception as e:  # extras only
        out["extras_error"] = f"{type(e).__name__}: {e}"
    return out


def run_reference(args):
    """CPU restatement of the reference's own CPU assembly, all host threads, on a bounded sample of the b200 arm's job:
    the first hex layers (poisson3d) / the first cells (other workloads) of the same mesh.  Rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    from arcanefem_b200 import mesh as gen
    from arcanefem_b200 import multigpu as mg
    from oracle import oracle as orc
    workload = args.workload
    dim, bs, n_default, opname, _ = WORKLOADS[workload]
    mode = "weak" if world == 1 else ("strong" if args.scaling in ("auto", "strong") else "weak")
    if workload == "poisson3d":
        if world > 1 and mode == "strong":
            n = nz_total = args.n_global
        else:
            n, nz_total = (args.n or n_default), (args.n or n_default) * world
        layers = max(1, min(nz_total, args.cpu_cells // (6 * n * n)))
        part = mg.slab_partition(n, 0, 1, nz_total=nz_total, max_layers=layers)  # the first `layers` hex layers of the job's mesh
        cells, coords = part["cells"], part["coords"]
        job_cells = 6 * n * n * nz_total
        nnz_job = None
        sample_txt = f"the first {layers} of {nz_total} hex layers of the job's mesh ({cells.shape[0]} of {job_cells} cells) per step"
    else:
        n = nz_total = args.n or n_default
        cells, coords = gen.cube_kuhn(n) if dim == 3 else gen.square(n)
        job_cells = cells.shape[0]
        keep = min(job_cells, args.cpu_cells if opname != "elasticity" else min(args.cpu_cells, 2_000_000))
        sample_txt = f"the first {keep} of {job_cells} cells of the job's mesh per step"
        # a prefix of the cells only touches a prefix of the nodes (lexicographic numbering): pattern over those
        cells = cells[:keep]
        nn_used = int(cells.max()) + 1
        coords = coords[:nn_used]
    nn = coords.shape[0]
    ri, co = orc.build_pattern(cells, nn)  # untimed: the metric is the values assembly on an existing pattern
    nnz_sample = co.shape[0] * bs * bs
    threads = best_cpu_threads(orc, workload, cells, coords, ri, co, host_threads())
    times = []
    warm = max(args.warmup, 1)
    for it in range(warm + args.steps):
        s, used = cpu_seconds(orc, workload, cells, coords, ri, co, threads)
        if it >= warm:
            times.append(s)
    sec = float(np.mean(times))
    value = nnz_sample / sec / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": round(value, 4), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
        "ms_per_step": round(sec * 1e3, 3), "higher_is_better": True, "scaling": mode, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_text(workload, mode, n, world, nz_total), "name": workload, "n_hex": n, "nz_hex_total": nz_total, "cells_job": job_cells,
                   "nnz_of_sample": nnz_sample, "cells_of_sample": int(cells.shape[0]),
                   "note": "CPU restatement of modules/poisson/FemModule.cc:300-323 (the reference needs Arcane and cannot be built here); value = nonzeros of the "
                           "sample's matrix / time of one pass over the sample: a throughput, so it compares with the b200 arm's whole-job GNNZ/s"},
        "cpu_baseline": {"value": round(value, 4), "unit": UNIT, "cores": used, "kind": "port", "host_cores": host_threads(), "sample": sample_txt + f", {used} threads"},
        "e2e": {"value": round(value, 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="poisson3d", choices=sorted(WORKLOADS))
    ap.add_argument("--scaling", default="auto", choices=["auto", "weak", "strong"], help="N > 1: auto = strong (configs[4]) with the weak run nested in the line")
    ap.add_argument("--n", type=int, default=0, help="hexes per side (per GPU in the weak mode); 0 = the workload's BASELINE size (203 -> 50.2M tets)")
    ap.add_argument("--n-global", type=int, default=406, help="hexes per side of the one cube of the strong mode (406 -> 401.5M tets)")
    ap.add_argument("--variant", type=int, default=3, help="1 cell atomic, 2 cell atomic warp-aggregated, 3 node gather")
    ap.add_argument("--fan-mode", type=int, default=0, help="node gather walk on tets: 0 chosen from the plan, 1 select walk, 2 lockstep walk")
    ap.add_argument("--no-overlap", action="store_true", help="N > 1: b2f_assemble then b2f_ghost_exchange in sequence instead of b2f_assemble_exchange")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity sample")
    ap.add_argument("--no-numa-bind", action="store_true", help="N > 1: do not pin the ranks to the NUMA node of their GPU")
    ap.add_argument("--sparsity-build", type=int, default=0, help="0 default (row-wise from the node->cell lists), 1 global edge sort (the reference's algorithm)")
    ap.add_argument("--cpu-cells", type=int, default=60_000_000, help="cells of the bounded CPU sample (the N = 1 mesh has 50.2 M: one whole pass)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
