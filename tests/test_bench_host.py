"""Host-side pieces of bench.py that need no GPU: the extended-precision row recomputation behind `parity_sample`, the
reference arm (CPU port) at N = 1 and under a torchrun-like environment, and the host thread count."""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_row_recomputation_agrees_with_the_oracle():
    sys.path.insert(0, ROOT)
    import bench
    from arcanefem_b200 import mesh as gen
    from oracle import oracle as orc
    cells, coords = gen.cube_kuhn(9)
    ri, co = orc.build_pattern(cells, coords.shape[0])
    vals, _ = orc.assemble(cells, coords, ri, co)
    rows = np.unique(np.random.default_rng(0).choice(coords.shape[0], 300, replace=False))
    ref, mask, missing = bench.laplace_tet_rows_reference(cells, coords, ri, co, rows)
    assert missing == 0
    got = np.zeros(ref.shape)
    for i, r in enumerate(rows):
        got[i, :ri[r + 1] - ri[r]] = vals[ri[r]:ri[r + 1]]
    rowmax = np.abs(ref).max(axis=1, keepdims=True)
    rel = (np.abs(got - ref) / np.maximum(np.abs(ref), 1e-3 * rowmax))[mask]
    assert float(rel.max()) <= 1e-12, float(rel.max())


def _reference_line(extra, env=None):
    e = dict(os.environ)
    e.update(env or {})
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1"] + extra, capture_output=True, text=True,
                         env=e, timeout=600, check=True).stdout.strip().splitlines()
    return json.loads(out[-1])


def test_reference_arm_single_rank():
    d = _reference_line(["--n", "16"])
    assert d["impl"] == "reference" and d["metric"] == "bsr_assembly_gnnz_per_s" and d["value"] > 0 and d["scaling"] == "weak"
    assert d["cpu_baseline"]["kind"] == "port" and d["e2e"]["h2d_bytes_per_step"] == 0
    assert d["config"]["workload"].startswith("BASELINE configs[1]")


def test_reference_arm_under_torchrun_uses_the_host_cores_and_the_jobs_mesh():
    """torchrun exports OMP_NUM_THREADS=1: the arm must still see the host cores, and at N > 1 describe (and sample) the
    N-rank job - the 1-thread / single-GPU-mesh defect of round 1."""
    env = {"OMP_NUM_THREADS": "1", "WORLD_SIZE": "4", "RANK": "0", "LOCAL_RANK": "0"}
    d = _reference_line(["--gpus", "4", "--n-global", "20", "--cpu-cells", "30000"], env)
    assert d["n_gpus"] == 4 and d["scaling"] == "strong"
    assert "cut into 4 z-slabs of 5/5/5/5" in d["config"]["workload"]
    assert d["cpu_baseline"]["host_cores"] == len(os.sched_getaffinity(0))
    assert d["config"]["cells_job"] == 6 * 20 ** 3 and d["config"]["cells_of_sample"] < d["config"]["cells_job"]
    # the other ranks exit without printing
    env["RANK"] = "1"
    e = dict(os.environ)
    e.update(env)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "4"], capture_output=True, text=True, env=e, timeout=120, check=True)
    assert out.stdout.strip() == ""


def test_workload_strings_are_shared_by_both_arms():
    sys.path.insert(0, ROOT)
    import bench
    a = bench.workload_text("poisson3d", "strong", 406, 8, 406)
    assert "51/51/51/51/51/51/50/50" in a and "401540496 cells" in a
    assert "102/102/101/101" in bench.workload_text("poisson3d", "strong", 406, 4, 406)
    t, src = bench.ncu_traffic("poisson3d")
    assert t is None or (2.0e9 < t < 4.0e9 and "git blob" in src)
