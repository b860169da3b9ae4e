"""The range-checked build of the library (tools/build_variant.sh checked -DB2F_CHECKED=1; stands in for compute-sanitizer, which is
closed on the measurement pool): every shared-memory access of the fan walk is checked against the thread's own ranges and every
gathered node id against [0, nb_node), violations are counted in b2f_stats.columns_not_found.  These tests only run when
B2F_LIBRARY points at that build (tools/gpu_checked.sh); the rest of the GPU suite, run against it, asserts the counter to be zero."""
import os

import numpy as np
import pytest

from tests.util import MESHES

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

CHECKED = os.environ.get("B2F_LIBRARY", "").endswith("_checked.so")


@pytest.mark.skipif(not CHECKED, reason="needs the B2F_CHECKED build (B2F_LIBRARY=.../libb200fem_checked.so)")
@pytest.mark.parametrize("mesh_name", ["cube5", "square9_shuffled"])
def test_checked_build_counts_a_corrupted_step_word(mesh_name):
    import arcanefem_b200 as fem
    from arcanefem_b200.bsr import Mesh
    dim, cells, coords = MESHES[mesh_name]()
    bsr = fem.BSRFormat(1)
    bsr.initialize(Mesh.from_numpy(dim, cells, coords), True)
    bsr.computeSparsity()
    op = fem.Operator.laplace()
    bsr.assembleBilinear(op)
    torch.cuda.synchronize()
    assert bsr.stats()["columns_not_found"] == 0
    good = bsr.matrix().values().clone()
    plan = bsr.plan()
    steps = plan["fan_steps"]
    fan_ptr = plan["fan_ptr"].cpu().numpy()
    row = int(np.argmax(np.diff(fan_ptr)))  # a row with the longest walk
    q = int(fan_ptr[row]) + 2  # an emit word inside its walk
    saved = steps[q].clone()
    w = int(saved.item()) & 0xFFFFFFFF
    assert not w & 0x80000000
    bad = (w & ~1023) | 1000  # slot 1000: outside the row's columns and accumulators
    steps[q] = torch.tensor(bad - (1 << 32) if bad >= (1 << 31) else bad, dtype=steps.dtype, device=steps.device)
    bsr.assembleBilinear(op)
    torch.cuda.synchronize()
    assert bsr.stats()["columns_not_found"] > 0, "the checked build did not notice the out-of-range slot"
    steps[q] = saved
    bsr.assembleBilinear(op)
    torch.cuda.synchronize()
    assert bsr.stats()["columns_not_found"] == 0
    assert torch.equal(bsr.matrix().values(), good)
