#!/usr/bin/env python
"""Statistics of the fan-walk plan on the headline mesh: how many steps change the pivot, how many
warp steps carry the pivot hint, START words per row, steps per row.  One JSON line."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import arcanefem_b200 as fem
    from arcanefem_b200 import mesh as gen
    from arcanefem_b200.bsr import Mesh
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 203
    cells, coords = gen.cube_kuhn(n)
    bsr = fem.BSRFormat(1)
    bsr.initialize(Mesh.from_numpy(3, cells, coords), True)
    bsr.computeSparsity()
    plan = bsr.plan()
    w = plan["fan_steps"]
    ptr = plan["fan_ptr"].to(torch.int64)
    start = w < 0  # bit 31
    emit = ~start
    pivot = emit & ((w >> 10) & 1 == 0)
    hint = emit & ((w >> 12) & 1 == 1)
    nstep = (ptr[1:] - ptr[:-1])
    # warp steps: rows are walked 32 at a time, a warp runs max(steps) of its rows
    nn = nstep.shape[0]
    pad = (-nn) % 32
    per_warp = torch.nn.functional.pad(nstep, (0, pad)).reshape(-1, 32)
    out = {"n": n, "rows": int(nn), "steps": int(w.numel()), "start_words": int(start.sum()), "emit_steps": int(emit.sum()),
           "pivot_changes": int(pivot.sum()), "pivot_change_fraction_of_emit_steps": round(float(pivot.sum()) / float(emit.sum()), 4),
           "steps_with_pivot_hint": int(hint.sum()), "hint_fraction_of_emit_steps": round(float(hint.sum()) / float(emit.sum()), 4),
           "steps_per_row_mean": round(float(nstep.double().mean()), 3), "steps_per_row_max": int(nstep.max()),
           "warp_steps": int(per_warp.max(dim=1).values.sum()), "lane_steps_idle_fraction": round(1.0 - float(nstep.sum()) / float(per_warp.max(dim=1).values.sum() * 32), 4)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
