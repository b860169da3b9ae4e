#!/usr/bin/env python
"""Summarise an .ncu-rep (raw page + source page) into text: key metrics, opcode mix, hottest SASS lines."""
import csv, subprocess, sys, io
from collections import Counter
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, vals = rows[0], rows[-1]
want = ['Kernel Name','gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','sm__throughput.avg.pct_of_peak_sustained_elapsed',
 'sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','launch__occupancy_limit_shared_mem','launch__occupancy_limit_registers','launch__grid_size','launch__block_size',
 'smsp__inst_executed.sum','sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct_of_peak_sustained_active','l1tex__t_sector_hit_rate.pct','lts__t_sector_hit_rate.pct',
 'lts__t_sectors_op_red.sum','lts__t_sectors_op_atom.sum','l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum','l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum','sm__cycles_elapsed.max','smsp__thread_inst_executed_per_inst_executed.ratio',
 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','lts__t_bytes.sum','l1tex__t_bytes.sum','sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active']
for i, h in enumerate(hdr):
    if h in want:
        print(f"{h} = {vals[i]} {rows[1][i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr, data = rows[hi], rows[hi + 1:]
ix = {h: i for i, h in enumerate(hdr)}
f = lambda r, k: float(r[ix[k]] or 0)
ti = sum(f(r, 'Instructions Executed') for r in data); ts = sum(f(r, '# Samples') for r in data)
print(f"total warp-inst {ti:.0f}, samples {ts:.0f}, sass lines {len(data)}")
c, s = Counter(), Counter()
for r in data:
    t = r[ix['Source']].split()
    if not t: continue
    op = t[1] if t[0].startswith('@') else t[0]
    op = op.split('.')[0]
    c[op] += f(r, 'Instructions Executed'); s[op] += f(r, '# Samples')
print("opcode mix (inst%, samples%):")
for op, v in c.most_common(22):
    print(f"  {op:10s} {v/ti*100:5.1f} {s[op]/ts*100:5.1f}")
stalls = [k for k in hdr if k.startswith('stall_') and '(' not in k]
tot = Counter()
for r in data:
    for k in stalls: tot[k] += f(r, k)
print("stall reasons (all samples):", ", ".join(f"{k[6:]} {v/ts*100:.1f}%" for k, v in tot.most_common(8)))
print("hottest SASS:")
for r in sorted(data, key=lambda r: -f(r, '# Samples'))[:int(sys.argv[2]) if len(sys.argv) > 2 else 25]:
    st = {k: f(r, k) for k in stalls}
    print(f"  {f(r,'# Samples')/ts*100:5.1f}% inst {f(r,'Instructions Executed')/ti*100:4.1f}% {max(st, key=st.get)[6:]:14s} {r[ix['Source']][:100]}")
