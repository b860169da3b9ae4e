#!/usr/bin/env python
"""Where a kernel's instructions and stall samples sit, by section of its SASS: tools/ncu_sections.py report.ncu-rep
Reads the source page of an ncu report (ncu --set full --import-source on), finds the loops (backward branches) and prints, for
every loop and for the stretches between the outermost ones, the share of executed warp instructions and of sampled warp states,
then the opcode mix and the stall reasons of the hottest loop and its top sample lines."""
import csv
import io
import re
import subprocess
import sys
from collections import Counter


def main():
    rep = sys.argv[1]
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr, data = rows[hi], rows[hi + 1:]
    ix = {h: i for i, h in enumerate(hdr)}
    f = lambda r, k: float(r[ix[k]] or 0)
    tot = sum(f(r, "Instructions Executed") for r in data)
    ts = sum(f(r, "# Samples") for r in data)
    addr = [int(r[ix["Address"]], 16) if r[ix["Address"]].startswith("0x") else int(r[ix["Address"]]) for r in data]
    pos = {a: i for i, a in enumerate(addr)}
    loops = []
    for i, r in enumerate(data):
        m = re.search(r"BRA(?:\.\w+)*\s+(?:\w+,\s*)?0x([0-9a-f]+)", r[ix["Source"]])
        if m:
            j = pos.get(int(m.group(1), 16))
            if j is not None and j <= i:
                loops.append((j, i))
    share = lambda a, b: (sum(f(r, "Instructions Executed") for r in data[a:b + 1]) / tot * 100, sum(f(r, "# Samples") for r in data[a:b + 1]) / ts * 100)
    print(f"{len(data)} SASS lines, {tot:.0f} warp instructions, {ts:.0f} samples")
    print("loops (first line, last line, instructions, inst %, samples %):")
    for a, b in sorted(loops):
        i_, s_ = share(a, b)
        if i_ >= 0.3 or s_ >= 0.3:
            print(f"  {a:5d} {b:5d} {b - a + 1:5d}  {i_:5.1f}  {s_:5.1f}")
    hot = max(loops, key=lambda ab: share(*ab)[0] if ab[1] - ab[0] < 600 else 0)
    a, b = hot
    W = data[a:b + 1]
    wi = sum(f(r, "Instructions Executed") for r in W)
    ws = sum(f(r, "# Samples") for r in W)
    print(f"hottest inner loop: lines {a}-{b}: {wi / tot * 100:.1f} % of the instructions, {ws / ts * 100:.1f} % of the samples")
    c, s = Counter(), Counter()
    for r in W:
        t = r[ix["Source"]].split()
        if not t:
            continue
        op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
        c[op] += f(r, "Instructions Executed")
        s[op] += f(r, "# Samples")
    print("  opcode mix (inst % of the loop, samples % of the loop):")
    for op, v in c.most_common(16):
        print(f"    {op:8s} {v / wi * 100:5.1f} {s[op] / ws * 100:5.1f}")
    stalls = [k for k in hdr if k.startswith("stall_") and "(" not in k]
    st = Counter()
    for r in W:
        for k in stalls:
            st[k] += f(r, k)
    print("  stall reasons of the loop:", ", ".join(f"{k[6:]} {v / ws * 100:.1f}%" for k, v in st.most_common(8)))
    print("top sample lines of the kernel:")
    for smp, i in sorted(((f(r, "# Samples"), i) for i, r in enumerate(data)), reverse=True)[:16]:
        r = data[i]
        why = max(stalls, key=lambda k: f(r, k))
        where = "loop" if a <= i <= b else "    "
        print(f"  {smp / ts * 100:5.2f}%  line {i:5d} {where} {why[6:]:12s} {r[ix['Source']][:64]}")


if __name__ == "__main__":
    main()
