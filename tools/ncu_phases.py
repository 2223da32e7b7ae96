#!/usr/bin/env python3
"""Per-window (default 100 SASS instructions) sample share of the main loop of a kernel, from
`ncu --page source --csv`: shows which phase of a long straight-line loop body the time goes to.
usage: ncu_phases.py file.csv first_instr last_instr clk_per_iteration [window]"""
import csv, sys
from collections import Counter
rows = list(csv.reader(open(sys.argv[1])))
for i, r in enumerate(rows):
    if r and r[0] == 'Address':
        hdr = r; body = rows[i + 1:]; break
col = {n: i for i, n in enumerate(hdr)}
lo, hi, clk = int(sys.argv[2]), int(sys.argv[3]), float(sys.argv[4])
win = int(sys.argv[5]) if len(sys.argv) > 5 else 100
tot = sum(int(body[k][col['# Samples']] or 0) for k in range(lo, hi))
stalls = [n for n in hdr if n.startswith('stall_') and 'Not Issued' not in n]
print(f"loop samples {tot}; clk/iter {clk}")
for a in range(lo, hi, win):
    b = min(a + win, hi)
    ops = Counter(); s = 0; st = Counter()
    for k in range(a, b):
        r = body[k]; src = r[col['Source']].strip()
        op = src.split()[0] if src else '?'
        if op.startswith('@'): op = src.split()[1]
        cls = 'dp' if op[0] == 'D' and op[:4] in ('DFMA', 'DADD', 'DMUL') else ('cvt' if op[:3] in ('I2F', 'F2I') else ('lds' if op.startswith('LDS') else ('sts' if op.startswith('STS') else ('shfl' if op.startswith('SHFL') else ('tm' if 'TM' in op else ('bar' if op.startswith('BAR') else 'int'))))))
        ops[cls] += 1
        n = int(r[col['# Samples']] or 0); s += n
        for c in stalls:
            v = r[col[c]]
            if v: st[c[6:]] += int(v)
    top = ", ".join(f"{k}={v * 100 // max(s, 1)}%" for k, v in st.most_common(4))
    print(f"{a:5d}-{b:5d} clk {s / tot * clk:7.0f} ({(b - a)} instr: " + " ".join(f"{k}{v}" for k, v in sorted(ops.items())) + f") | {top}")
