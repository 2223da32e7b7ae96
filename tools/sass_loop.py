#!/usr/bin/env python3
"""Opcode census of each blind_rotate kernel instantiation in a .o/.so (cuobjdump -sass): whole
kernel and the CMUX loop body (largest backward-branch span).  usage: sass_loop.py lib.so [filter]"""
import re, subprocess, sys
from collections import Counter
out = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout
flt = sys.argv[2] if len(sys.argv) > 2 else "blind_rotate_v2"
kern = None; ins = {}
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = m.group(1); ins[kern] = []; continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m and kern:
        ins[kern].append((int(m.group(1), 16), m.group(2).strip()))
for k, lst in ins.items():
    if flt not in k: continue
    addr = {a: i for i, (a, _) in enumerate(lst)}
    spans = []
    for i, (a, s) in enumerate(lst):
        m2 = re.search(r"BRA.*?(0x[0-9a-f]+)", s)
        if m2:
            t = int(m2.group(1), 16)
            if t in addr and addr[t] < i:
                spans.append((i - addr[t], addr[t], i))
    spans.sort(reverse=True)
    outer = spans[0]
    inner = [sp for sp in spans[1:] if sp[1] >= outer[1] and sp[2] <= outer[2]]
    n, lo, hi = inner[0] if inner else outer
    c = Counter()
    for a, s in lst[lo:hi + 1]:
        op = s.split()[1] if s.startswith("@") else s.split()[0]
        op = op.split(".")[0] if not op.startswith(("LDS", "STS", "LDL", "STL", "LDTM", "STTM")) else op
        c[op] += 1
    dp = c["DFMA"] + c["DADD"] + c["DMUL"]
    print(f"{k[-40:]}: total {len(lst)} instrs, loop {n} instrs, DP {dp}, "
          f"STL {sum(v for o, v in c.items() if o.startswith('STL'))} LDL {sum(v for o, v in c.items() if o.startswith('LDL'))} "
          f"LDS {sum(v for o, v in c.items() if o.startswith('LDS'))} STS {sum(v for o, v in c.items() if o.startswith('STS'))} "
          f"SHFL {c['SHFL']} I2F {c['I2F']} F2I {c['F2I']} MOV {c['MOV'] + c['IMAD']} BAR {c['BAR']} NOP {c['NOP']}")
