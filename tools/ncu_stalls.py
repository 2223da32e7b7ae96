#!/usr/bin/env python3
"""Summarise `ncu -i X.ncu-rep --page source --csv` for one kernel: stall samples per region
(regions split at BAR.SYNC) and per opcode, so the hot phase of a long straight-line kernel is
visible without the GUI.  usage: ncu_stalls.py file.csv"""
import csv
import sys
from collections import Counter, defaultdict

rows = list(csv.reader(open(sys.argv[1])))
hdr = None
for i, r in enumerate(rows):
    if r and r[0] == "Address":
        hdr = r
        body = rows[i + 1:]
        break
col = {n: i for i, n in enumerate(hdr)}
stall_cols = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
region = 0
reg_samples = defaultdict(int)
reg_stalls = defaultdict(Counter)
reg_ops = defaultdict(Counter)
reg_first = {}
op_samples = Counter()
total = 0
for r in body:
    if len(r) < len(hdr):
        continue
    sass = r[col["Source"]].strip()
    op = sass.split()[0] if sass else "?"
    if op.startswith("@"):
        op = sass.split()[1]
    opc = op.split(".")[0]
    try:
        n = int(float(r[col["# Samples"]] or 0))
    except ValueError:
        n = 0
    total += n
    reg_samples[region] += n
    reg_first.setdefault(region, r[col["Address"]])
    reg_ops[region][opc] += 1
    op_samples[opc] += n
    for s in stall_cols:
        try:
            v = int(float(r[col[s]] or 0))
        except ValueError:
            v = 0
        if v:
            reg_stalls[region][s] += v
    if opc == "BAR":
        region += 1
print("total samples", total)
for g in sorted(reg_samples):
    if reg_samples[g] < 0.005 * total:
        continue
    ops = reg_ops[g]
    dp = ops["DFMA"] + ops["DADD"] + ops["DMUL"]
    top = ", ".join(f"{k[6:]}={v}" for k, v in reg_stalls[g].most_common(5))
    print(f"region {g:2d} @{reg_first[g]} samples {reg_samples[g]:8d} ({100*reg_samples[g]/total:5.1f}%) "
          f"instrs {sum(ops.values()):5d} dp {dp:4d} lds {ops['LDS']:3d} sts {ops['STS']:3d} ldg {ops['LDG']:3d} | {top}")
print("by opcode:", ", ".join(f"{k}={v}" for k, v in op_samples.most_common(14)))
