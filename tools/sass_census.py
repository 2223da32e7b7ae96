#!/usr/bin/env python3
"""Opcode census per kernel of libfhs_cuda.so (cuobjdump -sass): the instructions that prove which
hardware paths a kernel uses -- UTCIMMA (tcgen05.mma kind::i8), LDTM / STTM (tcgen05.ld / st: TMEM),
UBLKCP (cp.async.bulk), UTMALDG (tensor-map TMA), STAS (st.async to a peer CTA's shared memory),
DFMA / DMUL / DADD (FP64 pipe), LDS / STS, SHFL, BAR, IMAD.  usage: sass_census.py [lib.so] > profiles/rN_sass_census.txt"""
import os
import re
import subprocess
import sys
from collections import Counter

lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.abspath(__file__)), "..",
                                                          "fhe_string_b200", "libfhs_cuda.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
dem = {}
kern, ins = None, {}
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = m.group(1)
        ins[kern] = Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(.*?);", line)
    if m and kern:
        s = m.group(1).strip()
        op = s.split()[1] if s.startswith("@") else s.split()[0]
        ins[kern][op.split(".")[0]] += 1
names = subprocess.run(["c++filt"] + list(ins), capture_output=True, text=True).stdout.splitlines()
COLS = ["UTCIMMA", "LDTM", "STTM", "UBLKCP", "UTMALDG", "STAS", "DFMA", "DMUL", "DADD", "LDS", "STS", "LDG", "STG",
        "SHFL", "BAR", "IMAD", "F2I", "I2F"]
print(f"# {os.path.basename(lib)}: SASS opcode counts per kernel (static instruction counts, whole kernel)")
print(f"{'kernel':58s} {'total':>7s} " + " ".join(f"{c:>7s}" for c in COLS))
tot = Counter()
for k, nm in sorted(zip(ins, names), key=lambda x: x[1]):
    c = ins[k]
    short = nm.replace("fhs::(anonymous namespace)::", "").replace("void ", "")
    short = re.sub(r"\(.*", "", short)
    print(f"{short[:58]:58s} {sum(c.values()):7d} " + " ".join(f"{c[x]:7d}" for x in COLS))
    tot.update(c)
print(f"{'ALL KERNELS':58s} {sum(tot.values()):7d} " + " ".join(f"{tot[x]:7d}" for x in COLS))
