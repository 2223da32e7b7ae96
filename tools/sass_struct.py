#!/usr/bin/env python3
"""Order of barriers / LDS.128 / STS.128 / LDTM / STTM runs in one kernel's SASS (where ptxas
placed the synchronisation relative to the memory phases).  usage: sass_struct.py file.sass key"""
import re, sys
out = open(sys.argv[1]).read(); key = sys.argv[2]
kern = None; ins = {}
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m: kern = m.group(1); ins[kern] = []; continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m and kern: ins[kern].append(m.group(2).strip())
full = ins[[x for x in ins if key in x][0]]
o = [(s.split()[1] if s.startswith('@') else s.split()[0]) for s in full]
runs = []
for i, x in enumerate(o):
    cls = 'BAR' if x.startswith('BAR') else ('LDS128' if x == 'LDS.128' else ('STS128' if x == 'STS.128' else ('LDTM' if x.startswith('LDTM') else ('STTM' if x.startswith('STTM') else None))))
    if cls:
        if runs and runs[-1][0] == cls and cls != 'BAR' and i - runs[-1][2] < 60: runs[-1][2] = i; runs[-1][3] += 1
        else: runs.append([cls, i, i, 1])
for r in runs:
    if r[1] > 1800: print(r, full[r[1]][:60] if r[0] == 'BAR' else '')
