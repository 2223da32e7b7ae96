#!/usr/bin/env python3
"""profiles/r2_ops_wall_vs_device.txt from the bench lines of N = 1, 2, 4, 8 (the op_latency block of bench.py).
usage: ops_table.py bench_n1.json bench_n2.json bench_n4.json bench_n8.json"""
import json, sys
runs = []
for f in sys.argv[1:]:
    runs.append(json.loads(open(f).read().strip().splitlines()[-1]))
runs.sort(key=lambda d: d["n_gpus"])
print("Round 2 string-op latency, wall vs device (ms; median of 5 after a warm-up call, max over ranks; bench.py op_latency, final tree)")
print("host = host operands + host result in page-locked buffers (H2D + D2H inside)   resident = operands and result in HBM   device = CUDA events around operand copies + levels")
print(f"{'op':<28}" + "".join(f"{'N=' + str(d['n_gpus']):>27}" for d in runs))
print(f"{'':<28}" + "".join(f"{'host / resident / device':>27}" for _ in runs))
for cfg, tag in (("config3_64char", "config3"), ("config4_256char", "config4")):
    for op in runs[0]["op_latency"][cfg]:
        row = f"{tag + ' ' + op:<28}"
        for d in runs:
            v = d["op_latency"][cfg][op]
            row += f"{v['ms']:9.2f} /{v['resident_ms']:8.2f} /{v['device_ms']:7.2f}"
        print(row)
for op in runs[0]["op_latency"]["config5_1024_comparisons"]:
    row = f"{'config5 1024x ' + op:<28}"
    for d in runs:
        v = d["op_latency"]["config5_1024_comparisons"][op]
        row += f"{v['ms']:13.1f} ms {v['comparisons_per_s']:7.0f}/s  "
    print(row)
d1, d8 = runs[0]["op_latency"], runs[-1]["op_latency"]
print("levels / LUT evaluations / levels split by rank at N=%d: " % runs[-1]["n_gpus"] + "; ".join(
    f"{tag} {op}: {d1[cfg][op]['levels']}/{d1[cfg][op]['lut_evals']}/{d8[cfg][op]['sharded_levels']}"
    for cfg, tag in (("config3_64char", "config3"), ("config4_256char", "config4")) for op in d1[cfg]))
print("exchange at N>1: " + runs[-1]["op_latency"].get("exchange", "?"))
print("headline (PBS/s device / e2e): " + "; ".join(f"N={d['n_gpus']}: {d['value']:.0f} / {d['e2e']['value']:.0f}" for d in runs))
