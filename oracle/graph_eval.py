"""Clear-text evaluator of exported level programs -- TEST INFRASTRUCTURE (lives under oracle/).

Evaluates the DAG that libfhs_string builds on PLAINTEXT block values with exactly the semantics
of the encrypted pipeline: a PBS node computes LUT[x] for x in [0,16) and -LUT[x-16] for
x in [16,32) (padding bit set: tfhe-rs' negacyclic convention, oracle/fhs_oracle.c orc_build_lut),
all arithmetic mod 32.  Used by tests/ to check every string op against Python/Rust str semantics
without a GPU, and (optionally) through the oracle's real PBS for small programs."""
import numpy as np


def eval_clear(ex, inputs, check=True):
    kind, lut, t0, tn, cst = ex["kind"], ex["lut"], ex["t0"], ex["tn"], ex["cst"]
    src, coef, luts = ex["term_src"], ex["term_coef"], ex["luts"]
    n = len(kind)
    val = np.zeros(n, dtype=np.int64)
    it = iter(inputs)
    for i in range(n):
        if kind[i] == 0:
            val[i] = next(it)
            continue
        a, b = int(t0[i]), int(t0[i] + tn[i])
        x = int(cst[i]) + int(np.dot(val[src[a:b]], coef[a:b]))
        if kind[i] == 2:
            val[i] = x % 32
            continue
        x %= 32
        if x < 16:
            val[i] = int(luts[lut[i]][x]) % 32
        else:
            val[i] = (-int(luts[lut[i]][x - 16])) % 32
    out = []
    for o in ex["outputs"]:
        v = int(val[o]) if o >= 0 else int(-1 - o)
        if check:
            assert 0 <= v < 16, f"output block out of the 4-bit space: {v}"
        out.append(v)
    return out
