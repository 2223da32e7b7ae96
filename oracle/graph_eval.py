"""Clear-text evaluator of exported level programs -- TEST INFRASTRUCTURE (lives under oracle/).

Evaluates the DAG that libfhs_string builds on PLAINTEXT block values with exactly the semantics
of the encrypted pipeline: a PBS node computes LUT[x] for x in [0,16) and -LUT[x-16] for
x in [16,32) (padding bit set: tfhe-rs' negacyclic convention, oracle/fhs_oracle.c orc_build_lut),
all arithmetic mod 32.  Used by tests/ to check every string op against Python/Rust str semantics
without a GPU, and (optionally) through the oracle's real PBS for small programs.

With `check` it also asserts the range claims the builder made (fhs_prog_export_ranges): the TRUE
integer value of every linear combination lies inside [in_lo, in_hi], an unsigned PBS input inside
[0, 16), a signed one inside (-16, 16), every node's output inside [out_lo, out_hi], so a wrong
`bounded()` override in radix.hpp / split.hpp / replace.hpp fails at the node where it happens instead
of as a decoded-string mismatch."""
import numpy as np


def eval_clear(ex, inputs, check=True):
    kind, lut, t0, tn, cst = ex["kind"], ex["lut"], ex["t0"], ex["tn"], ex["cst"]
    src, coef, luts = ex["term_src"], ex["term_coef"], ex["luts"]
    ranges = check and "in_lo" in ex
    n = len(kind)
    val = np.zeros(n, dtype=np.int64)      # true (signed, unreduced) value of every node
    it = iter(inputs)
    for i in range(n):
        if kind[i] == 0:
            val[i] = next(it)
            if ranges:
                assert ex["out_lo"][i] <= val[i] <= ex["out_hi"][i], f"input {i}: {val[i]} outside its declared range"
            continue
        a, b = int(t0[i]), int(t0[i] + tn[i])
        x = int(cst[i]) + int(np.dot(val[src[a:b]], coef[a:b]))
        if ranges:
            assert ex["in_lo"][i] <= x <= ex["in_hi"][i], \
                f"node {i}: linear combination {x} outside the claimed [{ex['in_lo'][i]}, {ex['in_hi'][i]}]"
        if kind[i] == 2:
            val[i] = x
        else:
            if ranges:
                if ex["signed_in"][i]:
                    assert -16 < x < 16, f"node {i}: signed PBS input {x}"
                else:
                    assert 0 <= x < 16, f"node {i}: PBS input {x} outside the 4-bit plaintext space"
            xm = x % 32
            if xm < 16:
                val[i] = int(luts[lut[i]][xm]) % 32
            else:
                val[i] = -(int(luts[lut[i]][xm - 16]) % 32)
            if val[i] <= -16 or val[i] >= 16:        # LUT entries >= 16 are negative values mod 32
                val[i] = ((val[i] + 16) % 32) - 16
        if ranges:
            assert ex["out_lo"][i] <= val[i] <= ex["out_hi"][i], \
                f"node {i} (kind {kind[i]}): value {val[i]} outside the claimed [{ex['out_lo'][i]}, {ex['out_hi'][i]}]"
    out = []
    for o in ex["outputs"]:
        v = int(val[o]) % 32 if o >= 0 else int(-1 - o)
        if check:
            assert 0 <= v < 16, f"output block out of the 4-bit space: {v}"
        out.append(v)
    return out
