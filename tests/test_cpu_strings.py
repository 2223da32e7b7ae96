"""CPU suite: every string op's level program, evaluated in the clear (oracle/graph_eval.py), against
Rust std::str semantics (= Python str for ASCII) on the reference's own test tables
(src/ciphertext/tests/*.rs) and on random inputs."""
import random

import pytest

from fhe_string_b200.program import Program
from oracle.graph_eval import eval_clear
from strutil import WS, decode, enc_chars, to_blocks, usize_blocks

_cache = {}
# Largest variance multiplier (in units of one PBS output, as fhs_prog_info reports it: rounded up) any
# level program may feed into a PBS.  test_noise_budget_failure_probability turns it into a per-evaluation
# failure probability (DESIGN.md 4 "Noise budget").
MAX_NOISE = 163


def run(op, strs, pads=None, usizes=(), scalars=(), nbu=4):
    pads = pads or [None] * len(strs)
    chars = [enc_chars(s, p) for s, p in zip(strs, pads)]
    key = (op, tuple(len(c) for c in chars), len(usizes), tuple(scalars), nbu)
    if key not in _cache:
        p = Program(op, [len(c) for c in chars], len(usizes), scalars, 4, nbu)
        _cache[key] = (p, p.export())
    p, ex = _cache[key]
    inputs = [b for c in chars for b in to_blocks(c)] + [b for u in usizes for b in usize_blocks(u, nbu)]
    assert len(inputs) == p.n_inputs
    assert p.max_noise <= MAX_NOISE, f"{op}: variance multiplier {p.max_noise} fed to a PBS"
    out = eval_clear(ex, inputs)        # asserts every range claim of the builder on the way
    assert all(v <= 3 for v in out), f"{op}: output digit above 3 (clean-block invariant for chained ops): {out}"
    return decode(p.desc, out, nbu)


# (input, pattern, pad) rows in the style of tests/search.rs:15-40, tests/compare.rs:67-110
PAIRS = [("abc", "abc", None, None), ("abc", "abd", None, None), ("abc", "ab", None, None), ("ab", "abc", None, None),
         ("", "", None, None), ("a", "", None, None), ("", "a", None, None), ("abc", "abc", 5, None),
         ("abc", "abc", None, 6), ("abc", "abc", 7, 4), ("abc", "b", 5, 3), ("abcabc", "bc", None, None),
         ("abcabc", "cab", 8, None), ("abc", "abcd", None, None), ("Hello", "hELLO", None, 7),
         (" A bcbc ", "bc", None, None), ("xxx", "x", None, None), ("zzz", "zzzz", 4, 4)]


@pytest.mark.parametrize("a,b,pa,pb", PAIRS)
def test_compare_ops(a, b, pa, pb):
    assert run("eq", [a, b], [pa, pb]) == (a == b)
    assert run("ne", [a, b], [pa, pb]) == (a != b)
    assert run("lt", [a, b], [pa, pb]) == (a < b)
    assert run("le", [a, b], [pa, pb]) == (a <= b)
    assert run("gt", [a, b], [pa, pb]) == (a > b)
    assert run("ge", [a, b], [pa, pb]) == (a >= b)
    assert run("eq_ignore_ascii_case", [a, b], [pa, pb]) == (a.lower() == b.lower())


@pytest.mark.parametrize("a,b,pa,pb", PAIRS)
def test_search_ops(a, b, pa, pb):
    assert run("contains", [a, b], [pa, pb]) == (b in a)
    f = a.find(b)
    assert run("find", [a, b], [pa, pb]) == (None if f < 0 else f)
    r = a.rfind(b)
    assert run("rfind", [a, b], [pa, pb]) == (None if r < 0 else r)
    assert run("starts_with", [a, b], [pa, pb]) == a.startswith(b)
    assert run("ends_with", [a, b], [pa, pb]) == a.endswith(b)
    assert run("strip_prefix", [a, b], [pa, pb]) == (a[len(b):] if a.startswith(b) else None)
    assert run("strip_suffix", [a, b], [pa, pb]) == (a[:len(a) - len(b)] if a.endswith(b) else None)


@pytest.mark.parametrize("s,pad", [("", None), ("a", None), ("abc", 5), ("  a b  ", None), ("\t\n a\r", 8),
                                   ("   ", None), ("aZ@[`{z", None), ("Hello, World! 09", 20), (" A bcbc ", None)])
def test_unary_ops(s, pad):
    assert run("is_empty", [s], [pad]) == (s == "")
    assert run("len", [s], [pad]) == len(s)
    assert run("to_uppercase", [s], [pad]) == s.upper()
    assert run("to_lowercase", [s], [pad]) == s.lower()
    assert run("trim", [s], [pad]) == s.strip(WS)
    assert run("trim_start", [s], [pad]) == s.lstrip(WS)
    assert run("trim_end", [s], [pad]) == s.rstrip(WS)


def test_substr():
    s = "hello world"
    for i in (0, 1, 5, 11, 12, 14):
        assert run("substr_to", [s], [13], usizes=[i]) == s[:i]
        assert run("substr_from", [s], [13], usizes=[i]) == s[i:]


def test_random_against_python_str():
    rng = random.Random(1234)
    alphabet = "abAB \t"
    for _ in range(40):
        a = "".join(rng.choice(alphabet) for _ in range(rng.randint(0, 7)))
        b = "".join(rng.choice(alphabet) for _ in range(rng.randint(0, 3)))
        pa, pb = 8, 4
        assert run("eq", [a, b], [pa, pb]) == (a == b)
        assert run("lt", [a, b], [pa, pb]) == (a < b)
        assert run("contains", [a, b], [pa, pb]) == (b in a)
        f, r = a.find(b), a.rfind(b)
        assert run("find", [a, b], [pa, pb]) == (None if f < 0 else f)
        assert run("rfind", [a, b], [pa, pb]) == (None if r < 0 else r)
        assert run("ends_with", [a, b], [pa, pb]) == a.endswith(b)
        assert run("strip_suffix", [a, b], [pa, pb]) == (a[:len(a) - len(b)] if a.endswith(b) else None)
        assert run("trim", [a], [pa]) == a.strip(WS)
        assert run("to_uppercase", [a], [pa]) == a.upper()
        assert run("len", [a], [pa]) == len(a)


def test_five_block_usize_long_string():
    """l >= 256 needs 5-block FheUsize (lib.rs:62; SURVEY.md 7 'parameter trap')."""
    s = "ab" * 130
    assert run("len", [s], [260], nbu=5) == 260
    assert run("find", [s + "c", "bc"], [262, None], nbu=5) == 259


def rust_split_inclusive(s, p):
    out, i = [], 0
    if p == "":   # Rust: "abc".split_inclusive("") == ["a", "b", "c"] -- matches at 0 close empty piece... see below
        return None
    while True:
        j = s.find(p, i)
        if j < 0:
            break
        out.append(s[i:j + len(p)])
        i = j + len(p)
    if i < len(s):
        out.append(s[i:])
    return out


SPLIT_ROWS = [("a,b,c", ",", None, None), ("xxx", "x", None, None), ("xxx", "xx", None, None), ("abc", "", None, None),
              ("", "x", None, None), ("", "", None, None), ("a,b,", ",", 6, 2), (",a", ",", None, 3),
              ("abcabca", "abc", 9, None), (" A bcbc ", "bc", None, None), ("aaaa", "aa", 5, 3)]


@pytest.mark.parametrize("s,p,ps,pp", SPLIT_ROWS)
def test_split_family(s, p, ps, pp):
    if p != "":
        assert run("split", [s, p], [ps, pp]) == s.split(p)
        assert run("rsplit", [s, p], [ps, pp]) == s.rsplit(p)[::-1]
        term = s.split(p)
        if term and term[-1] == "":
            term = term[:-1]
        assert run("split_terminator", [s, p], [ps, pp]) == term
        term_r = s.rsplit(p)           # right-to-left non-overlapping search (search.rs:114-152)
        if term_r and term_r[-1] == "":
            term_r = term_r[:-1]
        assert run("rsplit_terminator", [s, p], [ps, pp]) == term_r[::-1]
        assert run("split_inclusive", [s, p], [ps, pp]) == rust_split_inclusive(s, p)
        for n in (0, 1, 2, 3, 9):
            want = [] if n == 0 else s.split(p, n - 1)
            assert run("splitn", [s, p], [ps, pp], usizes=[n]) == want
            want = [] if n == 0 else s.rsplit(p, n - 1)[::-1]
            assert run("rsplitn", [s, p], [ps, pp], usizes=[n]) == want
        f = s.find(p)
        assert run("split_once", [s, p], [ps, pp]) == (None if f < 0 else (s[:f], s[f + len(p):]))
        r = s.rfind(p)
        assert run("rsplit_once", [s, p], [ps, pp]) == (None if r < 0 else (s[:r], s[r + len(p):]))
    else:
        # Rust: "abc".split("") == ["", "a", "b", "c", ""] (tests/split.rs:33-47)
        assert run("split", [s, p], [ps, pp]) == [""] + list(s) + [""]
        assert run("rsplit", [s, p], [ps, pp]) == ([""] + list(s) + [""])[::-1]


@pytest.mark.parametrize("s,pad", [("a b  c", None), ("  lead and trail \t", 20), ("", None), ("   ", 4), ("one", 5),
                                   ("a\tb\nc\rd", None)])
def test_split_ascii_whitespace(s, pad):
    assert run("split_ascii_whitespace", [s], [pad]) == s.split()


REPLACE_ROWS = [("abcabc", "bc", "X", None, None, None), ("abcabc", "bc", "", None, None, None),
                ("abcabc", "b", "xyz", 8, 2, 4), ("aaa", "aa", "b", None, None, None), ("abc", "", "-", None, None, None),
                ("", "", "x", None, None, None), ("abc", "abc", "", 5, 4, 2), ("abc", "d", "x", None, None, None),
                (" A bcbc ", "bc", "_", None, None, None), ("xxxx", "x", "yy", None, None, None)]


@pytest.mark.parametrize("a,p,s,pa,pp,ps", REPLACE_ROWS)
def test_replace(a, p, s, pa, pp, ps):
    want = a.replace(p, s)
    assert run("replace", [a, p, s], [pa, pp, ps], scalars=[len(want)]) == want
    assert run("replace", [a, p, s], [pa, pp, ps], scalars=[len(want) + 3]) == want
    for n in (0, 1, 2, 5):
        wn = a.replace(p, s, n)
        assert run("replacen", [a, p, s], [pa, pp, ps], usizes=[n], scalars=[len(wn) + 1]) == wn


# insert.rs:14-124 (tables in the style of tests/insert.rs:20-45,76-95)
@pytest.mark.parametrize("a,b,pa,pb", [("abc", "xy", None, None), ("", "", None, None), ("a", "", None, None),
                                       ("", "b", None, None), ("abc", "xy", 5, 4), ("hello", "world", 7, None)])
def test_add_and_insert(a, b, pa, pb):
    assert run("add", [a, b], [pa, pb]) == a + b
    for i in range(len(a) + 1):
        assert run("insert", [a, b], [pa, pb], usizes=[i]) == a[:i] + b + a[i:]


@pytest.mark.parametrize("a,pa", [("ab", None), ("abc", 4), ("", None), ("x", None), ("", 3)])
def test_repeat(a, pa):
    for n in (0, 1, 2, 3, 5, 8, 20):
        for l in (0, 1, 4, 7, 16):
            assert run("repeat", [a], [pa], usizes=[n], scalars=(l,)) == (a * n)[:l]


def test_slice_vector_len_and_get():
    """FheStringSliceVector::len / get (split.rs:32-83) chained after split: the split program's
    output blocks are the input of slices_len / slices_get."""
    for s, pat in ((" A bcbc ", "bc"), ("a,b,,c", ","), ("xxx", "x"), ("abc", "zz")):
        chars = [enc_chars(s, 9), enc_chars(pat)]
        sp = Program("split", [len(c) for c in chars], 0, (), 4, 4)
        vals = eval_clear(sp.export(), [b for c in chars for b in to_blocks(c)])
        n_entries = int(sp.desc.split(":")[1])
        want = s.split(pat)
        ln = Program("slices_len", [len(chars[0])], 0, (n_entries,), 4, 4)
        assert decode(ln.desc, eval_clear(ln.export(), vals)) == len(want)
        gt = Program("slices_get", [len(chars[0])], 0, (n_entries,), 4, 4)
        ex = gt.export()
        for i in range(len(want) + 2):
            got = decode(gt.desc, eval_clear(ex, vals + usize_blocks(i)))
            assert got == (want[i] if i < len(want) else None), (s, pat, i, got)


def _rust_rsplit(s, p):
    """Rust rsplit: non-overlapping matches found right to left, pieces listed right to left."""
    return s.rsplit(p)[::-1]


LONG_OVERLAP = [("a" * 30, "aa", 33, None), ("a" * 31, "aaa", None, 5), ("ab" * 14 + "a", "aba", 32, None),
                ("abcabcabcabcabcabcabcabcabcabc", "abcabc", None, 8), ("x" * 25 + "yx" * 4, "xx", 40, 4),
                ("aabaabaabaabaabaabaabaab", "aabaa", None, None), ("b" * 24, "a", 30, 3)]


@pytest.mark.parametrize("s,p,ps,pp", LONG_OVERLAP)
def test_split_replace_long_overlapping(s, p, ps, pp):
    """Strings long enough for the chunked (speculative) non-overlapping match selection and the
    last-set scan of split.hpp, with matches that overlap in every possible phase."""
    assert run("split", [s, p], [ps, pp]) == s.split(p)
    assert run("rsplit", [s, p], [ps, pp]) == _rust_rsplit(s, p)
    assert run("split_inclusive", [s, p], [ps, pp]) == rust_split_inclusive(s, p)
    for n in (1, 2, 4, 30):
        assert run("splitn", [s, p], [ps, pp], usizes=[n]) == s.split(p, n - 1)
        assert run("rsplitn", [s, p], [ps, pp], usizes=[n]) == s.rsplit(p, n - 1)[::-1]
    want = s.replace(p, "zz")
    assert run("replace", [s, p, "zz"], [ps, pp, None], scalars=[len(want) + 1]) == want
    wn = s.replace(p, "q", 3)
    assert run("replacen", [s, p, "q"], [ps, pp, None], usizes=[3], scalars=[len(wn) + 1]) == wn


def test_split_random_two_letter_alphabet():
    rng = random.Random(5)
    for _ in range(25):
        s = "".join(rng.choice("ab") for _ in range(rng.randint(20, 34)))
        p = "".join(rng.choice("ab") for _ in range(rng.randint(2, 4)))
        pp = rng.choice([None, 5, 7])
        assert run("split", [s, p], [36, pp]) == s.split(p)
        assert run("rsplit", [s, p], [36, pp]) == _rust_rsplit(s, p)


def test_hypothesis_generated_ascii_strings():
    """SURVEY 8(c) tier 1: hypothesis-generated ASCII inputs (small alphabet so that matches, overlaps
    and repeated separators actually occur), every result against Python str semantics."""
    from hypothesis import given, settings, strategies as st

    alpha = st.sampled_from("ab ,")
    text = st.text(alphabet=alpha, min_size=0, max_size=10)
    pat = st.text(alphabet=alpha, min_size=0, max_size=3)

    @settings(max_examples=60, deadline=None, derandomize=True)
    @given(s=text, p=pat, ps=st.sampled_from([None, 12]), pp=st.sampled_from([None, 4]))
    def check(s, p, ps, pp):
        assert run("contains", [s, p], [ps, pp]) == (p in s)
        f = s.find(p)
        assert run("find", [s, p], [ps, pp]) == (None if f < 0 else f)
        r = s.rfind(p)
        assert run("rfind", [s, p], [ps, pp]) == (None if r < 0 else r)
        assert run("lt", [s, p], [ps, pp]) == (s < p)
        assert run("eq", [s, p], [ps, pp]) == (s == p)
        assert run("trim", [s], [ps]) == s.strip(WS)
        assert run("to_uppercase", [s], [ps]) == s.upper()
        if p != "":
            assert run("split", [s, p], [ps, pp]) == s.split(p)
            assert run("rsplit", [s, p], [ps, pp]) == s.rsplit(p)[::-1]
            want = s.replace(p, "xy")
            assert run("replace", [s, p, "xy"], [ps, pp, None], scalars=[len(want) + 1]) == want
        assert run("split_ascii_whitespace", [s], [ps]) == s.split()

    check()


def test_ordering_long_strings_first_difference_everywhere():
    """Fan-in-4 signed sign reduction (radix.hpp reduce_sign) on 40-char strings: the first difference
    at every depth of the reduction tree, both directions, plus prefix / equal cases."""
    rng = random.Random(11)
    base = "".join(rng.choice("abcdefghijklmnopqrstuvwxyz") for _ in range(40))
    cases = [(base, base), (base, base[:-1]), (base[:-1], base), (base, base + "a"), ("", base), (base, "")]
    for pos in (0, 1, 3, 4, 7, 15, 16, 31, 39):
        for delta in (-1, 1):
            other = base[:pos] + chr(ord(base[pos]) + delta) + "".join(rng.choice("az") for _ in range(39 - pos))
            cases.append((base, other))
    for a, b in cases:
        assert run("lt", [a, b], [44, 41]) == (a < b), (a, b)
        assert run("le", [a, b], [44, 41]) == (a <= b), (a, b)
        assert run("gt", [a, b], [44, 41]) == (a > b), (a, b)
        assert run("ge", [a, b], [44, 41]) == (a >= b), (a, b)


def test_noise_budget_failure_probability():
    """The PBS input is  M x (PBS output variance) + keyswitch + modulus-switch noise; a LUT evaluation
    fails when that exceeds half a box (2^-6 of the torus).  With the oracle's measured PBS output variance
    (tests/golden/noise_golden.json) and the analytic keyswitch / modulus-switch terms
    (tests/test_cpu_oracle_first_principles.py checks the former against samples), M = MAX_NOISE must keep the
    per-evaluation failure probability below tfhe-rs' 2^-40 target for this parameter set."""
    import json
    import math
    import os
    gold = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "noise_golden.json")))
    var_pbs = gold["variance"]
    n, kN, lwe_std = 742, 2048, 7.069849454709433e-6
    var_ks = kN * 5 * 66 / 12.0 * lwe_std**2 + (kN / 2.0) * 2.0**-30 / 12.0
    var_ms = (1.0 / 4096) ** 2 / 12.0 * (1 + n / 2.0)
    def log2_pfail(m):
        z = (1.0 / 64) / math.sqrt(m * var_pbs + var_ks + var_ms)
        return math.log2(math.erfc(z / math.sqrt(2.0)))
    # fresh input 2^-40.3 (the parameter set itself, tfhe-rs advertises 2^-40 for it); tfhe-rs' bivariate
    # pack 4x + y (17) 2^-40.2; one-hot digit sums here (<= 64) 2^-40.0; fan-in-4 sign reduction (85)
    # 2^-39.9; the 8-candidate selections of the chunked match scan in split / replace (<= 163) 2^-39.5
    assert -40.4 < log2_pfail(1) < -40.2
    assert log2_pfail(17) < -40.2
    assert log2_pfail(64) < -39.95
    assert log2_pfail(MAX_NOISE) < -39.4, log2_pfail(MAX_NOISE)
    assert log2_pfail(400) > -38.5         # round 1's test bound (400) was well outside the budget


def test_noise_multiplier_per_op_class():
    """Which programs reach which multiplier (64-char operands, 8-char pattern)."""
    def mx(op, nc, nu=0, sc=()):
        return Program(op, nc, nu, sc, 4, 4).max_noise
    for op in ("eq", "contains", "starts_with", "strip_suffix"):
        assert mx(op, [65, 9]) <= 34
    for op, nc in (("find", [65, 9]), ("rfind", [65, 9]), ("len", [65]), ("trim", [65]), ("add", [65, 9])):
        assert mx(op, nc) <= 64
    assert mx("to_uppercase", [65]) <= 17
    assert mx("lt", [65, 65]) <= 85
    assert mx("split", [65, 9]) <= MAX_NOISE and mx("replace", [65, 9, 3], 0, (64,)) <= MAX_NOISE


# ---- the rest of the public surface (VERDICT r1 missing-4): mod.rs:198,217, compare.rs:119, convert.rs:14,22,
# trim.rs:15, split.rs:86 ----
def test_char_at_and_substr_end():
    s = "hello world"
    for i in (0, 4, 10, 11, 12, 15):
        assert run("char_at", [s], [13], usizes=[i]) == (ord(s[i]) if i < len(s) else 0)
    for a, b in ((0, 5), (6, 11), (3, 3), (5, 2), (0, 11), (9, 14)):
        assert run("substr_end", [s], [13], usizes=[a, b]) == s[a:b]


def test_substr_eq_clear_offset():
    for a, b, pa, pb in (("abcabc", "bc", None, None), ("abcabc", "bc", 8, 4), ("abc", "", None, None), ("abc", "abcd", None, None),
                         ("ab", "b", None, 3)):
        for i in range(len(a) + 2):
            assert run("substr_eq", [a, b], [pa, pb], scalars=(i,)) == (a[i:i + len(b)] == b), (a, b, i)


def test_ascii_char_predicates():
    p = {k: Program(k, [1], 0, (), 4, 4) for k in ("char_is_uppercase", "char_is_lowercase", "char_is_whitespace")}
    ex = {k: v.export() for k, v in p.items()}
    for c in range(128):
        blocks = to_blocks([c])
        assert bool(eval_clear(ex["char_is_uppercase"], blocks)[0]) == (65 <= c <= 90), c
        assert bool(eval_clear(ex["char_is_lowercase"], blocks)[0]) == (97 <= c <= 122), c
        assert bool(eval_clear(ex["char_is_whitespace"], blocks)[0]) == (9 <= c <= 13 or c == 32), c
    assert all(v.depth == 2 and v.pbs_count == 3 for v in p.values())


def test_slice_vector_truncate():
    for s, pat in (("a,b,,c", ","), ("xxx", "x"), ("abc", "zz")):
        chars = [enc_chars(s, 8), enc_chars(pat)]
        sp = Program("split", [len(c) for c in chars], 0, (), 4, 4)
        vals = eval_clear(sp.export(), [b for c in chars for b in to_blocks(c)])
        n_entries = int(sp.desc.split(":")[1])
        want = s.split(pat)
        tr = Program("slices_truncate", [len(chars[0])], 0, (n_entries,), 4, 4)
        ex = tr.export()
        for i in range(len(want) + 2):
            flags = decode(tr.desc, eval_clear(ex, vals + usize_blocks(i)))
            # rebuild the vector with the new flags and decode it like split's output
            p = 4 * len(chars[0])
            new = list(vals)
            for e in range(n_entries):
                new[p] = int(flags[e])
                p += 1 + 2 * 4
            assert decode(sp.desc, new) == want[:i], (s, pat, i)


def test_fuzz_every_op_against_rust_semantics():
    """Seeded fuzz over every op family (empty patterns, padded and unpadded operands, truncating output
    capacities, splitn / replacen counts) against tests/rust_str.py; 400 iterations of this loop (9 323 checks)
    were run once by hand when it was written, the suite keeps 50."""
    import rust_str as R
    rng = random.Random(2024)
    alpha = "ab A,\t"
    for _ in range(50):
        a = "".join(rng.choice(alpha) for _ in range(rng.randint(0, 9)))
        p = "".join(rng.choice(alpha) for _ in range(rng.randint(0, 3)))
        s = "".join(rng.choice("xyz") for _ in range(rng.randint(0, 3)))
        pa = rng.choice([None, len(a) + rng.randint(0, 3)])
        pp = rng.choice([None, len(p) + rng.randint(0, 2)])
        ps = rng.choice([None, len(s) + 1])
        k, l = rng.randint(0, 4), rng.randint(0, 12)
        checks = [("split", [a, p], [pa, pp], (), (), R.split(a, p)), ("rsplit", [a, p], [pa, pp], (), (), R.rsplit(a, p)),
                  ("split_terminator", [a, p], [pa, pp], (), (), R.split_terminator(a, p)),
                  ("split_inclusive", [a, p], [pa, pp], (), (), R.split_inclusive(a, p)),
                  ("splitn", [a, p], [pa, pp], [k], (), R.splitn(a, k, p)), ("rsplitn", [a, p], [pa, pp], [k], (), R.rsplitn(a, k, p)),
                  ("split_once", [a, p], [pa, pp], (), (), R.split_once(a, p)), ("rsplit_once", [a, p], [pa, pp], (), (), R.rsplit_once(a, p)),
                  ("find", [a, p], [pa, pp], (), (), R.find(a, p)), ("rfind", [a, p], [pa, pp], (), (), R.rfind(a, p)),
                  ("strip_prefix", [a, p], [pa, pp], (), (), R.strip_prefix(a, p)), ("strip_suffix", [a, p], [pa, pp], (), (), R.strip_suffix(a, p)),
                  ("ends_with", [a, p], [pa, pp], (), (), a.endswith(p)), ("lt", [a, p], [pa, pp], (), (), a < p),
                  ("eq_ignore_ascii_case", [a, p], [pa, pp], (), (), a.lower() == p.lower()),
                  ("trim", [a], [pa], (), (), a.strip(R.WS)), ("split_ascii_whitespace", [a], [pa], (), (), R.split_ascii_whitespace(a)),
                  ("add", [a, p], [pa, pp], (), (), a + p),
                  ("replace", [a, p, s], [pa, pp, ps], (), (l,), R.replace(a, p, s)[:l]),
                  ("replacen", [a, p, s], [pa, pp, ps], [k], (l,), R.replacen(a, p, s, k)[:l]),
                  ("repeat", [a], [pa], [k], (l,), (a * k)[:l])]
        for op, strs, pads, us, sc, want in checks:
            got = run(op, strs, pads, usizes=us, scalars=sc)
            if isinstance(want, tuple):
                got = tuple(got) if got is not None else None
            assert got == want, (op, strs, pads, us, sc, got, want)
