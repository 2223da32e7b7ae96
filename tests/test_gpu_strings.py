"""GPU parity: string ops executed as level programs on the B200 (libfhs_string -> libfhs_cuda),
encrypted with the oracle's client key, decrypted results against std::str semantics
(the reference's own acceptance criterion, src/ciphertext/tests/*.rs)."""
import numpy as np
import pytest

from fhe_string_b200.program import Program
from strutil import WS, decode, enc_chars, to_blocks, usize_blocks

pytestmark = pytest.mark.gpu


def run_gpu(engine, keyset, op, strs, pads=None, usizes=(), nbu=4, seed=0):
    pads = pads or [None] * len(strs)
    chars = [enc_chars(s, p) for s, p in zip(strs, pads)]
    p = Program(op, [len(c) for c in chars], len(usizes), (), 4, nbu)
    vals = [b for c in chars for b in to_blocks(c)] + [b for u in usizes for b in usize_blocks(u, nbu)]
    cts = keyset.encrypt(np.array(vals, dtype=np.uint64), seed=seed + 17)
    out, stats = p.run(engine, cts)
    dec = [int(x) for x in keyset.decrypt(out)]
    assert all(0 <= v < 16 for v in dec), dec
    assert stats["pbs"] == p.pbs_count and stats["levels"] == p.depth
    return decode(p.desc, dec, nbu), stats


def test_cmd_example_config(engine, keyset):
    """BASELINE configs[0]: examples/cmd --input " A bcbc " --pattern "bc"."""
    s, pat = " A bcbc ", "bc"
    assert run_gpu(engine, keyset, "contains", [s, pat])[0] is True
    assert run_gpu(engine, keyset, "find", [s, pat])[0] == s.find(pat)
    assert run_gpu(engine, keyset, "rfind", [s, pat])[0] == s.rfind(pat)
    assert run_gpu(engine, keyset, "starts_with", [s, pat])[0] == s.startswith(pat)
    assert run_gpu(engine, keyset, "ends_with", [s, pat])[0] == s.endswith(pat)
    assert run_gpu(engine, keyset, "eq", [s, pat])[0] is False
    assert run_gpu(engine, keyset, "lt", [s, pat])[0] == (s < pat)
    assert run_gpu(engine, keyset, "ge", [s, pat])[0] == (s >= pat)
    assert run_gpu(engine, keyset, "len", [s])[0] == len(s)
    assert run_gpu(engine, keyset, "is_empty", [s])[0] is False
    assert run_gpu(engine, keyset, "to_uppercase", [s])[0] == s.upper()
    assert run_gpu(engine, keyset, "to_lowercase", [s])[0] == s.lower()
    assert run_gpu(engine, keyset, "trim", [s])[0] == s.strip(WS)
    assert run_gpu(engine, keyset, "trim_start", [s])[0] == s.lstrip(WS)
    assert run_gpu(engine, keyset, "trim_end", [s])[0] == s.rstrip(WS)
    assert run_gpu(engine, keyset, "strip_prefix", [s, " A"])[0] == s[2:]
    assert run_gpu(engine, keyset, "strip_suffix", [s, "bc "])[0] == s[:-3]
    assert run_gpu(engine, keyset, "eq_ignore_ascii_case", [s, " a BCBC "])[0] is True


def test_padded_64_char_ops(engine, keyset):
    """BASELINE configs[2]: eq / contains / find on a 64-char padded string, 8-char pattern."""
    rng = np.random.default_rng(7)
    s = "".join(chr(int(c)) for c in rng.integers(33, 127, 40))
    hit, miss = s[13:21], "~~zz~~zz"
    for pat in (hit, miss):
        v, st = run_gpu(engine, keyset, "contains", [s, pat], [64, None])
        assert v == (pat in s)
        f = s.find(pat)
        assert run_gpu(engine, keyset, "find", [s, pat], [64, None])[0] == (None if f < 0 else f)
    assert run_gpu(engine, keyset, "eq", [s, s], [64, 64])[0] is True
    assert run_gpu(engine, keyset, "eq", [s, s[:-1] + "!"], [64, 64])[0] is False
    assert run_gpu(engine, keyset, "lt", [s, s[:20] + "~"], [64, 64])[0] == (s < s[:20] + "~")


def test_substr_encrypted_index(engine, keyset):
    s = "hello world"
    assert run_gpu(engine, keyset, "substr_from", [s], [12], usizes=[6])[0] == s[6:]
    assert run_gpu(engine, keyset, "substr_to", [s], [12], usizes=[5])[0] == s[:5]


def test_split_and_replace_cmd_config(engine, keyset):
    """split / replace rows of examples/cmd on " A bcbc " with pattern "bc", substitution "_"."""
    s, pat, sub = " A bcbc ", "bc", "_"
    assert run_gpu(engine, keyset, "split", [s, pat])[0] == s.split(pat)
    assert run_gpu(engine, keyset, "rsplit", [s, pat])[0] == s.rsplit(pat)[::-1]
    assert run_gpu(engine, keyset, "split_inclusive", [s, pat])[0] == [" A bc", "bc", " "]
    assert run_gpu(engine, keyset, "split_terminator", [s, pat])[0] == s.split(pat)
    assert run_gpu(engine, keyset, "split_ascii_whitespace", [s])[0] == s.split()
    assert run_gpu(engine, keyset, "splitn", [s, pat], usizes=[2])[0] == s.split(pat, 1)
    assert run_gpu(engine, keyset, "rsplitn", [s, pat], usizes=[2])[0] == s.rsplit(pat, 1)[::-1]
    assert run_gpu(engine, keyset, "split_once", [s, pat])[0] == tuple(s.split(pat, 1))
    assert run_gpu(engine, keyset, "rsplit_once", [s, pat])[0] == tuple(s.rsplit(pat, 1))
    want = s.replace(pat, sub)
    p_chars = [enc_chars(x) for x in (s, pat, sub)]
    for op, usz, w in (("replace", (), want), ("replacen", (1,), s.replace(pat, sub, 1))):
        p = Program(op, [len(c) for c in p_chars], len(usz), (len(w),), 4, 4)
        vals = [b for c in p_chars for b in to_blocks(c)] + [b for u in usz for b in usize_blocks(u)]
        out, _ = p.run(engine, keyset.encrypt(np.array(vals, dtype=np.uint64), seed=5))
        assert decode(p.desc, [int(x) for x in keyset.decrypt(out)]) == w


def test_run_many_independent_comparisons(engine, keyset):
    """BASELINE configs[4] in miniature: M independent (a, b) pairs through fhs_prog_run_many (every
    level evaluated for all instances at once) give the same booleans as Python's str compare."""
    rng = np.random.default_rng(13)
    pairs = []
    for i in range(12):
        a = "".join(chr(int(c)) for c in rng.integers(97, 123, 6))
        b = a if i % 3 == 0 else a[:3] + "".join(chr(int(c)) for c in rng.integers(97, 123, 3))
        pairs.append((a, b))
    for op, f in (("eq", lambda a, b: a == b), ("lt", lambda a, b: a < b), ("contains", lambda a, b: b[:3] in a)):
        rhs = [(a, b[:3]) if op == "contains" else (a, b) for a, b in pairs]
        chars = [[enc_chars(a, 8), enc_chars(b, 8 if op != "contains" else None)] for a, b in rhs]
        p = Program(op, [len(c) for c in chars[0]], 0, (), 4, 4)
        vals = np.array([[v for c in inst for v in to_blocks(c)] for inst in chars], dtype=np.uint64)
        cts = keyset.encrypt(vals.reshape(-1), seed=23).reshape(len(pairs), p.n_inputs, -1)
        out, st = p.run_many(engine, cts)
        assert st["pbs"] == p.pbs_count * len(pairs) and st["levels"] == p.depth
        got = [decode(p.desc, [int(x) for x in keyset.decrypt(out[m])], 4) for m in range(len(pairs))]
        assert got == [f(a, b) for a, b in pairs]


def test_add_insert_repeat(engine, keyset):
    """insert.rs:14-124 on the GPU: add / insert / repeat of the cmd config strings."""
    s, pat = " A bcbc ", "bc"
    assert run_gpu(engine, keyset, "add", [s, pat])[0] == s + pat
    assert run_gpu(engine, keyset, "insert", [s, pat], usizes=[3])[0] == s[:3] + pat + s[3:]
    chars = [enc_chars(pat)]
    p = Program("repeat", [len(chars[0])], 1, (6,), 4, 4)
    vals = to_blocks(chars[0]) + usize_blocks(3)
    out, _ = p.run(engine, keyset.encrypt(np.array(vals, dtype=np.uint64), seed=9))
    assert decode(p.desc, [int(x) for x in keyset.decrypt(out)]) == pat * 3


def test_split_long_overlapping_under_noise(engine, keyset):
    """The chunked (speculative) non-overlapping match selection, the last-set scan and the fan-in-4
    sign reduction on real ciphertexts: 40-char strings padded to 48, matches overlapping in every phase."""
    s = "ab" * 10 + "aabaabaab" + "xyzxyzxyzab"
    for pat, pp in (("aba", 5), ("ab", None), ("aab", 8)):
        assert run_gpu(engine, keyset, "split", [s, pat], [48, pp])[0] == s.split(pat)
        assert run_gpu(engine, keyset, "rsplit", [s, pat], [48, pp])[0] == s.rsplit(pat)[::-1]
    assert run_gpu(engine, keyset, "splitn", [s, "ab"], [48, 4], usizes=[3])[0] == s.split("ab", 2)
    assert run_gpu(engine, keyset, "split", [s, "b"], [48, None])[0] == s.split("b")
    f = s.find("xyzab")
    assert run_gpu(engine, keyset, "find", [s, "xyzab"], [48, 8])[0] == f
    assert run_gpu(engine, keyset, "rfind", [s, "ab"], [48, 8])[0] == s.rfind("ab")
    t = s[:30] + "b" + s[31:]
    for a, b in ((s, t), (t, s), (s, s), (s[:20], s)):
        assert run_gpu(engine, keyset, "lt", [a, b], [48, 48])[0] == (a < b)
        assert run_gpu(engine, keyset, "ge", [a, b], [48, 48])[0] == (a >= b)
