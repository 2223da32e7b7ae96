"""GPU parity, tier 1 of SURVEY.md 8(c): the table-driven tests of tests/test_cpu_strings.py (rows in
the style of the reference's src/ciphertext/tests/*.rs, expectations = Rust std::str semantics) run
again on REAL ciphertexts: inputs encrypted with the oracle's client key, every level evaluated by
libfhs_cuda on the B200, outputs decrypted.  The test bodies are the CPU suite's own functions; only
their `run` (clear evaluation of the level program) is swapped for the encrypted one."""
import numpy as np
import pytest

import test_cpu_strings as T
from fhe_string_b200.program import Program
from strutil import decode, enc_chars, to_blocks, usize_blocks

pytestmark = pytest.mark.gpu

_programs = {}


@pytest.fixture
def encrypted_run(engine, keyset, monkeypatch):
    counter = [0]

    def run(op, strs, pads=None, usizes=(), scalars=(), nbu=4):
        pads = pads or [None] * len(strs)
        chars = [enc_chars(s, p) for s, p in zip(strs, pads)]
        key = (op, tuple(len(c) for c in chars), len(usizes), tuple(scalars), nbu)
        if key not in _programs:
            _programs[key] = Program(op, [len(c) for c in chars], len(usizes), scalars, 4, nbu)
        p = _programs[key]
        vals = [b for c in chars for b in to_blocks(c)] + [b for u in usizes for b in usize_blocks(u, nbu)]
        counter[0] += 1
        cts = keyset.encrypt(np.array(vals, dtype=np.uint64), seed=7000 + counter[0])
        out, stats = p.run(engine, cts)
        dec = [int(x) for x in keyset.decrypt(out)]
        assert all(0 <= v < 16 for v in dec), dec
        assert stats["pbs"] == p.pbs_count and stats["levels"] == p.depth
        return decode(p.desc, dec, nbu)

    monkeypatch.setattr(T, "run", run)
    return run


@pytest.mark.parametrize("a,b,pa,pb", T.PAIRS)
def test_compare_ops_encrypted(encrypted_run, a, b, pa, pb):
    T.test_compare_ops(a, b, pa, pb)


@pytest.mark.parametrize("a,b,pa,pb", T.PAIRS)
def test_search_ops_encrypted(encrypted_run, a, b, pa, pb):
    T.test_search_ops(a, b, pa, pb)


@pytest.mark.parametrize("s,pad", [("", None), ("a", None), ("abc", 5), ("  a b  ", None), ("\t\n a\r", 8),
                                   ("   ", None), ("aZ@[`{z", None), ("Hello, World! 09", 20), (" A bcbc ", None)])
def test_unary_ops_encrypted(encrypted_run, s, pad):
    T.test_unary_ops(s, pad)


def test_substr_encrypted(encrypted_run):
    T.test_substr()


@pytest.mark.parametrize("s,p,ps,pp", T.SPLIT_ROWS)
def test_split_family_encrypted(encrypted_run, s, p, ps, pp):
    T.test_split_family(s, p, ps, pp)


@pytest.mark.parametrize("s,pad", [("a b  c", None), ("  lead and trail \t", 20), ("", None), ("   ", 4), ("one", 5),
                                   ("a\tb\nc\rd", None)])
def test_split_ascii_whitespace_encrypted(encrypted_run, s, pad):
    T.test_split_ascii_whitespace(s, pad)


@pytest.mark.parametrize("a,p,s,pa,pp,ps", T.REPLACE_ROWS)
def test_replace_encrypted(encrypted_run, a, p, s, pa, pp, ps):
    T.test_replace(a, p, s, pa, pp, ps)


@pytest.mark.parametrize("a,b,pa,pb", [("abc", "xy", None, None), ("", "", None, None), ("a", "", None, None),
                                       ("", "b", None, None), ("abc", "xy", 5, 4), ("hello", "world", 7, None)])
def test_add_and_insert_encrypted(encrypted_run, a, b, pa, pb):
    T.test_add_and_insert(a, b, pa, pb)


def test_repeat_encrypted(encrypted_run):
    for a, pa, n, l in (("ab", None, 3, 7), ("abc", 4, 2, 4), ("", 3, 5, 4), ("x", None, 8, 16), ("ab", None, 0, 4)):
        assert encrypted_run("repeat", [a], [pa], usizes=[n], scalars=(l,)) == (a * n)[:l]


def test_five_block_usize_long_string_encrypted(encrypted_run):
    T.test_five_block_usize_long_string()


@pytest.mark.parametrize("s,p,ps,pp", T.LONG_OVERLAP)
def test_split_replace_long_overlapping_encrypted(encrypted_run, s, p, ps, pp):
    T.test_split_replace_long_overlapping(s, p, ps, pp)


def test_random_against_python_str_encrypted(encrypted_run):
    T.test_random_against_python_str()


def test_split_random_two_letter_alphabet_encrypted(encrypted_run):
    T.test_split_random_two_letter_alphabet()


def test_hypothesis_generated_ascii_strings_encrypted(encrypted_run):
    T.test_hypothesis_generated_ascii_strings()


def test_ordering_long_strings_encrypted(encrypted_run):
    T.test_ordering_long_strings_first_difference_everywhere()
