"""Tier 1 parity on the vectors the REFERENCE holds: every TestCase row of
/root/reference/src/ciphertext/tests/*.rs (105 rows, 24 test fns; committed as
tests/golden/reference_cases.json) through the level program of the method that test asserts, evaluated
in the clear, against the std::str answer."""
import pytest

import refcases
import rust_str
from test_cpu_strings import run

CALLS = refcases.calls()


def test_fixture_is_complete():
    d = refcases.load()
    assert d["n_rows"] == 105 and sum(len(t["rows"]) for t in d["tests"]) == 105
    assert len(d["tests"]) == 24
    ops = {o["op"] for t in d["tests"] for o in t["ops"]}
    assert len(ops) == 35, sorted(ops)
    # the stored expectations are what the Python statement of Rust's semantics gives today
    for c in CALLS:
        args = [c["strs"][0]] + ([c["usizes"][0]] if c["op"] in ("splitn", "rsplitn") else []) + c["strs"][1:] \
            + (c["usizes"] if c["op"] in ("replacen", "repeat") else [])
        got = rust_str.STD[c["op"]](*args)
        assert (tuple(got) if isinstance(c["expected"], tuple) else got) == c["expected"], c["id"]


def test_pinned_edge_cases_of_rust_std():
    """Rows whose std answer differs from Python's str method of the same name."""
    assert rust_str.split("abc", "") == ["", "a", "b", "c", ""]
    assert rust_str.split("xxx", "xx") == ["", "x"]
    assert rust_str.rsplit("xxx", "xx") == ["", "x"]
    assert rust_str.rsplit("axbxc", "x") == ["c", "b", "a"]
    assert rust_str.split_terminator("axbx", "x") == ["a", "b"]
    assert rust_str.split_inclusive("axbx", "x") == ["ax", "bx"]
    assert rust_str.splitn("axbxc", 2, "x") == ["a", "bxc"]
    assert rust_str.rsplitn("axbxc", 2, "x") == ["c", "axb"]
    assert rust_str.splitn("axbxc", 0, "x") == []


@pytest.mark.parametrize("c", CALLS, ids=[c["id"] for c in CALLS])
def test_reference_row_clear(c):
    scalars = () if c["l"] is None else (c["l"],)
    got = run(c["op"], c["strs"], c["pads"], usizes=c["usizes"], scalars=scalars)
    assert got == c["expected"]
