#!/usr/bin/env python
"""Extracts every `TestCase { .. }` row the reference's own unit tests hold
(/root/reference/src/ciphertext/tests/{mod,compare,convert,insert,replace,search,split,trim}.rs) into
tests/golden/reference_cases.json.  Run in the build container (the reference does not travel to the GPU
box); the JSON is committed.

Per `#[test] fn`: the file and line span of its table, the rows verbatim (strings, `None`/`Some(n)`
pads, integers), and the FheString methods the test calls on the encrypted operands (each with the line
of the call).  The reference computes the expected value at test time with the Rust `std::str` method of
the same name; `expected` here is that value from tests/rust_str.py (the Python statement of Rust's
semantics), stored so that the fixtures are self-contained and a change of the model shows up as a diff.
Nothing else of the reference is copied: no code, only the table data."""
import json
import os
import re
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import rust_str  # noqa: E402

REF = "/root/reference/src/ciphertext/tests"
FILES = ["mod.rs", "compare.rs", "convert.rs", "insert.rs", "replace.rs", "search.rs", "split.rs", "trim.rs"]


def parse_value(v):
    v = v.strip()
    if v == "None":
        return None
    m = re.fullmatch(r"Some\((\d+)\)", v)
    if m:
        return int(m.group(1))
    if re.fullmatch(r"\d+", v):
        return int(v)
    m = re.fullmatch(r'"((?:[^"\\]|\\.)*)"', v)
    if m:
        return bytes(m.group(1), "ascii").decode("unicode_escape")
    raise ValueError(f"unparsed value {v!r}")


def args_of(op, row):
    """Clear arguments of the std method, from the row's field names (the tests use two conventions:
    input/pattern/pad and a/a_pad/b/b_pad)."""
    first = row["input"] if "input" in row else row["a"]
    second = row.get("pattern", row.get("b"))
    if op in ("splitn", "rsplitn"):
        return [first, row["n"], second]
    if op == "replace":
        return [first, row["pattern"], row["replace"]]
    if op == "replacen":
        return [first, row["pattern"], row["replace"], row["n"]]
    if op == "repeat":
        return [first, row["n"]]
    if op in ("is_empty", "len", "to_lowercase", "to_uppercase", "trim", "trim_start", "trim_end",
              "split_ascii_whitespace"):
        return [first]
    return [first, second]


def main():
    tests = []
    n_rows = 0
    for fn in FILES:
        lines = open(os.path.join(REF, fn)).read().split("\n")
        starts = [i for i, l in enumerate(lines) if re.match(r"\s*fn \w+\(\)", l) and i and "#[test]" in lines[i - 1]]
        for k, s in enumerate(starts):
            e = starts[k + 1] if k + 1 < len(starts) else len(lines)
            name = re.match(r"\s*fn (\w+)\(\)", lines[s]).group(1)
            body = lines[s:e]
            rows, row, first_line, last_line, in_vec = [], None, None, None, False
            for off, l in enumerate(body):
                ln = s + off + 1
                if "let test_cases = vec![" in l:
                    in_vec = True
                    continue
                if not in_vec:
                    continue
                if re.match(r"\s*\];", l):
                    break
                if re.match(r"\s*TestCase \{", l):
                    row = {}
                    first_line = first_line or ln
                    continue
                if re.match(r"\s*\},?\s*$", l) and row is not None:
                    rows.append(row)
                    row = None
                    last_line = ln
                    continue
                m = re.match(r"\s*(\w+): (.*?),?\s*$", l)
                if m and row is not None:
                    row[m.group(1)] = parse_value(m.group(2))
            if not rows:
                continue
            text = "\n".join(body)
            ops = []
            for m in re.finditer(r"(\w+)_enc\s*\.\s*(\w+)\(\s*&server_key", text):
                line = s + text[:m.start()].count("\n") + 1
                ops.append({"op": m.group(2), "line": line})
            for r in rows:
                r["expected"] = {}
                for o in ops:
                    r["expected"][o["op"]] = rust_str.STD[o["op"]](*args_of(o["op"], r))
            n_rows += len(rows)
            tests.append({"file": f"src/ciphertext/tests/{fn}", "test": name, "rows_lines": [first_line, last_line],
                          "ops": ops, "rows": rows})
    out = {"source": "matthiasgeihs/fhe_string src/ciphertext/tests/*.rs (TestCase tables only)",
           "expected_by": "tests/rust_str.py (Rust std::str semantics; the reference evaluates std at test time)",
           "setup": "generate_keys_with_params(PARAM_MESSAGE_2_CARRY_2_KS_PBS, 255)  (tests/mod.rs:15-18)",
           "n_rows": n_rows, "tests": tests}
    path = os.path.join(HERE, "reference_cases.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1)
        f.write("\n")
    print(f"{n_rows} rows in {len(tests)} tests -> {path}")


if __name__ == "__main__":
    main()
