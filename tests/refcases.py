"""The reference's own TestCase tables (tests/golden/reference_cases.json, extracted from
/root/reference/src/ciphertext/tests/*.rs by tests/golden/make_reference_cases.py) as a flat list of
calls: one entry per (test fn, row, asserted method), with the operands encrypted the way that test
encrypts them."""
import json
import os

_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_cases.json")

UNARY = ("is_empty", "len", "to_lowercase", "to_uppercase", "trim", "trim_start", "trim_end", "split_ascii_whitespace")


def load():
    return json.load(open(_PATH))


def _tup(x):
    if isinstance(x, list):
        return [_tup(y) for y in x]
    return x


def calls():
    """-> list of dict(id, op, strs, pads, usizes, l, expected).  Pads follow each test's own
    encrypt_string calls: input/pattern/replace share `pad` (tests/split.rs:50-52, replace.rs:46-48),
    a/b have a_pad/b_pad (tests/compare.rs:113-114); `l` is result.len() for replace/replacen
    (tests/replace.rs:52,107-113) and the row's l for repeat (tests/insert.rs:103)."""
    out = []
    for t in load()["tests"]:
        for ri, r in enumerate(t["rows"]):
            for o in t["ops"]:
                op = o["op"]
                exp = r["expected"][op]
                if "input" in r:
                    first, fpad = r["input"], r["pad"]
                    second, spad = r.get("pattern"), r["pad"]
                else:
                    first, fpad = r["a"], r["a_pad"]
                    second, spad = r.get("b"), r.get("b_pad")
                c = dict(id=f"{t['file'].split('/')[-1]}:{o['line']}:{t['test']}[{ri}].{op}", op=op, usizes=[], l=None)
                if op in UNARY:
                    c.update(strs=[first], pads=[fpad])
                elif op in ("replace", "replacen"):
                    c.update(strs=[first, r["pattern"], r["replace"]], pads=[r["pad"]] * 3, l=len(exp))
                    if op == "replacen":
                        c["usizes"] = [r["n"]]
                elif op == "repeat":
                    c.update(strs=[first], pads=[fpad], usizes=[r["n"]], l=r["l"])
                else:
                    c.update(strs=[first, second], pads=[fpad, spad])
                    if op in ("splitn", "rsplitn"):
                        c["usizes"] = [r["n"]]
                if op in ("split_once", "rsplit_once") and exp is not None:
                    exp = tuple(exp)
                c["expected"] = exp
                out.append(c)
    return out
