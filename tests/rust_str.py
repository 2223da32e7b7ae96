"""Rust `std::str` semantics for ASCII strings, as the reference's tests assert them
(/root/reference/src/ciphertext/tests/*.rs compute every expected value with the std method of the same
name at test time).  Python's str methods differ in a few places (empty patterns, splitn counts, the
direction of rsplit, split_inclusive / split_terminator do not exist); this module is the one place
where those differences are written down.  Used by tests/golden/make_reference_cases.py and the tests."""

WS = "\t\n\x0b\x0c\r "   # char::is_whitespace restricted to ASCII = trim.rs:15-25 (9..13 or 32)


def _matches(s, p):
    """Start offsets of the non-overlapping matches of p in s, left to right (str::match_indices).
    The empty pattern matches at every char boundary, 0..=len."""
    if p == "":
        return list(range(len(s) + 1))
    out, i = [], 0
    while True:
        j = s.find(p, i)
        if j < 0:
            return out
        out.append(j)
        i = j + len(p)


def _rmatches(s, p):
    """Same, searched right to left (str::rmatch_indices), listed right to left."""
    if p == "":
        return list(range(len(s), -1, -1))
    out, e = [], len(s)
    while True:
        j = s.rfind(p, 0, e)
        if j < 0:
            return out
        out.append(j)
        e = j
    return out


def split(s, p):
    out, i = [], 0
    for j in _matches(s, p):
        out.append(s[i:j])
        i = j + len(p)
    out.append(s[i:])
    return out


def rsplit(s, p):
    out, e = [], len(s)
    for j in _rmatches(s, p):
        out.append(s[j + len(p):e])
        e = j
    out.append(s[:e])
    return out


def split_terminator(s, p):
    r = split(s, p)
    if r and r[-1] == "":
        r.pop()
    return r


def rsplit_terminator(s, p):
    r = rsplit(s, p)
    if r and r[0] == "":
        r.pop(0)
    return r


def split_inclusive(s, p):
    out, i = [], 0
    for j in _matches(s, p):
        out.append(s[i:j + len(p)])
        i = j + len(p)
    if i < len(s):
        out.append(s[i:])
    return out


def splitn(s, n, p):
    if n == 0:
        return []
    out, i = [], 0
    for j in _matches(s, p):
        if len(out) == n - 1:
            break
        out.append(s[i:j])
        i = j + len(p)
    out.append(s[i:])
    return out


def rsplitn(s, n, p):
    if n == 0:
        return []
    out, e = [], len(s)
    for j in _rmatches(s, p):
        if len(out) == n - 1:
            break
        out.append(s[j + len(p):e])
        e = j
    out.append(s[:e])
    return out


def split_once(s, p):
    j = s.find(p)
    return None if j < 0 else (s[:j], s[j + len(p):])


def rsplit_once(s, p):
    j = s.rfind(p)
    return None if j < 0 else (s[:j], s[j + len(p):])


def split_ascii_whitespace(s):
    return [x for x in "".join(" " if c in WS else c for c in s).split(" ") if x]


def find(s, p):
    j = s.find(p)
    return None if j < 0 else j


def rfind(s, p):
    j = s.rfind(p)
    return None if j < 0 else j


def strip_prefix(s, p):
    return s[len(p):] if s.startswith(p) else None


def strip_suffix(s, p):
    return s[:len(s) - len(p)] if s.endswith(p) else None


def replace(s, p, r):
    return replacen(s, p, r, len(s) + 2)


def replacen(s, p, r, n):
    out, i, k = [], 0, 0
    for j in _matches(s, p):
        if k == n:
            break
        out.append(s[i:j])
        out.append(r)
        i = j + len(p)
        k += 1
    out.append(s[i:])
    return "".join(out)


STD = {
    "is_empty": lambda a: a == "",
    "len": len,
    "eq": lambda a, b: a == b, "ne": lambda a, b: a != b,
    "lt": lambda a, b: a < b, "le": lambda a, b: a <= b, "gt": lambda a, b: a > b, "ge": lambda a, b: a >= b,
    "eq_ignore_ascii_case": lambda a, b: a.lower() == b.lower(),
    "to_lowercase": lambda a: a.lower(), "to_uppercase": lambda a: a.upper(),
    "trim": lambda a: a.strip(WS), "trim_start": lambda a: a.lstrip(WS), "trim_end": lambda a: a.rstrip(WS),
    "find": find, "rfind": rfind, "contains": lambda a, b: b in a,
    "starts_with": lambda a, b: a.startswith(b), "ends_with": lambda a, b: a.endswith(b),
    "strip_prefix": strip_prefix, "strip_suffix": strip_suffix,
    "split": split, "rsplit": rsplit, "split_terminator": split_terminator,
    "rsplit_terminator": rsplit_terminator, "split_inclusive": split_inclusive,
    "split_ascii_whitespace": split_ascii_whitespace, "split_once": split_once, "rsplit_once": rsplit_once,
    "splitn": lambda a, n, p: splitn(a, n, p), "rsplitn": lambda a, n, p: rsplitn(a, n, p),
    "replace": replace, "replacen": replacen,
    "add": lambda a, b: a + b, "repeat": lambda a, n: a * n,
    "insert": lambda a, i, b: a[:i] + b + a[i:],
}
