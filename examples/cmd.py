#!/usr/bin/env python
"""Driver equivalent of the reference's examples/cmd/main.rs: same flags (main.rs:8-38), runs the
string operations on an encrypted input, prints Std vs Fhe result with timings and an Equal mark
(main.rs:517-546).  Fhe results come from the GPU (libfhs_string -> libfhs_cuda)."""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fhe_string_b200.api import generate_keys  # noqa: E402

WS = "\t\n\x0b\x0c\r "


def rsplit_rev(s, p, n=-1):
    return s.rsplit(p, n)[::-1] if p else ([""] + list(s) + [""])[::-1]


def split_fwd(s, p, n=-1):
    return s.split(p, n) if p else [""] + list(s) + [""]


def split_term(parts):
    return parts[:-1] if parts and parts[-1] == "" else parts


def split_incl(s, p):
    out, i = [], 0
    while p:
        j = s.find(p, i)
        if j < 0:
            break
        out.append(s[i:j + len(p)])
        i = j + len(p)
    if i < len(s):
        out.append(s[i:])
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--input", required=True)
    ap.add_argument("--pattern", required=True)
    ap.add_argument("--substitution", default="_")
    ap.add_argument("--pad", type=int, default=None)
    ap.add_argument("--n", type=int, default=3)
    ap.add_argument("--filter", default=None)
    ap.add_argument("--l", type=int, default=255)
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--seed", type=int, default=0)
    a = ap.parse_args()

    t0 = time.time()
    ck, sk = generate_keys(a.l, a.seed, a.device)
    print(f"Generated keys ({time.time() - t0:.2f}s); blocks per usize: {ck.num_blocks_usize}")
    s, p, sub, n = a.input, a.pattern, a.substitution, a.n
    es, ep, esub = ck.encrypt_str(s, a.pad), ck.encrypt_str(p, a.pad), ck.encrypt_str(sub, a.pad)
    en = ck.encrypt_usize(n)
    opt = lambda v: None if v < 0 else v  # noqa: E731
    L = min(a.l, ck.max_len)
    cases = [
        ("len", lambda: len(s), lambda: es.len(sk)),
        ("is_empty", lambda: s == "", lambda: es.is_empty(sk)),
        ("eq", lambda: s == p, lambda: es.eq(sk, ep)),
        ("ne", lambda: s != p, lambda: es.ne(sk, ep)),
        ("le", lambda: s <= p, lambda: es.le(sk, ep)),
        ("lt", lambda: s < p, lambda: es.lt(sk, ep)),
        ("ge", lambda: s >= p, lambda: es.ge(sk, ep)),
        ("gt", lambda: s > p, lambda: es.gt(sk, ep)),
        ("eq_ignore_ascii_case", lambda: s.lower() == p.lower(), lambda: es.eq_ignore_ascii_case(sk, ep)),
        ("contains", lambda: p in s, lambda: es.contains(sk, ep)),
        ("find", lambda: opt(s.find(p)), lambda: es.find(sk, ep)),
        ("rfind", lambda: opt(s.rfind(p)), lambda: es.rfind(sk, ep)),
        ("starts_with", lambda: s.startswith(p), lambda: es.starts_with(sk, ep)),
        ("ends_with", lambda: s.endswith(p), lambda: es.ends_with(sk, ep)),
        ("to_lowercase", lambda: s.lower(), lambda: es.to_lowercase(sk)),
        ("to_uppercase", lambda: s.upper(), lambda: es.to_uppercase(sk)),
        ("add", lambda: (s + p)[:L], lambda: es.add(sk, ep)),
        ("repeat", lambda: (s * n)[:L], lambda: es.repeat(sk, en, min(L, len(s * n)))),
        ("trim", lambda: s.strip(WS), lambda: es.trim(sk)),
        ("trim_start", lambda: s.lstrip(WS), lambda: es.trim_start(sk)),
        ("trim_end", lambda: s.rstrip(WS), lambda: es.trim_end(sk)),
        ("strip_prefix", lambda: s[len(p):] if s.startswith(p) else None, lambda: es.strip_prefix(sk, ep)),
        ("strip_suffix", lambda: s[:len(s) - len(p)] if s.endswith(p) else None, lambda: es.strip_suffix(sk, ep)),
        ("replace", lambda: s.replace(p, sub)[:L], lambda: es.replace(sk, ep, esub, min(L, len(s.replace(p, sub))))),
        ("replacen", lambda: s.replace(p, sub, n)[:L], lambda: es.replacen(sk, ep, esub, en, min(L, len(s.replace(p, sub, n))))),
        ("split", lambda: split_fwd(s, p), lambda: es.split(sk, ep)),
        ("rsplit", lambda: rsplit_rev(s, p), lambda: es.rsplit(sk, ep)),
        ("split_once", lambda: tuple(s.split(p, 1)) if p in s else None, lambda: es.split_once(sk, ep)),
        ("rsplit_once", lambda: tuple(s.rsplit(p, 1)) if p in s else None, lambda: es.rsplit_once(sk, ep)),
        ("splitn", lambda: [] if n == 0 else split_fwd(s, p, n - 1), lambda: es.splitn(sk, en, ep)),
        ("rsplitn", lambda: [] if n == 0 else rsplit_rev(s, p, n - 1), lambda: es.rsplitn(sk, en, ep)),
        ("split_inclusive", lambda: split_incl(s, p), lambda: es.split_inclusive(sk, ep)),
        ("split_terminator", lambda: split_term(split_fwd(s, p)), lambda: es.split_terminator(sk, ep)),
        ("rsplit_terminator", lambda: split_term(s.rsplit(p))[::-1] if p else None, lambda: es.rsplit_terminator(sk, ep)),
        ("split_ascii_whitespace", lambda: s.split(), lambda: es.split_ascii_whitespace(sk)),
    ]
    ok_all, t_all = True, time.time()
    for name, std, fhe in cases:
        if a.filter and a.filter not in name:
            continue
        if name == "rsplit_terminator" and not p:
            continue
        t = time.time(); want = std(); d_std = time.time() - t
        t = time.time(); enc = fhe(); d_fhe = time.time() - t
        got = enc.decrypt(ck)
        st = sk.last_stats
        ok = got == want
        ok_all &= ok
        print(f"{name}\n  Std: {want!r} ({d_std * 1e6:.0f}us)\n  Fhe: {got!r} ({d_fhe * 1e3:.1f}ms, "
              f"{st['pbs']} LUT evals in {st['levels']} levels)\n  Equal: {'✅' if ok else '❌'}")
    print(f"Total: {time.time() - t_all:.2f}s  all equal: {ok_all}")
    return 0 if ok_all else 1


if __name__ == "__main__":
    sys.exit(main())
