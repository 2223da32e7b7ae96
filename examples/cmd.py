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

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import rust_str as R  # noqa: E402  (Rust std::str semantics: what the reference prints as "Std")


def rust_debug(v):
    """{:?} of the Rust value the reference prints (main.rs:534-535): "str", true, Some(3), ["a", "b"], ("a", "b")."""
    if v is None:
        return "None"
    if isinstance(v, bool):
        return "true" if v else "false"
    if isinstance(v, str):
        return '"' + v.replace("\\", "\\\\").replace('"', '\\"') + '"'
    if isinstance(v, (list, tuple)):
        body = ", ".join(rust_debug(x) for x in v)
        return "[" + body + "]" if isinstance(v, list) else "(" + body + ")"
    return str(v)


def some(v, wrap):
    """Option<T> results: the reference prints Some(..) / None."""
    return v if not wrap or v is None else _Some(v)


class _Some:
    def __init__(self, v):
        self.v = v

    def __eq__(self, o):
        return isinstance(o, _Some) and o.v == self.v

    def __repr__(self):
        return "Some(" + rust_debug(self.v) + ")"


def dur(sec):
    """Rust Duration {:?}: ns / µs / ms / s with the unit that keeps the integer part non-zero."""
    for unit, k in (("s", 1.0), ("ms", 1e-3), ("µs", 1e-6), ("ns", 1e-9)):
        if sec >= k or unit == "ns":
            return f"{sec / k:.6g}{unit}"


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

    print("Generating keys...")
    ck, sk = generate_keys(a.l, a.seed, a.device)
    print("Done.")
    s, p, sub, n = a.input, a.pattern, a.substitution, a.n
    print("Encrypting input...")
    es = ck.encrypt_str(s, a.pad)
    print("Done.\nEncrypting pattern...")
    ep = ck.encrypt_str(p, a.pad)
    print("Done.\nEncrypting substitution...")
    esub = ck.encrypt_str(sub, a.pad)
    print("Done.\nEncrypting n...")
    en = ck.encrypt_usize(n)
    print("Done.")
    L = min(a.l, ck.max_len)
    q = rust_debug
    # (name as main.rs prints it, std result, fhe call, result is an Option)  -- order of main.rs:91-508
    cases = [
        (f"{q(s)}.len()", lambda: len(s), lambda: es.len(sk), False),
        (f"{q(s)}.contains({q(p)})", lambda: p in s, lambda: es.contains(sk, ep), False),
        (f"{q(s)}.starts_with({q(p)})", lambda: s.startswith(p), lambda: es.starts_with(sk, ep), False),
        (f"{q(s)}.ends_with({q(p)})", lambda: s.endswith(p), lambda: es.ends_with(sk, ep), False),
        (f"{q(s)}.find({q(p)})", lambda: R.find(s, p), lambda: es.find(sk, ep), True),
        (f"{q(s)}.rfind({q(p)})", lambda: R.rfind(s, p), lambda: es.rfind(sk, ep), True),
        (f"{q(s)}.eq({q(p)})", lambda: s == p, lambda: es.eq(sk, ep), False),
        (f"{q(s)}.le({q(p)})", lambda: s <= p, lambda: es.le(sk, ep), False),
        (f"{q(s)}.lt({q(p)})", lambda: s < p, lambda: es.lt(sk, ep), False),
        (f"{q(s)}.ge({q(p)})", lambda: s >= p, lambda: es.ge(sk, ep), False),
        (f"{q(s)}.gt({q(p)})", lambda: s > p, lambda: es.gt(sk, ep), False),
        (f"{q(s)}.ne({q(p)})", lambda: s != p, lambda: es.ne(sk, ep), False),
        (f"{q(s)}.eq_ignore_ascii_case({q(p)})", lambda: s.lower() == p.lower(), lambda: es.eq_ignore_ascii_case(sk, ep), False),
        (f"{q(s)}.is_empty()", lambda: s == "", lambda: es.is_empty(sk), False),
        (f"{q(s)}.add({q(p)})", lambda: (s + p)[:L], lambda: es.add(sk, ep), False),
        (f"{q(s)}.repeat({n})", lambda: (s * n)[:L], lambda: es.repeat(sk, en, min(L, len(s * n))), False),
        (f"{q(s)}.replace({q(p)}, {q(sub)})", lambda: R.replace(s, p, sub)[:L],
         lambda: es.replace(sk, ep, esub, min(L, len(R.replace(s, p, sub)))), False),
        (f"{q(s)}.replacen({q(p)}, {q(sub)}, {n})", lambda: R.replacen(s, p, sub, n)[:L],
         lambda: es.replacen(sk, ep, esub, en, min(L, len(R.replacen(s, p, sub, n)))), False),
        (f"{q(s)}.split({q(p)})", lambda: R.split(s, p), lambda: es.split(sk, ep), False),
        (f"{q(s)}.rsplit({q(p)})", lambda: R.rsplit(s, p), lambda: es.rsplit(sk, ep), False),
        (f"{q(s)}.split_once({q(p)})", lambda: R.split_once(s, p), lambda: es.split_once(sk, ep), True),
        (f"{q(s)}.rsplit_once({q(p)})", lambda: R.rsplit_once(s, p), lambda: es.rsplit_once(sk, ep), True),
        (f"{q(s)}.splitn({n}, {q(p)})", lambda: R.splitn(s, n, p), lambda: es.splitn(sk, en, ep), False),
        (f"{q(s)}.rsplitn({n}, {q(p)})", lambda: R.rsplitn(s, n, p), lambda: es.rsplitn(sk, en, ep), False),
        (f"{q(s)}.split_terminator({q(p)})", lambda: R.split_terminator(s, p), lambda: es.split_terminator(sk, ep), False),
        (f"{q(s)}.rsplit_terminator({q(p)})", lambda: R.rsplit_terminator(s, p), lambda: es.rsplit_terminator(sk, ep), False),
        (f"{q(s)}.split_inclusive({q(p)})", lambda: R.split_inclusive(s, p), lambda: es.split_inclusive(sk, ep), False),
        (f"{q(s)}.split_ascii_whitespace()", lambda: R.split_ascii_whitespace(s), lambda: es.split_ascii_whitespace(sk), False),
        (f"{q(s)}.trim()", lambda: s.strip(R.WS), lambda: es.trim(sk), False),
        (f"{q(s)}.trim_start()", lambda: s.lstrip(R.WS), lambda: es.trim_start(sk), False),
        (f"{q(s)}.trim_end()", lambda: s.rstrip(R.WS), lambda: es.trim_end(sk), False),
        (f"{q(s)}.strip_prefix({q(p)})", lambda: R.strip_prefix(s, p), lambda: es.strip_prefix(sk, ep), True),
        (f"{q(s)}.strip_suffix({q(p)})", lambda: R.strip_suffix(s, p), lambda: es.strip_suffix(sk, ep), True),
        (f"{q(s)}.to_lowercase()", lambda: s.lower(), lambda: es.to_lowercase(sk), False),
        (f"{q(s)}.to_uppercase()", lambda: s.upper(), lambda: es.to_uppercase(sk), False),
    ]
    ok_all, t_all = True, time.time()
    for name, std, fhe, is_opt in cases:
        if a.filter and a.filter not in name:
            continue
        if ".rsplit_terminator(" in name and not p:
            continue                                  # the reference's empty-pattern rsplit_terminator is unspecified
        t = time.time(); want = std(); d_std = time.time() - t
        t = time.time(); got = fhe().decrypt(ck); d_fhe = time.time() - t      # like main.rs: op + decrypt
        if isinstance(want, tuple) or isinstance(got, tuple):
            want, got = (tuple(want) if want is not None else None), (tuple(got) if got is not None else None)
        want, got = some(want, is_opt), some(got, is_opt)
        st = sk.last_stats
        ok = got == want
        ok_all &= ok
        print(f"\n{name}")
        print(f"Std: {rust_debug(want) if not isinstance(want, _Some) else repr(want)} ({dur(d_std)})")
        print(f"Fhe: {rust_debug(got) if not isinstance(got, _Some) else repr(got)} ({dur(d_fhe)}; "
              f"{st['pbs']} LUT evals, {st['levels']} levels, {st['device_ms']:.2f} ms on the GPU)")
        print(f"Equal: {'✅' if ok else '❌'}")
    print(f"\nDuration (total): {dur(time.time() - t_all)}")
    print(f"all equal: {ok_all}")
    return 0 if ok_all else 1


if __name__ == "__main__":
    sys.exit(main())
