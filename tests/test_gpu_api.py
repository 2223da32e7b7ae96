"""GPU: the product's own client (keygen / encrypt / decrypt) + server API, no oracle involved."""
import subprocess
import sys
import os

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def keys():
    from fhe_string_b200.api import generate_keys
    return generate_keys(l=255, seed=11)


def test_api_roundtrip_and_ops(keys):
    ck, sk = keys
    s = ck.encrypt_str("Hello, B200!", 16)
    p = ck.encrypt_str("B2")
    assert s.decrypt(ck) == "Hello, B200!"
    assert s.contains(sk, p).decrypt(ck) is True
    assert s.find(sk, p).decrypt(ck) == 7
    assert s.to_uppercase(sk).decrypt(ck) == "HELLO, B200!"
    assert s.len(sk).decrypt(ck) == 12
    assert s.split(sk, ck.encrypt_str("l")).decrypt(ck) == "Hello, B200!".split("l")
    assert s.replace(sk, ck.encrypt_str("l"), ck.encrypt_str("LL"), 14).decrypt(ck) == "HeLLLLo, B200!"
    assert s.strip_prefix(sk, ck.encrypt_str("Hello")).decrypt(ck) == ", B200!"
    assert s.strip_prefix(sk, p).decrypt(ck) is None
    assert s.substr_from(sk, ck.encrypt_usize(7)).decrypt(ck) == "B200!"
    # insert.rs + FheStringSliceVector::len / get (split.rs:32-83)
    assert s.add(sk, p).decrypt(ck) == "Hello, B200!B2"
    assert p.repeat(sk, ck.encrypt_usize(3), 6).decrypt(ck) == "B2B2B2"
    sv = s.split(sk, ck.encrypt_str("l"))
    assert sv.len(sk).decrypt(ck) == 3
    assert sv.get(sk, ck.encrypt_usize(2)).decrypt(ck) == "o, B200!"
    assert sv.get(sk, ck.encrypt_usize(5)).decrypt(ck) is None


def test_cmd_driver_all_equal():
    """BASELINE configs[0]: examples/cmd --input " A bcbc " --pattern "bc"."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "cmd.py"), "--input", " A bcbc ",
                        "--pattern", "bc", "--seed", "3"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert "❌" not in r.stdout and "all equal: True" in r.stdout


def test_config4_256_char_ops_full_size():
    """BASELINE configs[3] at full size: keys for l = 256 (5 usize blocks), a 200-char string of words
    padded to 256 (257 FheAsciiChar = 1 028 blocks); split(" "), replace(2-char pattern, "c", l=256),
    to_uppercase / to_lowercase decrypt to the Python str answers; round trip and idempotence hold."""
    import numpy as np
    from fhe_string_b200.api import generate_keys
    ck, sk = generate_keys(l=256, seed=5)
    assert ck.num_blocks_usize == 5
    rng = np.random.default_rng(11)
    words = ["".join(chr(int(c)) for c in rng.integers(97, 123, rng.integers(2, 7))) for _ in range(80)]
    s = " ".join(words)[:200]
    es = ck.encrypt_str(s, 256)
    assert es.decrypt(ck) == s
    up = es.to_uppercase(sk)
    assert up.decrypt(ck) == s.upper()
    assert up.to_uppercase(sk).decrypt(ck) == s.upper()              # idempotent
    assert up.to_lowercase(sk).decrypt(ck) == s                      # round trip (input is lowercase)
    assert es.split(sk, ck.encrypt_str(" ")).decrypt(ck) == s.split(" ")
    pat = s[50:52]
    assert es.replace(sk, ck.encrypt_str(pat), ck.encrypt_str("c"), 256).decrypt(ck) == s.replace(pat, "c")
    assert es.len(sk).decrypt(ck) == 200
    sk.engine.close()


def test_resident_results_chain_without_host_copies(keys):
    """Results stay in HBM (device-resident arrays) and feed the next op; the host copy is made at the first
    `.blocks` / decrypt.  Same answers as the host-array mode; cached execution state on the second call."""
    from fhe_string_b200.api import ServerKey
    ck, sk = keys
    s = ck.encrypt_str("Hello, B200!", 16)
    up = s.to_uppercase(sk)
    assert up.is_resident() and up._host is None                       # nothing downloaded yet
    st1 = dict(sk.last_stats)
    hit = up.contains(sk, ck.encrypt_str("B200"))                      # operand straight from device memory
    assert hit.decrypt(ck) is True
    assert up.decrypt(ck) == "HELLO, B200!"
    assert s.to_uppercase(sk).decrypt(ck) == "HELLO, B200!"
    st2 = sk.last_stats
    assert st2["cached"] and st2["plan_ms"] < st1["plan_ms"] + 0.5 and st2["device_ms"] > 0
    assert st2["levels"] == st1["levels"] and st2["pbs"] == st1["pbs"]
    # split result: 1 + 3 x entries slices of one device array, chained into slices_get
    sv = s.split(sk, ck.encrypt_str("l"))
    assert sv.get(sk, ck.encrypt_usize(1)).decrypt(ck) == ""
    assert sv.decrypt(ck) == "Hello, B200!".split("l")
    # host-array mode gives the same answers
    hk = ServerKey(ck, engine=sk.engine, resident=False)
    up2 = s.to_uppercase(hk)
    assert not up2.is_resident() and up2.decrypt(ck) == "HELLO, B200!"
    assert up2.contains(hk, ck.encrypt_str("B200")).decrypt(ck) is True
    n, nbytes = sk.engine.cache_stats()
    assert n >= 3 and nbytes > 0


def test_execution_cache_follows_program_and_budget(keys):
    import gc
    from fhe_string_b200.program import Program
    ck, sk = keys
    eng = sk.engine
    n0, _ = eng.cache_stats()
    p = Program("to_lowercase", [5], 0, (), 4, 4)
    cts = ck.encrypt_str("AbCd").blocks
    out, st = p.run(eng, cts)
    assert "".join(chr(c) for c in ck._radix(out, 4) if c) == "abcd"
    assert eng.cache_stats()[0] == n0 + 1
    out2, st2 = p.run(eng, cts)
    assert st2["ms"] < st["ms"] + 1.0
    del p
    gc.collect()
    assert eng.cache_stats()[0] == n0                                   # fhs_prog_destroy dropped its entry
    eng.cache_limit(1)                                                  # 1 byte: everything but the newest goes
    assert sk.program("is_empty", [3]) is not None
    ck.encrypt_str("ab").is_empty(sk)
    assert eng.cache_stats()[0] <= 1
    eng.cache_limit(16 << 30)


def test_bad_lut_id_is_rejected(keys):
    import numpy as np
    ck, sk = keys
    cts = ck.encrypt_blocks([1, 2, 3])
    tables = np.arange(32, dtype=np.uint64).reshape(2, 16) % 16
    with pytest.raises(Exception) as e:
        lut = np.array([0, 1, 2], dtype=np.uint32)                      # 2 is out of range
        out = np.empty_like(cts)
        sk.engine._check(sk.engine.L.fhs_ks_pbs_batch(sk.engine.h, cts.ctypes.data, lut.ctypes.data,
                                                       tables.ctypes.data, 2, 3, out.ctypes.data))
    assert "lut_id" in str(e.value)


def test_remaining_public_surface(keys):
    """mod.rs:96,198,217,240,426-437, compare.rs:119, convert.rs:14,22, trim.rs:15, split.rs:86,229 on real
    ciphertexts: new_trivial / char_at / substr_end / substr_eq / FheAsciiChar predicates / FheUsize
    constructors / FheStringSliceVector::truncate + reverse."""
    from fhe_string_b200.api import FheString, FheUsize
    ck, sk = keys
    text = "Hello, B200!"
    s = ck.encrypt_str(text, 16)
    c = s.char_at(sk, ck.encrypt_usize(7))
    assert chr(c.decrypt(ck)) == "B"
    assert c.is_uppercase(sk).decrypt(ck) is True and c.is_lowercase(sk).decrypt(ck) is False
    assert s.char_at(sk, ck.encrypt_usize(5)).is_whitespace(sk).decrypt(ck) is False
    assert s.char_at(sk, ck.encrypt_usize(6)).is_whitespace(sk).decrypt(ck) is True
    assert s.char_at(sk, ck.encrypt_usize(14)).decrypt(ck) == 0                   # out of bounds -> 0
    assert s.substr_end(sk, ck.encrypt_usize(7), ck.encrypt_usize(11)).decrypt(ck) == "B200"
    assert s.substr_eq(sk, 7, ck.encrypt_str("B2")).decrypt(ck) is True
    assert s.substr_eq(sk, 6, ck.encrypt_str("B2")).decrypt(ck) is False
    # trivial ciphertexts made by the server key are accepted by every op (mod.rs:96, :426)
    t = FheString.new_trivial(sk, "B200")
    assert t.decrypt(ck) == "B200"
    assert s.contains(sk, t).decrypt(ck) is True
    assert s.find(sk, t).decrypt(ck) == 7
    assert s.substr_from(sk, FheUsize.new_trivial(sk, 7)).decrypt(ck) == "B200!"
    assert FheUsize.new_from_bool(sk, s.contains(sk, t)).decrypt(ck) == 1
    assert FheUsize.max(sk) == 255 == FheUsize.max(ck) == FheString.max_len_with_key(sk)
    with pytest.raises(ValueError) as e:
        FheString.new_trivial(sk, "x" * 256)
    assert str(e.value) == "string length exceeds maximum length"
    sv = s.split(sk, ck.encrypt_str("l"))
    sv.truncate(sk, ck.encrypt_usize(2))
    assert sv.decrypt(ck) == text.split("l")[:2]
    sv.reverse()
    assert sv.decrypt(ck) == text.split("l")[:2][::-1]
