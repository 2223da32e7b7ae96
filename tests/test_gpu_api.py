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
