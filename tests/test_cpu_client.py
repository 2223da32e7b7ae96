"""CPU: the product's client (libfhs_string: keygen / encrypt / decrypt, client.cpp) -- generator
known answer, OS-seeded keys differ between calls, encrypt/decrypt round trip, FheString::new errors."""
import ctypes

import numpy as np
import pytest

from fhe_string_b200.api import ClientKey
from fhe_string_b200.program import load_host_library


def test_chacha20_block_rfc8439_known_answer():
    L = load_host_library()
    L.fhs_chacha20_block.argtypes = [ctypes.c_void_p] * 3
    L.fhs_chacha20_block.restype = None
    key = np.frombuffer(bytes(range(32)), dtype=np.uint32).copy()
    cn = np.array([1, 0x09000000, 0x4A000000, 0], dtype=np.uint32)      # RFC 8439 2.3.2
    out = np.zeros(64, np.uint8)
    L.fhs_chacha20_block(key.ctypes.data, cn.ctypes.data, out.ctypes.data)
    assert out.tobytes().hex() == ("10f1e7e4d13b5915500fdd1fa32071c4c7d1f4c733c068030422aa9ac3d46c4e"
                                   "d2826446079faa0914c2d705d98b02a2b5129cd1de164eb9cbd083e8a2503c4e")


@pytest.fixture(scope="module")
def small_params():
    from fhe_string_b200.engine import default_params
    p = default_params()
    p.lwe_dim, p.poly_size = 64, 256          # client-side only: keygen in milliseconds
    return p


def test_os_seeded_keys_are_independent(small_params):
    a, b = ClientKey(l=15, seed=0, params=small_params), ClientKey(l=15, seed=0, params=small_params)
    assert not np.array_equal(a.ksk()[:4096], b.ksk()[:4096])
    msgs = np.arange(32) % 32
    ca, cb = a.encrypt_blocks(msgs), b.encrypt_blocks(msgs)
    assert np.array_equal(a.decrypt_blocks(ca), msgs) and np.array_equal(b.decrypt_blocks(cb), msgs)
    assert not np.array_equal(a.decrypt_blocks(cb), msgs)             # the other key does not decrypt
    # fresh masks per ciphertext and per call
    assert not np.array_equal(ca[0, :8], ca[16, :8]) and not np.array_equal(ca[0, :8], a.encrypt_blocks(msgs[:1])[0, :8])


def test_seeded_mode_is_reproducible(small_params):
    a, b = ClientKey(l=15, seed=7, params=small_params), ClientKey(l=15, seed=7, params=small_params)
    assert np.array_equal(a.ksk(), b.ksk()) and np.array_equal(a.bsk(), b.bsk())
    assert not np.array_equal(a.ksk()[:512], ClientKey(l=15, seed=8, params=small_params).ksk()[:512])


def test_masks_look_uniform_and_noise_is_small(small_params):
    k = ClientKey(l=15, seed=0, params=small_params)
    c = k.encrypt_blocks(np.zeros(256, dtype=np.uint64))
    bits = np.unpackbits(c[:, :-1].copy().view(np.uint8))
    assert abs(bits.mean() - 0.5) < 0.005
    # keyswitch-key bodies carry sigma_lwe noise: phase - plaintext is small but not zero
    assert len(np.unique(c[:, -1])) == 256


def test_string_layout_errors(small_params):
    k = ClientKey(l=15, seed=3, params=small_params)
    for s, l, msg in (("é", None, "string is not ascii"), ("a\0b", None, "string contains terminator character"),
                      ("a" * 16, None, "string length exceeds maximum length"),
                      ("ab", 16, "pad length exceeds maximum length"), ("abc", 2, "string length exceeds pad length")):
        with pytest.raises(ValueError) as e:
            k.encrypt_str(s, l)
        assert str(e.value) == msg                                    # mod.rs:59-71
    assert k.encrypt_str("abc", 5).decrypt(k) == "abc"


def test_tfhe_wire_layout_round_trip(small_params):
    """fhe_string_b200/tfhe_wire.py = the bincode field order INTEGRATION.md expects for the reference's serde
    types (recalled, unverified against tfhe-rs): self-consistent, size per block as documented, strict about
    trailing bytes so that a real file falsifies it in one call."""
    from fhe_string_b200 import tfhe_wire as W
    k = ClientKey(l=15, seed=5, params=small_params)
    s = k.encrypt_str("abc", 4)
    raw = W.dump_fhe_string(s.blocks)
    lwe = small_params.poly_size + 1
    assert len(raw) == 8 + 5 * (8 + 4 * (8 + 8 * lwe + 60))            # 60 = 24 (modulus, scalar bits) + 4 x 8 + 4 (pbs_order)
    back = W.parse_fhe_string(raw)
    assert np.array_equal(back, s.blocks)
    from fhe_string_b200.api import FheString
    assert FheString(back).decrypt(k) == "abc"
    with pytest.raises(W.WireError):
        W.parse_fhe_string(raw + b"\0")
    with pytest.raises(W.WireError):
        W.parse_fhe_string(raw[:-3])
    # the documented size for the real parameter set: 16 460 bytes per block
    assert 8 + 8 * 2049 + W.BLOCK_META.size == 16460
