"""CPU suite: the oracle (CPU restatement of the tfhe-rs 0.5 KS-PBS path) against its golden vectors
and against the algebraic properties SURVEY.md Appendix A states.  No GPU needed."""
import hashlib
import os

import numpy as np
import pytest

from oracle import oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "hotpath_golden.npz")


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def test_signed_decomposer_properties():
    rng = np.random.default_rng(0)
    for base_log, level in ((3, 5), (23, 1), (4, 3)):
        xs = rng.integers(0, 2**64, 200, dtype=np.uint64)
        for x in xs:
            d = O.decompose(int(x), base_log, level)
            half = 1 << (base_log - 1)
            assert np.all(d >= -half) and np.all(d <= half)
            # digits come least significant first: level `level` first
            recon = sum(int(d[t]) << (64 - base_log * (level - t)) for t in range(level)) % 2**64
            r = 64 - base_log * level
            closest = ((int(x) >> r) + ((int(x) >> (r - 1)) & 1)) << r
            assert recon == closest % 2**64


def test_golden_decomposer_and_modswitch(gold, keyset):
    for x, kd, pd, ms in zip(gold["dec_x"], gold["ks_digits"], gold["pbs_digits"], gold["modswitch"]):
        assert np.array_equal(O.decompose(int(x), 3, 5), kd)
        assert np.array_equal(O.decompose(int(x), 23, 1), pd)
        assert keyset.modswitch(int(x)) == int(ms)
    assert keyset.modswitch(0xFFFFFFFFFFFFFFFF) == 4096  # 2N, equivalent to 0


def test_golden_keys_are_deterministic(gold, keyset):
    assert sha(keyset.ksk()) == str(gold["ksk_sha"][0])
    assert sha(keyset.bsk()) == str(gold["bsk_sha"][0])
    assert sha(keyset.glwe_sk()) == str(gold["glwe_sk_sha"][0])


def test_golden_keyswitch(gold, keyset):
    cts = gold["cts"]
    assert np.array_equal(keyset.encrypt(gold["msgs"], seed=77), cts)
    small = keyset.keyswitch(cts)
    assert np.array_equal(small, gold["small"])
    # keyswitch preserves the message: phase under the small key decodes to the same value
    assert np.array_equal(keyset.decode(keyset.phase_small(small)), gold["msgs"])


def test_golden_lut_and_pbs(gold, keyset):
    table = gold["table"]
    lut = keyset.build_lut(table)
    assert sha(lut) == str(gold["lut_body_sha"][0])
    out = keyset.ks_pbs_batch(gold["cts"], np.zeros(len(gold["cts"]), np.uint32), table[None, :])
    assert np.array_equal(keyset.decrypt(out), gold["pbs_decrypted"])
    assert np.array_equal(gold["pbs_decrypted"], table[gold["msgs"]])
    assert sha(out) == str(gold["pbs_out_sha"][0])  # same machine arithmetic (IEEE f64, no FMA contraction)


def test_lut_structure(keyset):
    """a1.1: boxes of N/16, first half box negated and rotated to the end; mask polynomial zero."""
    table = np.arange(16, dtype=np.uint64)
    acc = keyset.build_lut(table)
    N = keyset.N
    assert np.all(acc[:N] == 0)
    body = acc[N:]
    d = np.uint64(keyset.delta)
    assert np.all(body[:64] == table[0] * d)
    for i in range(1, 16):
        assert np.all(body[64 + 128 * (i - 1): 64 + 128 * i] == table[i] * d)
    with np.errstate(over="ignore"):
        assert np.all(body[N - 64:] == np.uint64(0) - table[0] * d)
    # closed form used by the CUDA kernel, rotation 0
    j = np.arange(N)
    v = (j + 64) % (2 * N)
    f = table[(v % N) >> 7] * d
    with np.errstate(over="ignore"):
        want = np.where(v < N, f, np.uint64(0) - f)
    assert np.array_equal(body, want)


def test_sample_extract_and_blind_rotate_trivial(keyset):
    """Blind rotation with an all-zero mask is a pure negacyclic rotation of the LUT by -b~."""
    table = (np.arange(16, dtype=np.uint64) * 3 + 1) % 16
    for m in (0, 1, 7, 15, 16 + 3):
        small = np.zeros(keyset.small, dtype=np.uint64)
        small[-1] = np.uint64(m * keyset.delta)
        acc = keyset.blind_rotate(small, keyset.build_lut(table))
        out = keyset.sample_extract(acc)
        dec = int(keyset.decrypt(out[None, :])[0])
        want = int(table[m]) if m < 16 else (32 - int(table[m - 16])) % 32
        assert dec == want


def test_pbs_noise_matches_estimate(keyset):
    rng = np.random.default_rng(1)
    msgs = rng.integers(0, 16, 32)
    cts = keyset.encrypt(msgs, seed=5)
    table = np.arange(16, dtype=np.uint64)
    out = keyset.ks_pbs_batch(cts, np.zeros(32, np.uint32), table[None, :])
    assert np.array_equal(keyset.decrypt(out), msgs)
    err = keyset.phase_error(out, msgs)
    assert np.std(err) < 4e-5 and np.abs(err).max() < 2.0**-10


def test_noise_golden_matches_live_oracle(keyset):
    """tests/golden/noise_golden.json (32 768 oracle KS-PBS, make_noise_golden.py) is the reference of
    the GPU noise tier; a small live sample must agree with it within its own sampling error."""
    import json
    from conftest import bench_tables
    gold = json.load(open(os.path.join(os.path.dirname(GOLD), "noise_golden.json")))
    assert gold["n_samples"] >= 30000 and abs(gold["mean"]) < 1e-6
    tables = bench_tables()
    m = 96
    rng = np.random.default_rng(21)
    msgs = rng.integers(0, 16, m)
    cts = keyset.encrypt(msgs, seed=21)
    lid = (np.arange(m) % 8).astype(np.uint32)
    out = keyset.ks_pbs_batch(cts, lid, tables)
    want = tables[lid, msgs]
    assert np.array_equal(keyset.decrypt(out), want)
    var = float(np.var(keyset.phase_error(out, want)))
    assert 0.55 * gold["variance"] < var < 1.6 * gold["variance"], (var, gold["variance"])


def test_fast_build_same_algorithm_different_rounding():
    """oracle/Makefile `fast` (-DORC_FAST -march=native: SoA radix-4 FFT without permutations, per-thread
    scratch) is the CPU arm bench.py times.  It is the same algorithm: integer tiers bit-identical to the
    portable build, PBS decrypt-exact with the same noise level, its external product within FFT rounding
    of the exact integer convolution -- but not bit-identical (summation order, FMA contraction)."""
    from test_cpu_oracle_first_principles import signed
    slow, fast = O.Keyset(seed=1), O.Keyset(seed=1, fast=True)
    assert fast.L.orc_is_fast_build() == 1 and slow.L.orc_is_fast_build() == 0
    assert np.array_equal(slow.ksk(), fast.ksk()) and np.array_equal(slow.bsk(), fast.bsk())
    rng = np.random.default_rng(3)
    msgs = rng.integers(0, 16, 24)
    cts = slow.encrypt(msgs, seed=9)
    assert np.array_equal(fast.encrypt(msgs, seed=9), cts)
    assert np.array_equal(fast.keyswitch(cts[:4]), slow.keyswitch(cts[:4]))
    table = (np.arange(16, dtype=np.uint64) * 7 + 3) % 16
    out = fast.ks_pbs_batch(cts, np.zeros(len(cts), np.uint32), table[None, :])
    assert np.array_equal(fast.decrypt(out), table[msgs])
    assert np.std(fast.phase_error(out, table[msgs])) < 4e-5
    # one CMUX (a~_5 = 1, everything else 0) through both builds: equal up to FFT rounding
    small = np.zeros(fast.small, dtype=np.uint64)
    small[5] = np.uint64(1) << np.uint64(52)
    lut = fast.build_lut(table)
    err = signed(fast.blind_rotate(small, lut) - slow.blind_rotate(small, lut))
    assert np.abs(err).max() < 2.0**-20
