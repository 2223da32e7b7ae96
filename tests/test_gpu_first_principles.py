"""GPU: the CUDA kernels against first principles (no recollection of tfhe-rs involved), at sample
counts the CPU oracle cannot reach in seconds.  Complements tests/test_cpu_oracle_first_principles.py."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
U = np.uint64


def _signed(x):
    return np.asarray(x, dtype=U).view(np.int64).astype(np.float64) / 2.0**64


def _synth(keyset, n, seed, space=32):
    """Valid big-key encryptions made with numpy (fast): uniform masks, Gaussian noise."""
    rng = np.random.default_rng(seed)
    msgs = rng.integers(0, space, n)
    s = keyset.glwe_sk().astype(bool)
    mask = rng.integers(0, 2**64, size=(n, keyset.big - 1), dtype=U)
    noise = np.rint(rng.normal(0.0, keyset.p.glwe_std * 2.0**64, n)).astype(np.int64).astype(U)
    with np.errstate(over="ignore"):
        body = mask[:, s].sum(axis=1, dtype=U) + msgs.astype(U) * U(keyset.delta) + noise
    return msgs, np.concatenate([mask, body[:, None]], axis=1)


def test_keyswitch_preserves_phase_1e4_samples(engine, keyset):
    """phase under the small key == phase under the big key within the analytic keyswitch noise, over
    10 240 ciphertexts; tcgen05 kernel, messages use the padding bit too."""
    p = keyset.p
    n = 10240
    msgs, cts = _synth(keyset, n, 123)
    small = engine.keyswitch(cts)
    with np.errstate(over="ignore"):
        err = _signed(keyset.phase_small(small) - keyset.phase(cts))
    kN = p.glwe_dim * p.poly_size
    sigma = (kN * p.ks_level * 66 / 12.0 * p.lwe_std**2 + (kN / 2.0) * 2.0**-30 / 12.0) ** 0.5
    assert np.abs(err).max() < 6.5 * sigma, (np.abs(err).max(), sigma)
    assert 0.93 * sigma < np.std(err) < 1.07 * sigma, (np.std(err), sigma)
    # the mean is NOT zero for a fixed key: the most significant digit of the balanced base-8 decomposition
    # is never -4 (no carry out of the top), so it averages +1/2 and the error carries the key-dependent
    # offset  1/2 * sum_i e_(i, top level)  ~ N(0, kN / 4 * sigma_lwe^2)  (1.6e-4 rms, 1 % of half a box)
    bias_sigma = 0.5 * p.lwe_std * kN**0.5
    assert abs(np.mean(err)) < 4 * bias_sigma + 4 * sigma / n**0.5, (np.mean(err), bias_sigma)
    assert np.array_equal(keyset.decode(keyset.phase_small(small)), msgs)


def test_pbs_output_is_lut_of_input_phase(engine, keyset):
    """KS-PBS from its definition: the output decrypts to f(m) for m < 16 and to -f(m - 16) for an input
    with the padding bit set (negacyclic extension), for random tables, on 2 048 ciphertexts."""
    rng = np.random.default_rng(8)
    tables = rng.integers(0, 16, size=(8, 16)).astype(U)
    n = 2048
    msgs, cts = _synth(keyset, n, 321)
    lut_id = rng.integers(0, 8, n).astype(np.uint32)
    out = engine.ks_pbs(cts, lut_id, tables)
    f = tables[lut_id, msgs % 16].astype(np.int64)
    want = np.where(msgs < 16, f, (32 - f) % 32)
    assert np.array_equal(keyset.decrypt(out).astype(np.int64), want)
