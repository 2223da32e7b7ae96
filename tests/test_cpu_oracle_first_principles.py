"""Independent checks of the oracle that do NOT rest on a recollection of tfhe-rs: every step of the
KS-PBS path is compared with the mathematical object it is defined to compute, in exact integer
arithmetic (numpy u64 wraps = arithmetic mod 2^64; Python ints where needed).

* external product (oracle/fhs_oracle.c add_external_product, f64 negacyclic FFT) vs the exact
  negacyclic convolution  sum_r digit_r(ct1) * GGSW_i[r][c]  mod (X^N + 1, 2^64);
* the same product decrypts to  s_i * phase(ct1)  under the GLWE key (GGSW definition);
* keyswitch: phase under the small key == phase under the big key within the analytic noise;
* modulus switch == round(x * 2N / 2^64); sample extract == coefficient 0 of the GLWE phase;
* one CMUX at a time, a whole blind rotation tracks  LUT * X^{-(b~ - sum a~_i s_i)}.
What these do not pin: choices where several conventions are self-consistent (KSK level order in
memory, decomposer tie rule, Fourier layout) -- those stay "unpinned vs tfhe-rs", DESIGN.md 2."""
import numpy as np
import pytest

U = np.uint64


def negacyclic_mul_u64(d, b):
    """(d * b) mod (X^N + 1) with coefficients mod 2^64.  d: int64 digits, b: uint64."""
    N = len(d)
    out = np.zeros(N, dtype=U)
    du = d.astype(np.int64).view(U)
    with np.errstate(over="ignore"):
        nb = U(0) - b
        ext = np.concatenate([nb, b])        # ext[N + j - t] = b[j - t] for j >= t, -b[N + j - t] for j < t
        for t in range(N):
            if du[t]:
                out += du[t] * ext[N - t: 2 * N - t]
    return out


def digit23(x):
    """One-level signed decomposition from its definition: the multiple of 2^41 closest to x (ties up),
    as a signed 23-bit number (balanced representative of x / 2^41 mod 2^23)."""
    x = np.asarray(x, dtype=U)
    q = ((x >> U(41)) + ((x >> U(40)) & U(1))) & U((1 << 23) - 1)
    q = q.astype(np.int64)
    return np.where(q > (1 << 22), q - (1 << 23), q)    # 2^22 itself stays positive (tfhe: digit in [-B/2, B/2])


def glwe_phase(ks, ct):
    """body - mask * S  (negacyclic), exact."""
    N = ks.N
    s = ks.glwe_sk().astype(np.int64)
    return ct[N:] - negacyclic_mul_u64(s, ct[:N])


def signed(x):
    return np.asarray(x, dtype=U).view(np.int64).astype(np.float64) / 2.0**64


def test_digit23_matches_oracle_decomposer(keyset):
    from oracle import oracle as O
    rng = np.random.default_rng(5)
    xs = np.concatenate([rng.integers(0, 2**64, 300, dtype=U),
                         np.array([0, 1 << 40, (1 << 40) - 1, (1 << 63), (1 << 63) + (1 << 40), 2**64 - 1,
                                   (1 << 63) - (1 << 40), 3 << 62], dtype=U)])
    want = digit23(xs)
    for x, w in zip(xs, want):
        got = int(O.decompose(int(x), 23, 1)[0])
        # the two representatives of the tie +-2^22 are the same element mod 2^23
        assert (got - int(w)) % (1 << 23) == 0 and abs(got) <= 1 << 22


@pytest.mark.parametrize("i", [0, 1, 371, 741])
def test_external_product_vs_exact_integer_convolution(keyset, i):
    ks = keyset
    N = ks.N
    rng = np.random.default_rng(100 + i)
    ct1 = rng.integers(0, 2**64, 2 * N, dtype=U)
    acc0 = rng.integers(0, 2**64, 2 * N, dtype=U)
    got = ks.external_product(i, ct1, acc0)
    bsk = ks.bsk().reshape(ks.n, 1, 2, 2, N)
    digs = [digit23(ct1[:N]), digit23(ct1[N:])]
    with np.errstate(over="ignore"):
        for c in range(2):
            exact = acc0[c * N:(c + 1) * N].copy()
            for r in range(2):
                exact += negacyclic_mul_u64(digs[r], bsk[i, 0, r, c])
            err = signed(got[c * N:(c + 1) * N] - exact)
            # f64 FFT rounding only: products reach 2^95 units, 53-bit mantissa => ~2^-26 of the torus
            # per coefficient (DESIGN.md 4 "Noise tier"); a wrong twist / sign / row order gives O(1).
            assert np.abs(err).max() < 2.0**-20, np.abs(err).max()
            assert np.std(err) < 2.0**-23, np.std(err)
            assert abs(np.mean(err)) < 2.0**-26


def test_external_product_is_multiplication_by_the_key_bit(keyset):
    """GGSW_i encrypts s_i: phase(GGSW_i (x) ct1) = s_i * phase(ct1) + noise, for both values of s_i."""
    ks = keyset
    N = ks.N
    sk = ks.lwe_sk()
    seen = set()
    rng = np.random.default_rng(9)
    for i in range(40):
        bit = int(sk[i])
        if bit in seen and len(seen) == 2:
            break
        seen.add(bit)
        ct1 = rng.integers(0, 2**64, 2 * N, dtype=U)
        out = ks.external_product(i, ct1, np.zeros(2 * N, dtype=U))
        with np.errstate(over="ignore"):
            err = signed(glwe_phase(ks, out) - U(bit) * glwe_phase(ks, ct1))
        # decomposition rounding (2^-24 per mask coefficient x ~1024 key bits) dominates: sigma ~ 2^-20.5
        assert np.abs(err).max() < 2.0**-15, (i, bit, np.abs(err).max())
        assert np.std(err) < 2.0**-18
    assert seen == {0, 1}


def test_keyswitch_preserves_phase_within_analytic_noise(keyset):
    ks = keyset
    p = ks.p
    rng = np.random.default_rng(17)
    m = 384
    msgs = rng.integers(0, 32, m)
    cts = ks.encrypt(msgs, seed=31)
    small = ks.keyswitch(cts)
    with np.errstate(over="ignore"):
        err = signed(ks.phase_small(small) - ks.phase(cts))
    kN = p.glwe_dim * p.poly_size
    # every KSK row is used with a balanced base-8 digit of a uniform 15-bit value: E[d^2] = (8^2 + 2) / 12;
    # rounding the mask to 15 bits adds 2^-30 / 12 per coefficient whose key bit is set (~ kN / 2 of them)
    var_rows = kN * p.ks_level * (64 + 2) / 12.0 * p.lwe_std**2
    var_round = (kN / 2.0) * 2.0**(-2 * 15) / 12.0
    sigma = (var_rows + var_round) ** 0.5
    assert np.abs(err).max() < 6 * sigma, (np.abs(err).max(), sigma)
    assert 0.75 * sigma < np.std(err) < 1.25 * sigma, (np.std(err), sigma)
    # key-dependent offset: the top digit of the balanced decomposition averages +1/2 (never -4), see
    # tests/test_gpu_first_principles.py
    assert abs(np.mean(err)) < 4 * 0.5 * p.lwe_std * kN**0.5 + 4 * sigma / m**0.5


def test_modswitch_is_rounding_to_2N(keyset):
    rng = np.random.default_rng(3)
    for x in list(rng.integers(0, 2**64, 200, dtype=U)) + [0, 2**64 - 1, 1 << 51, (1 << 51) - 1, 1 << 63]:
        x = int(x)
        want = (x * 4096 + (1 << 63)) >> 64          # round-half-up of x * 2N / 2^64, in [0, 2N]
        assert keyset.modswitch(x) == want


def test_sample_extract_is_coefficient_zero_of_the_phase(keyset):
    ks = keyset
    rng = np.random.default_rng(4)
    acc = rng.integers(0, 2**64, 2 * ks.N, dtype=U)
    lwe = ks.sample_extract(acc)
    assert int(ks.phase(lwe[None, :])[0]) == int(glwe_phase(ks, acc)[0])


def test_blind_rotation_tracks_the_predicted_rotation(keyset):
    """After the whole blind rotation the accumulator's phase is LUT * X^{-(b~ - sum a~_i s_i)} up to
    noise: the exponent is computed here from the key bits, the LUT rotated exactly."""
    ks = keyset
    N = ks.N
    table = (np.arange(16, dtype=U) * 5 + 2) % 16
    cts = ks.encrypt([11], seed=77)
    small = ks.keyswitch(cts)[0]
    acc = ks.blind_rotate(small, ks.build_lut(table))
    ms = np.array([ks.modswitch(int(x)) % (2 * N) for x in small], dtype=np.int64)
    e = int((ms[-1] - np.dot(ms[:-1], ks.lwe_sk().astype(np.int64))) % (2 * N))
    lut = ks.build_lut(table)[N:]
    with np.errstate(over="ignore"):
        ext = np.concatenate([lut, U(0) - lut])
        j = (np.arange(N) + e) % (2 * N)              # (LUT * X^{-e})[j] = LUT_ext[j + e]
        want = ext[j]
        err = signed(glwe_phase(ks, acc) - want)
    assert np.abs(err).max() < 2.0**-11, np.abs(err).max()
    assert np.std(err) < 6e-5
