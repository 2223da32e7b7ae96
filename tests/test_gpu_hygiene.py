"""GPU hygiene: compute-sanitizer is closed on this pool, so races and protocol bugs
are hunted with what the kernels promise instead -- no atomics and no order-dependent reductions, hence
BIT-IDENTICAL outputs for the same input however often and in whatever launch mix it is run -- plus a
stress loop over random sizes, and direct tests of the modulus switch / LUT box edges
through all three blind-rotation kernels."""
import numpy as np
import pytest

from conftest import bench_tables

pytestmark = pytest.mark.gpu
U = np.uint64


def _enc(keyset, n, seed):
    rng = np.random.default_rng(seed)
    msgs = rng.integers(0, 16, n)
    return msgs, keyset.encrypt(msgs, seed=seed)


def test_stress_400_random_size_launches_pair_and_wide(engine, keyset):
    """400 back-to-back device-buffer launches of random sizes 1..296 (pair kernel <= 74: 2-CTA clusters +
    st.async to the peer's shared memory; wide kernel <= 296), no host synchronisation between them; a hang
    or a wrong result here is a protocol bug (mbarrier phase, double buffering, TMEM allocation)."""
    import torch
    tables = bench_tables()
    B = 296
    msgs, cts = _enc(keyset, B, 77)
    lut = (np.arange(B) % 8).astype(np.uint32)
    d_in = torch.from_numpy(cts.view(np.int64)).cuda()
    d_lut = torch.from_numpy(lut.view(np.int32)).cuda()
    d_tab = torch.from_numpy(tables.view(np.int64).copy()).cuda()
    d_out = torch.zeros_like(d_in)
    rng = np.random.default_rng(5)
    want = tables[lut, msgs]
    for it in range(400):
        b = int(rng.integers(1, 75)) if it % 3 else int(rng.integers(75, 297))
        engine.ks_pbs_dev(d_in.data_ptr(), d_lut.data_ptr(), d_tab.data_ptr(), len(tables), b, d_out.data_ptr())
        if it % 40 == 39:
            engine.synchronize()
            got = d_out.cpu().numpy().view(U)[:b]
            assert np.array_equal(keyset.decrypt(got), want[:b]), (it, b)
    engine.synchronize()


@pytest.mark.parametrize("n", [1, 37, 74, 148, 296, 592, 700])
def test_repeated_launches_are_bit_identical(engine, keyset, n):
    """Same input, five launches (with other sizes in between): every output bit equal.  The kernels have
    no atomics and a fixed summation order per ciphertext, so ANY difference is a race (shared-memory
    tile reuse, TMEM double buffering, DSMEM receive buffers, the stagger token)."""
    tables = bench_tables()
    msgs, cts = _enc(keyset, n, 900 + n)
    lut = ((np.arange(n) * 3) % 8).astype(np.uint32)
    first = engine.ks_pbs(cts, lut, tables)
    assert np.array_equal(keyset.decrypt(first), tables[lut, msgs])
    other = keyset.trivial(np.arange(50) % 16)
    for k in range(4):
        engine.ks_pbs(other[: 10 + 13 * k], np.zeros(10 + 13 * k, np.uint32), tables)   # interleave another wave shape
        again = engine.ks_pbs(cts, lut, tables)
        assert np.array_equal(again, first), f"launch {k + 2} differs from the first in {(again != first).sum()} words"


def test_batch_position_does_not_change_a_ciphertext(engine, keyset):
    """The same ciphertext evaluated alone (pair kernel), inside a wide wave and inside a 4-slot round:
    the three kernels use different FFT factorizations, so bits differ, but the PHASE agrees within FFT
    rounding and the decryption is the same; inside one kernel the slot it lands in changes nothing."""
    tables = bench_tables()
    msgs, cts = _enc(keyset, 700, 31)
    lut = np.full(700, 6, dtype=np.uint32)
    big = engine.ks_pbs(cts, lut, tables)                 # 592 in 4-slot rounds + 108 wide
    for idx in (0, 5, 591):
        solo = engine.ks_pbs(cts[idx:idx + 1], lut[:1], tables)      # pair kernel
        e1 = keyset.phase_error(solo, tables[6, msgs[idx:idx + 1]])
        e2 = keyset.phase_error(big[idx:idx + 1], tables[6, msgs[idx:idx + 1]])
        assert abs(float(e1[0] - e2[0])) < 2e-4
    # same kernel, different slot: rotate the batch by one 4-slot round
    rolled = np.roll(cts[:592], 148, axis=0)
    out_r = engine.ks_pbs(np.concatenate([rolled, cts[592:]]), lut, tables)
    assert np.array_equal(np.roll(out_r[:592], -148, axis=0), big[:592])


@pytest.mark.parametrize("n", [8, 100, 200, 640])
def test_box_edges_and_modswitch_boundaries(engine, keyset, n):
    """a1.5 directly: TRIVIAL small-key ciphertexts (mask 0, so the blind rotation is a pure rotation by
    the modulus-switched body) placed at every LUT box boundary -1, 0, +1 modulus-switch steps and at the
    rounding threshold of the modulus switch itself, through the pair (n = 8), wide (100, 200) and
    4-slot (640) kernels via fhs_pbs_batch.  Box i covers b~ in [128 i - 64, 128 i + 64) of 4096; the
    switch rounds ((x >> 51) + 1) >> 1, i.e. half-up at 2^51."""
    table = (np.arange(16, dtype=U) * 7 + 1) % 16
    cases = []                                    # (body, expected decryption mod 32)
    step = 1 << 52                                # one modulus-switch step of the torus
    for i in range(32):
        for d in (-1, 0, 1):
            bt = (128 * i - 64 + d) % 4096        # lower edge of box i, +-1 step
            val = ((bt + 64) // 128) % 32
            cases.append(((bt * step) % 2**64, val))
    # rounding of the switch: x just below / at half a step rounds down / up
    for bt in (64, 63, 4095, 2047, 2048):
        val_up = (((bt + 1) % 4096 + 64) // 128) % 32
        val_dn = ((bt + 64) // 128) % 32
        cases.append(((bt * step + (1 << 51) - 1) % 2**64, val_dn))
        cases.append(((bt * step + (1 << 51)) % 2**64, val_up))
    if n >= len(cases):                           # one launch of exactly n ciphertexts
        chunks = [(cases * ((n + len(cases) - 1) // len(cases)))[:n]]
    else:                                         # every case, in launches of n ciphertexts
        chunks = [cases[i:i + n] for i in range(0, len(cases), n)]
    for ch in chunks:
        small = np.zeros((len(ch), keyset.small), dtype=U)
        small[:, -1] = np.array([c[0] for c in ch], dtype=U)
        out = engine.pbs(small, np.zeros(len(ch), np.uint32), table[None, :])
        got = keyset.decrypt(out).astype(np.int64)
        want = np.array([int(table[v]) if v < 16 else (32 - int(table[v - 16])) % 32 for _, v in ch])
        assert np.array_equal(got, want), np.nonzero(got != want)[0][:10]
        # a pure rotation of an exact LUT: no noise at all
        assert np.abs(keyset.phase_error(out, want)).max() < 2.0**-30
