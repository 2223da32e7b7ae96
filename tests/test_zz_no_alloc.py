"""GPU: fhs_ks_pbs_batch_dev_noalloc (include/fhs_cuda.h) -- the allocation-free variant of the KS-PBS
device entry point, scratch in file-scope device arrays of the library.  No counterpart in the reference.

Collected last on purpose.  Passes alone in a fresh process and after the
other tests: it makes its own device buffers and, for the "allocates nothing" check, its own context."""
import ctypes

import numpy as np
import pytest

from conftest import bench_tables

pytestmark = pytest.mark.gpu
U = np.uint64
MAX_BATCH = 384          # FHS_NOALLOC_MAX_BATCH


def _device_operands(keyset, n, seed):
    import torch
    tables = bench_tables()
    rng = np.random.default_rng(seed)
    msgs = rng.integers(0, 16, n)
    cts = keyset.encrypt(msgs, seed=seed)
    lut = ((np.arange(n) * 5 + seed) % 8).astype(np.uint32)
    d_in = torch.from_numpy(cts.view(np.int64)).cuda()
    d_lut = torch.from_numpy(lut.view(np.int32)).cuda()
    d_tab = torch.from_numpy(tables.view(np.int64).copy()).cuda()
    return tables, msgs, lut, d_in, d_lut, d_tab


def _call(L, name, eng, d_in, d_lut, d_tab, n_luts, b, d_out):
    rc = getattr(L, name)(eng.h, d_in.data_ptr(), d_lut.data_ptr(), d_tab.data_ptr(), n_luts, b, d_out.data_ptr())
    assert rc == 0, (name, b, L.fhs_last_error())


@pytest.mark.parametrize("n", [1, 74, 129, 296, MAX_BATCH])
def test_noalloc_matches_the_existing_entry_point_twice_in_a_row(engine, keyset, n):
    """Through the C ABI, same device inputs: two calls of the new entry point in a row (the second one on
    other ciphertexts, so stale scratch would show), each bit-identical to fhs_ks_pbs_batch_dev -- the
    tolerance of tests/test_gpu_hygiene.py for one kernel on one input -- and decrypting to LUT[msg]."""
    import torch
    L = engine.L
    for call, seed in enumerate((4000 + n, 5000 + n)):
        tables, msgs, lut, d_in, d_lut, d_tab = _device_operands(keyset, n, seed)
        want = torch.zeros_like(d_in)
        got = torch.full_like(d_in, -1)
        _call(L, "fhs_ks_pbs_batch_dev", engine, d_in, d_lut, d_tab, len(tables), n, want)
        _call(L, "fhs_ks_pbs_batch_dev_noalloc", engine, d_in, d_lut, d_tab, len(tables), n, got)
        engine.synchronize()
        w, g = want.cpu().numpy().view(U), got.cpu().numpy().view(U)
        assert np.array_equal(g, w), f"call {call + 1}: {(g != w).sum()} words differ"
        assert np.array_equal(keyset.decrypt(g), tables[lut, msgs])


def test_noalloc_smaller_batch_after_a_larger_one(engine, keyset):
    """The arrays keep the last call's contents: a 5-ciphertext call after a 384-ciphertext one must not see
    them (digit rows past the batch are rewritten as zero by the call itself)."""
    import torch
    L = engine.L
    tables, _, _, d_in, d_lut, d_tab = _device_operands(keyset, MAX_BATCH, 61)
    big = torch.zeros_like(d_in)
    _call(L, "fhs_ks_pbs_batch_dev_noalloc", engine, d_in, d_lut, d_tab, len(tables), MAX_BATCH, big)
    tables, msgs, lut, d_in, d_lut, d_tab = _device_operands(keyset, 5, 62)
    want, got = torch.zeros_like(d_in), torch.zeros_like(d_in)
    _call(L, "fhs_ks_pbs_batch_dev_noalloc", engine, d_in, d_lut, d_tab, len(tables), 5, got)
    _call(L, "fhs_ks_pbs_batch_dev", engine, d_in, d_lut, d_tab, len(tables), 5, want)
    engine.synchronize()
    assert np.array_equal(got.cpu().numpy(), want.cpu().numpy())
    assert np.array_equal(keyset.decrypt(got.cpu().numpy().view(U)), tables[lut, msgs])


def test_noalloc_rejects_sizes_beyond_the_static_scratch(engine, keyset):
    import torch
    L = engine.L
    L.fhs_last_error.restype = ctypes.c_char_p
    tables, _, _, d_in, d_lut, d_tab = _device_operands(keyset, 2, 7)
    d_out = torch.zeros_like(d_in)
    rc = L.fhs_ks_pbs_batch_dev_noalloc(engine.h, d_in.data_ptr(), d_lut.data_ptr(), d_tab.data_ptr(), len(tables),
                                        MAX_BATCH + 1, d_out.data_ptr())
    assert rc == -1                                          # FHS_ERR_INVALID, nothing launched
    msg = L.fhs_last_error().decode()
    assert "FHS_NOALLOC_MAX_BATCH = 384" in msg and "385" in msg
    assert L.fhs_ks_pbs_batch_dev_noalloc(engine.h, d_in.data_ptr(), d_lut.data_ptr(), d_tab.data_ptr(), len(tables),
                                          0, d_out.data_ptr()) == 0          # empty batch: no-op, as the other entry
    assert L.fhs_ks_pbs_batch_dev_noalloc(engine.h, None, d_lut.data_ptr(), d_tab.data_ptr(), len(tables), 2,
                                          d_out.data_ptr()) == -1


def test_noalloc_allocates_nothing_on_a_fresh_context(engine, keyset):
    """A context that has never run a batch has no scratch of its own (fhs_ks_pbs_batch_dev would grow it by
    ~19 MB at 384 ciphertexts).  After every kernel has been loaded once through the session's context, the
    free device memory is the same before and after the new entry point's calls on a fresh context."""
    import torch
    import fhe_string_b200 as F
    L = engine.L
    sizes = (MAX_BATCH, 74, 200)
    ops = {n: _device_operands(keyset, n, 300 + n) for n in sizes}
    outs = {n: torch.zeros_like(ops[n][3]) for n in sizes}
    for n in sizes:                                          # loads the kernels of every wave shape
        _call(L, "fhs_ks_pbs_batch_dev", engine, *ops[n][3:], 8, n, outs[n])
        _call(L, "fhs_ks_pbs_batch_dev_noalloc", engine, *ops[n][3:], 8, n, outs[n])
    engine.synchronize()
    fresh = F.Engine(0)
    try:
        fresh.upload_keys(keyset.p, keyset.ksk(), keyset.bsk(), ksk_level_order=0)
        fresh.synchronize()
        torch.cuda.synchronize()
        free0, _ = torch.cuda.mem_get_info()
        for n in sizes:
            _call(L, "fhs_ks_pbs_batch_dev_noalloc", fresh, *ops[n][3:], 8, n, outs[n])
        fresh.synchronize()
        free1, _ = torch.cuda.mem_get_info()
        assert free1 == free0, f"{free0 - free1} bytes of device memory taken by an allocation-free call"
        for n in sizes:
            tables, msgs, lut = ops[n][:3]
            assert np.array_equal(keyset.decrypt(outs[n].cpu().numpy().view(U)), tables[lut, msgs])
    finally:
        fresh.close()
