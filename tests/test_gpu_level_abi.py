"""GPU: the pool / level entry points of include/fhs_cuda.h called directly (the level scheduler reaches them
through fhs_pool_load_program / fhs_level_run; a Rust host would call these): fhs_pool_create / write / read,
fhs_level and its descriptor form fhs_level_d, argument validation."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


class LevelDesc(ctypes.Structure):
    _fields_ = [("B", ctypes.c_size_t), ("term_off", ctypes.c_void_p), ("term_idx", ctypes.c_void_p),
                ("term_coef", ctypes.c_void_p), ("const_body", ctypes.c_void_p), ("lut_id", ctypes.c_void_p),
                ("lut_tables", ctypes.c_void_p), ("n_luts", ctypes.c_size_t), ("out_slot", ctypes.c_size_t)]


def _bind(L):
    vp, sz = ctypes.c_void_p, ctypes.c_size_t
    L.fhs_pool_create.argtypes = [vp, sz, ctypes.POINTER(vp)]
    L.fhs_pool_destroy.argtypes = [vp]
    L.fhs_pool_destroy.restype = None
    L.fhs_pool_write.argtypes = [vp, sz, sz, vp]
    L.fhs_pool_read.argtypes = [vp, sz, sz, vp]
    L.fhs_level.argtypes = [vp, sz, vp, vp, vp, vp, vp, vp, sz, sz]
    L.fhs_level_d.argtypes = [vp, ctypes.POINTER(LevelDesc)]
    L.fhs_pool_capacity.argtypes = [vp]
    L.fhs_pool_capacity.restype = sz


def test_level_bivariate_pack_and_linear_level(engine, keyset):
    """One level = 'in_b = sum coef * pool[idx] + const; out = LUT[in_b]': the bivariate pack 4x + y of tfhe-rs
    (unchecked_apply_lookup_table_bivariate) for 40 pairs, through fhs_level and fhs_level_d, then a purely
    linear level (lut_id = NULL) that subtracts two results."""
    L = engine.L
    _bind(L)
    rng = np.random.default_rng(11)
    n = 40
    x, y = rng.integers(0, 4, n), rng.integers(0, 4, n)
    cts = keyset.encrypt(np.concatenate([x, y]), seed=3)
    pool = ctypes.c_void_p()
    assert L.fhs_pool_create(engine.h, 4 * n, ctypes.byref(pool)) == 0
    assert L.fhs_pool_capacity(pool) == 4 * n
    assert L.fhs_pool_write(pool, 0, 2 * n, cts.ctypes.data) == 0
    off = (np.arange(n + 1) * 2).astype(np.uint32)
    idx = np.stack([np.arange(n), n + np.arange(n)], axis=1).astype(np.uint32).reshape(-1)
    coef = np.tile(np.array([4, 1], dtype=np.int32), n)
    cst = np.zeros(n, dtype=np.uint64)
    tables = np.zeros((2, 16), dtype=np.uint64)
    for v in range(16):
        tables[0, v] = ((v >> 2) * (v & 3)) % 4          # f(x, y) = x * y mod 4
        tables[1, v] = 1 if (v >> 2) == (v & 3) else 0   # x == y
    lut = (np.arange(n) % 2).astype(np.uint32)
    assert L.fhs_level(pool, n, off.ctypes.data, idx.ctypes.data, coef.ctypes.data, cst.ctypes.data, lut.ctypes.data,
                       tables.ctypes.data, 2, 2 * n) == 0, L.fhs_last_error()
    d = LevelDesc(n, off.ctypes.data, idx.ctypes.data, coef.ctypes.data, cst.ctypes.data, lut.ctypes.data,
                  tables.ctypes.data, 2, 3 * n)
    assert L.fhs_level_d(pool, ctypes.byref(d)) == 0
    out = np.empty((2 * n, keyset.big), dtype=np.uint64)
    assert L.fhs_pool_read(pool, 2 * n, 2 * n, out.ctypes.data) == 0
    want = tables[lut, 4 * x + y]
    assert np.array_equal(keyset.decrypt(out[:n]), want) and np.array_equal(keyset.decrypt(out[n:]), want)
    # linear level: slot[2n + b] - slot[3n + b] + 5  ->  5 (the two runs agree), written over the inputs
    off2 = (np.arange(n + 1) * 2).astype(np.uint32)
    idx2 = np.stack([2 * n + np.arange(n), 3 * n + np.arange(n)], axis=1).astype(np.uint32).reshape(-1)
    coef2 = np.tile(np.array([1, -1], dtype=np.int32), n)
    cst2 = np.full(n, 5 * keyset.delta, dtype=np.uint64)
    assert L.fhs_level(pool, n, off2.ctypes.data, idx2.ctypes.data, coef2.ctypes.data, cst2.ctypes.data, None, None, 0, 0) == 0
    lin = np.empty((n, keyset.big), dtype=np.uint64)
    assert L.fhs_pool_read(pool, 0, n, lin.ctypes.data) == 0
    assert np.array_equal(keyset.decrypt(lin), np.full(n, 5))
    # validation: a LUT id without a table, a term outside the pool, an output range outside the pool
    bad_lut = lut.copy()
    bad_lut[7] = 2
    assert L.fhs_level(pool, n, off.ctypes.data, idx.ctypes.data, coef.ctypes.data, cst.ctypes.data, bad_lut.ctypes.data,
                       tables.ctypes.data, 2, 2 * n) == -1
    assert b"lut_id[7]" in L.fhs_last_error()
    bad_idx = idx.copy()
    bad_idx[3] = 4 * n
    assert L.fhs_level(pool, n, off.ctypes.data, bad_idx.ctypes.data, coef.ctypes.data, cst.ctypes.data, lut.ctypes.data,
                       tables.ctypes.data, 2, 2 * n) == -1
    assert b"term_idx[3]" in L.fhs_last_error()
    assert L.fhs_level(pool, n, off.ctypes.data, idx.ctypes.data, coef.ctypes.data, cst.ctypes.data, lut.ctypes.data,
                       tables.ctypes.data, 2, 3 * n + 1) == -1
    L.fhs_pool_destroy(pool)


def test_upload_keys_fmt_flag(keyset):
    """fhs_upload_keys_fmt (one entry point, bsk_format flag) with the standard-domain key."""
    import fhe_string_b200 as F
    from fhe_string_b200.engine import Params
    eng = F.Engine(0)
    p = Params()
    for name, _ in Params._fields_:
        setattr(p, name, int(getattr(keyset.p, name)))
    ksk, bsk = np.ascontiguousarray(keyset.ksk()), np.ascontiguousarray(keyset.bsk())
    eng.L.fhs_upload_keys_fmt.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    assert eng.L.fhs_upload_keys_fmt(eng.h, ctypes.byref(p), ksk.ctypes.data, 0, bsk.ctypes.data, 7) == -1   # bad flag
    assert eng.L.fhs_upload_keys_fmt(eng.h, ctypes.byref(p), ksk.ctypes.data, 0, bsk.ctypes.data, -1) == 0
    eng._keys_uploaded(p, p.glwe_dim * p.poly_size)
    msgs = np.arange(16)
    out = eng.ks_pbs(keyset.encrypt(msgs, seed=1), np.zeros(16, np.uint32), np.arange(16, dtype=np.uint64)[None, :])
    assert np.array_equal(keyset.decrypt(out), msgs)
    eng.close()
