"""ctypes binding of include/fhs_cuda.h (libfhs_cuda.so).  Host-side plumbing only."""
import ctypes
import os
import weakref

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libfhs_cuda.so")

u64p = ctypes.POINTER(ctypes.c_uint64)
u32p = ctypes.POINTER(ctypes.c_uint32)


class FhsError(RuntimeError):
    pass


class Params(ctypes.Structure):
    """fhs_params: tfhe-rs ClassicPBSParameters subset (reference src/lib.rs:31,45,50)."""
    _fields_ = [(n, ctypes.c_uint32) for n in (
        "lwe_dim", "glwe_dim", "poly_size", "pbs_base_log", "pbs_level", "ks_base_log", "ks_level",
        "message_modulus", "carry_modulus", "reserved")]


# every symbol include/fhs_cuda.h declares
# include/fhs_cuda.h FHS_FOURIER_*: exponent sign of a caller-supplied Fourier BSK
FOURIER_NATURAL_POS = 0
FOURIER_NATURAL_NEG = 1

EXPORTS = [
    "fhs_default_params", "fhs_last_error", "fhs_version", "fhs_ctx_create", "fhs_ctx_destroy",
    "fhs_ctx_set_stream", "fhs_ctx_synchronize", "fhs_upload_keys", "fhs_upload_keys_fourier", "fhs_ks_pbs_batch",
    "fhs_comm_unique_id", "fhs_comm_init", "fhs_comm_info", "fhs_comm_destroy", "fhs_comm_allgather",
    "fhs_comm_broadcast_host",
    "fhs_ks_pbs_batch_dev", "fhs_ks_pbs_batch_dev_noalloc", "fhs_keyswitch_batch", "fhs_keyswitch_batch_dev", "fhs_build_lut",
    "fhs_pbs_batch", "fhs_pbs_batch_dev", "fhs_last_kernel_ms", "fhs_launch_count",
    "fhs_measure_fp64_tflops",
    "fhs_pool_enqueue_write", "fhs_pool_enqueue_read", "fhs_pool_copy_in", "fhs_pool_copy_out",
    "fhs_dev_alloc", "fhs_dev_free", "fhs_dev_write", "fhs_dev_read", "fhs_host_alloc", "fhs_host_free",
    "fhs_upload_keys_fmt", "fhs_level_d", "fhs_host_is_pinned", "fhs_ctx_cache_get", "fhs_ctx_cache_put", "fhs_cache_drop_owner",
    "fhs_ctx_cache_stats", "fhs_ctx_cache_limit", "fhs_ctx_timer_start", "fhs_ctx_timer_stop", "fhs_ctx_timer_ms",
    "fhs_pool_create_shared", "fhs_pool_is_shared", "fhs_pool_peer_allgather", "fhs_pool_peer_run_done",
    "fhs_pool_peer_status", "fhs_comm_peer_pushes",
]

_lib = None


def lib_path():
    return _LIB


def load_library():
    """Loads libfhs_cuda.so; raises FhsError if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB):
        raise FhsError(f"{_LIB} is missing: build it with `python -m fhe_string_b200.build` "
                       "(there is no CPU fallback)")
    L = ctypes.CDLL(_LIB)
    vp = ctypes.c_void_p
    sz = ctypes.c_size_t
    L.fhs_default_params.argtypes = [ctypes.POINTER(Params)]
    L.fhs_default_params.restype = None
    L.fhs_last_error.restype = ctypes.c_char_p
    L.fhs_version.restype = ctypes.c_int
    L.fhs_ctx_create.argtypes = [ctypes.c_int, ctypes.POINTER(vp)]
    L.fhs_ctx_destroy.argtypes = [vp]
    L.fhs_ctx_destroy.restype = None
    L.fhs_ctx_set_stream.argtypes = [vp, vp]
    L.fhs_ctx_synchronize.argtypes = [vp]
    L.fhs_upload_keys.argtypes = [vp, ctypes.POINTER(Params), vp, ctypes.c_int, vp]
    L.fhs_upload_keys_fourier.argtypes = [vp, ctypes.POINTER(Params), vp, ctypes.c_int, vp, ctypes.c_int]
    L.fhs_comm_unique_id.argtypes = [vp]
    L.fhs_comm_init.argtypes = [vp, vp, ctypes.c_int, ctypes.c_int]
    L.fhs_comm_info.argtypes = [vp, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int)]
    L.fhs_comm_destroy.argtypes = [vp]
    L.fhs_comm_destroy.restype = None
    L.fhs_comm_allgather.argtypes = [vp, vp, ctypes.c_size_t]
    L.fhs_comm_broadcast_host.argtypes = [vp, vp, ctypes.c_size_t, ctypes.c_int]
    L.fhs_ks_pbs_batch.argtypes = [vp, vp, vp, vp, sz, sz, vp]
    L.fhs_ks_pbs_batch_dev.argtypes = [vp, vp, vp, vp, sz, sz, vp]
    L.fhs_ks_pbs_batch_dev_noalloc.argtypes = [vp, vp, vp, vp, sz, sz, vp]
    L.fhs_keyswitch_batch.argtypes = [vp, vp, sz, vp]
    L.fhs_keyswitch_batch_dev.argtypes = [vp, vp, sz, vp]
    L.fhs_build_lut.argtypes = [vp, vp, vp]
    L.fhs_pbs_batch.argtypes = [vp, vp, vp, vp, sz, sz, vp]
    L.fhs_pbs_batch_dev.argtypes = [vp, vp, vp, vp, sz, sz, vp]
    L.fhs_last_kernel_ms.argtypes = [vp, ctypes.POINTER(ctypes.c_float)]
    L.fhs_launch_count.argtypes = [vp]
    L.fhs_launch_count.restype = ctypes.c_uint64
    L.fhs_measure_fp64_tflops.argtypes = [vp, ctypes.POINTER(ctypes.c_double)]
    L.fhs_dev_alloc.argtypes = [vp, sz, ctypes.POINTER(vp)]
    L.fhs_dev_free.argtypes = [vp, vp]
    L.fhs_dev_free.restype = None
    L.fhs_dev_write.argtypes = [vp, vp, sz, vp]
    L.fhs_dev_read.argtypes = [vp, vp, sz, vp]
    L.fhs_host_alloc.argtypes = [vp, sz, ctypes.POINTER(vp)]
    L.fhs_host_free.argtypes = [vp, vp]
    L.fhs_host_free.restype = None
    L.fhs_host_is_pinned.argtypes = [vp]
    L.fhs_ctx_cache_stats.argtypes = [vp, ctypes.POINTER(sz), ctypes.POINTER(sz)]
    L.fhs_ctx_cache_limit.argtypes = [vp, sz]
    L.fhs_comm_peer_pushes.argtypes = [vp]
    L.fhs_comm_peer_pushes.restype = ctypes.c_ulonglong
    _lib = L
    return L


def default_params():
    p = Params()
    load_library().fhs_default_params(ctypes.byref(p))
    return p


def _np(a, dtype):
    a = np.ascontiguousarray(a, dtype=dtype)
    return a, a.ctypes.data


class Engine:
    """One context on one GPU (fhs_ctx).  Mirrors the role of the reference's ServerKey
    (src/server_key.rs:12-33): holds the device-resident key material and evaluates levels."""

    def __init__(self, device=0):
        self.L = load_library()
        h = ctypes.c_void_p()
        self._check(self.L.fhs_ctx_create(int(device), ctypes.byref(h)))
        self.h = h
        self.device = device
        self.p = None
        self.comm = None   # (rank, world) once comm_init has run

    def _check(self, rc):
        if rc != 0:
            raise FhsError(f"fhs error {rc}: {self.L.fhs_last_error().decode()}")

    def close(self):
        if getattr(self, "h", None):
            for ch in getattr(self, "_pinned", []):
                self.L.fhs_host_free(self.h, ctypes.c_void_p(ch))
            self._pinned = []
            for lst in getattr(self, "_pin_free", {}).values():     # result buffers nobody holds any more
                while lst:
                    self.L.fhs_host_free(self.h, ctypes.c_void_p(lst.pop()))
            self.L.fhs_ctx_destroy(self.h)    # also frees cached executions and device arrays' allocator
            self.h = None

    # ---- residency: device-resident ciphertext arrays, page-locked host buffers, cache ----
    def dev_alloc(self, n_blocks):
        d = ctypes.c_void_p()
        self._check(self.L.fhs_dev_alloc(self.h, int(n_blocks), ctypes.byref(d)))
        return int(d.value)

    def dev_free(self, ptr):
        if getattr(self, "h", None) and ptr:
            self.L.fhs_dev_free(self.h, ctypes.c_void_p(ptr))

    def dev_write(self, ptr, blocks):
        blocks = np.ascontiguousarray(blocks, dtype=np.uint64).reshape(-1, self.big)
        self._check(self.L.fhs_dev_write(self.h, ctypes.c_void_p(ptr), len(blocks), blocks.ctypes.data))

    def dev_read(self, ptr, n_blocks, out=None):
        if out is None:
            out = np.empty((n_blocks, self.big), dtype=np.uint64)
        self._check(self.L.fhs_dev_read(self.h, ctypes.c_void_p(ptr), int(n_blocks), out.ctypes.data))
        return out

    def pinned_blocks(self, n_blocks):
        """A page-locked u64 [n_blocks][big] array that lives until close() (operand staging arenas)."""
        q = ctypes.c_void_p()
        nbytes = max(1, int(n_blocks)) * self.big * 8
        self._check(self.L.fhs_host_alloc(self.h, nbytes, ctypes.byref(q)))
        if not hasattr(self, "_pinned"):
            self._pinned = []
        self._pinned.append(int(q.value))
        buf = (ctypes.c_uint64 * (max(1, int(n_blocks)) * self.big)).from_address(q.value)
        return np.ctypeslib.as_array(buf).reshape(-1, self.big)

    def pinned_array(self, n_blocks):
        """A page-locked u64 [n_blocks][big] array for RESULTS: recycled through a size-class free list when
        the last numpy view of it is garbage-collected (host-array mode returns these, so a result can be
        passed back as an operand, or read by the client, without another staging copy)."""
        if not hasattr(self, "_pin_free"):
            self._pin_free = {}
        n = max(1, int(n_blocks))
        cap = 1 << max(4, (n - 1).bit_length())
        lst = self._pin_free.setdefault(cap, [])
        if lst:
            ptr = lst.pop()
        else:
            q = ctypes.c_void_p()
            self._check(self.L.fhs_host_alloc(self.h, cap * self.big * 8, ctypes.byref(q)))
            ptr = int(q.value)
        buf = (ctypes.c_uint64 * (cap * self.big)).from_address(ptr)
        weakref.finalize(buf, lst.append, ptr)       # back to the free list when every view is gone
        return np.ctypeslib.as_array(buf).reshape(cap, self.big)[:n_blocks]

    def is_pinned(self, arr):
        return bool(self.L.fhs_host_is_pinned(ctypes.c_void_p(arr.ctypes.data)))

    def cache_stats(self):
        n, b = ctypes.c_size_t(), ctypes.c_size_t()
        self._check(self.L.fhs_ctx_cache_stats(self.h, ctypes.byref(n), ctypes.byref(b)))
        return int(n.value), int(b.value)

    def cache_limit(self, device_bytes):
        self._check(self.L.fhs_ctx_cache_limit(self.h, int(device_bytes)))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- keys ----
    def upload_keys(self, params, ksk, bsk, ksk_level_order=0):
        p = Params()
        for name, _ in Params._fields_:
            setattr(p, name, int(getattr(params, name)))
        ksk, kp = _np(ksk, np.uint64)
        bsk, bp = _np(bsk, np.uint64)
        big = p.glwe_dim * p.poly_size
        assert ksk.size == big * p.ks_level * (p.lwe_dim + 1), "ksk size"
        assert bsk.size == p.lwe_dim * p.pbs_level * (p.glwe_dim + 1) ** 2 * p.poly_size, "bsk size"
        self._check(self.L.fhs_upload_keys(self.h, ctypes.byref(p), kp, int(ksk_level_order), bp))
        self._keys_uploaded(p, big)

    def upload_keys_fourier(self, params, ksk, bsk_fourier, ksk_level_order=0, convention=FOURIER_NATURAL_NEG):
        """fhs_upload_keys_fourier: `bsk_fourier` complex128 [n][level][k+1][k+1][N/2], natural order."""
        p = Params()
        for name, _ in Params._fields_:
            setattr(p, name, int(getattr(params, name)))
        ksk, kp = _np(ksk, np.uint64)
        bf = np.ascontiguousarray(bsk_fourier, dtype=np.complex128)
        big = p.glwe_dim * p.poly_size
        assert ksk.size == big * p.ks_level * (p.lwe_dim + 1), "ksk size"
        assert bf.size == p.lwe_dim * p.pbs_level * (p.glwe_dim + 1) ** 2 * p.poly_size // 2, "bsk size"
        self._check(self.L.fhs_upload_keys_fourier(self.h, ctypes.byref(p), kp, int(ksk_level_order),
                                                   bf.ctypes.data_as(ctypes.c_void_p), int(convention)))
        self._keys_uploaded(p, big)

    # ---- multi-GPU (one process per GPU): the library's own NCCL communicator ----
    def comm_unique_id(self):
        """fhs_comm_unique_id: 128 bytes made on one rank, to be shipped to the others."""
        buf = (ctypes.c_uint8 * 128)()
        self._check(self.L.fhs_comm_unique_id(buf))
        return bytes(buf)

    def comm_init(self, unique_id, rank, world):
        buf = (ctypes.c_uint8 * 128).from_buffer_copy(unique_id)
        self._check(self.L.fhs_comm_init(self.h, buf, int(rank), int(world)))
        self.comm = (int(rank), int(world))

    def comm_init_from_torch(self):
        """Bootstrap over an initialised torch.distributed group: rank 0's id is broadcast as an object."""
        import torch.distributed as dist
        rank, world = dist.get_rank(), dist.get_world_size()
        box = [self.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, 0)
        self.comm_init(box[0], rank, world)

    def comm_broadcast(self, arr, root=0):
        """fhs_comm_broadcast_host on a numpy array that has the same shape/dtype on every rank (in place)."""
        assert arr.flags["C_CONTIGUOUS"] and arr.flags["WRITEABLE"]
        self._check(self.L.fhs_comm_broadcast_host(self.h, arr.ctypes.data_as(ctypes.c_void_p), arr.nbytes, int(root)))
        return arr

    def comm_allgather_dev(self, dev_ptr, bytes_per_rank):
        self._check(self.L.fhs_comm_allgather(self.h, ctypes.c_void_p(dev_ptr), int(bytes_per_rank)))

    def comm_destroy(self):
        self.L.fhs_comm_destroy(self.h)
        self.comm = None

    def _keys_uploaded(self, p, big):
        self.p = p
        self.big = big + 1
        self.small = p.lwe_dim + 1
        self.space = p.message_modulus * p.carry_modulus

    def set_stream(self, cuda_stream_ptr):
        self._check(self.L.fhs_ctx_set_stream(self.h, ctypes.c_void_p(cuda_stream_ptr or 0)))

    def synchronize(self):
        self._check(self.L.fhs_ctx_synchronize(self.h))

    # ---- host-buffer entry points ----
    def ks_pbs(self, cts, lut_id, tables, out=None):
        cts, cp = _np(cts, np.uint64)
        cts = cts.reshape(-1, self.big)
        lut_id, lp = _np(lut_id, np.uint32)
        tables, tp = _np(tables, np.uint64)
        tables = tables.reshape(-1, self.space)
        assert lut_id.size == len(cts)
        assert lut_id.size == 0 or int(lut_id.max()) < len(tables)
        if out is None:
            out = np.empty_like(cts)
        self._check(self.L.fhs_ks_pbs_batch(self.h, cp, lp, tp, len(tables), len(cts),
                                             out.ctypes.data))
        return out

    def ks_pbs_ptr(self, in_ptr, lut_ptr, tab_ptr, n_luts, batch, out_ptr):
        """Host pointers (e.g. pinned torch tensors)."""
        self._check(self.L.fhs_ks_pbs_batch(self.h, in_ptr, lut_ptr, tab_ptr, n_luts, batch, out_ptr))

    def keyswitch(self, cts):
        cts, cp = _np(cts, np.uint64)
        cts = cts.reshape(-1, self.big)
        out = np.empty((len(cts), self.small), dtype=np.uint64)
        self._check(self.L.fhs_keyswitch_batch(self.h, cp, len(cts), out.ctypes.data))
        return out

    def build_lut(self, table):
        table, tp = _np(table, np.uint64)
        assert table.size == self.space
        out = np.empty((self.p.glwe_dim + 1) * self.p.poly_size, dtype=np.uint64)
        self._check(self.L.fhs_build_lut(self.h, tp, out.ctypes.data))
        return out

    def pbs(self, small, lut_id, tables):
        small, sp = _np(small, np.uint64)
        small = small.reshape(-1, self.small)
        lut_id, lp = _np(lut_id, np.uint32)
        tables, tp = _np(tables, np.uint64)
        tables = tables.reshape(-1, self.space)
        out = np.empty((len(small), self.big), dtype=np.uint64)
        self._check(self.L.fhs_pbs_batch(self.h, sp, lp, tp, len(tables), len(small),
                                          out.ctypes.data))
        return out

    # ---- device-pointer entry points (ints from torch .data_ptr()) ----
    def ks_pbs_dev(self, d_in, d_lut_id, d_tables, n_luts, batch, d_out):
        self._check(self.L.fhs_ks_pbs_batch_dev(self.h, d_in, d_lut_id, d_tables, n_luts, batch, d_out))

    def ks_pbs_dev_noalloc(self, d_in, d_lut_id, d_tables, n_luts, batch, d_out):
        """fhs_ks_pbs_batch_dev_noalloc: same results, scratch in the library's static device arrays, at most
        NOALLOC_MAX_BATCH ciphertexts per call (include/fhs_cuda.h)."""
        self._check(self.L.fhs_ks_pbs_batch_dev_noalloc(self.h, d_in, d_lut_id, d_tables, n_luts, batch, d_out))

    def keyswitch_dev(self, d_in, batch, d_out):
        self._check(self.L.fhs_keyswitch_batch_dev(self.h, d_in, batch, d_out))

    def pbs_dev(self, d_small, d_lut_id, d_tables, n_luts, batch, d_out):
        self._check(self.L.fhs_pbs_batch_dev(self.h, d_small, d_lut_id, d_tables, n_luts, batch, d_out))

    def last_kernel_ms(self):
        ms = (ctypes.c_float * 2)()
        self._check(self.L.fhs_last_kernel_ms(self.h, ms))
        return float(ms[0]), float(ms[1])

    def measure_fp64_tflops(self):
        v = ctypes.c_double()
        self._check(self.L.fhs_measure_fp64_tflops(self.h, ctypes.byref(v)))
        return float(v.value)

    def peer_pushes(self):
        """Split levels exchanged by the library's own peer-window kernel (0: ncclAllGather or single GPU)."""
        return int(self.L.fhs_comm_peer_pushes(self.h))

    def launch_count(self):
        return int(self.L.fhs_launch_count(self.h))
