"""ctypes binding of include/fhs_string.h (libfhs_string.so): level programs of fhe_string ops."""
import ctypes
import os

import numpy as np

from .engine import FhsError, load_library

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libfhs_string.so")
_lib = None

HOST_EXPORTS = ["fhs_prog_run_ex", "fhs_prog_export_ranges", "fhs_prog_run_many", "fhs_prog_run_sharded", "fhs_prog_plan", "fhs_prog_build", "fhs_prog_destroy", "fhs_host_last_error", "fhs_prog_info",
                "fhs_prog_output_desc", "fhs_prog_export", "fhs_prog_run"]


def load_host_library():
    global _lib
    if _lib is not None:
        return _lib
    load_library()  # libfhs_cuda.so first (dependency)
    if not os.path.exists(_LIB):
        raise FhsError(f"{_LIB} is missing: build it with `python -m fhe_string_b200.build`")
    L = ctypes.CDLL(_LIB)
    vp, sz = ctypes.c_void_p, ctypes.c_size_t
    L.fhs_prog_build.argtypes = [ctypes.c_char_p, vp, sz, sz, vp, sz, ctypes.c_uint32, ctypes.c_uint32,
                                 ctypes.POINTER(vp)]
    L.fhs_prog_destroy.argtypes = [vp]
    L.fhs_prog_destroy.restype = None
    L.fhs_host_last_error.restype = ctypes.c_char_p
    L.fhs_prog_info.argtypes = [vp, vp]
    L.fhs_prog_output_desc.argtypes = [vp]
    L.fhs_prog_output_desc.restype = ctypes.c_char_p
    L.fhs_prog_export.argtypes = [vp] * 11
    L.fhs_prog_export_ranges.argtypes = [vp] * 6
    L.fhs_prog_run.argtypes = [vp, vp, vp, vp, vp]
    L.fhs_prog_run_ex.argtypes = [vp, vp, vp, sz, vp, vp, vp, ctypes.c_int, ctypes.c_int, sz, vp, vp]
    L.fhs_prog_run_many.argtypes = [vp, vp, sz, vp, vp, vp]
    L.fhs_prog_run_sharded.argtypes = [vp, vp, vp, vp, vp, ctypes.c_int, ctypes.c_int, sz, vp, vp]
    L.fhs_prog_plan.argtypes = [vp, ctypes.c_int, ctypes.c_int, sz, vp, sz, ctypes.POINTER(sz)]
    L.fhs_prog_export_schedule.argtypes = [vp, ctypes.c_int, sz, vp, vp]
    L.fhs_prog_output_noise.argtypes = [vp]
    L.fhs_prog_output_noise.restype = ctypes.c_double
    _lib = L
    return L


class Operand(ctypes.Structure):
    """fhs_operand: a piece of a program's input, in host memory or resident on the device."""
    _fields_ = [("host", ctypes.c_void_p), ("dev", ctypes.c_void_p), ("n_blocks", ctypes.c_size_t)]


STAT_KEYS = ("levels", "pbs", "launches", "ms", "device_ms", "plan_ms", "sharded_levels", "cached")


def _stats(st):
    d = dict(levels=int(st[0]), pbs=int(st[1]), launches=int(st[2]), ms=float(st[3]))
    if len(st) > 4:
        d.update(device_ms=float(st[4]), plan_ms=float(st[5]), sharded_levels=int(st[6]), cached=bool(st[7]))
    return d


class Program:
    """One fhe_string operation compiled for fixed operand shapes."""

    def __init__(self, op, n_chars=(), n_usize_args=0, scalars=(), num_blocks_char=4, num_blocks_usize=4):
        self.L = load_host_library()
        self.op = op
        nc = np.asarray(list(n_chars), dtype=np.uint32)
        sc = np.asarray(list(scalars), dtype=np.uint64)
        h = ctypes.c_void_p()
        rc = self.L.fhs_prog_build(op.encode(), nc.ctypes.data if len(nc) else None, len(nc), int(n_usize_args),
                                   sc.ctypes.data if len(sc) else None, len(sc), num_blocks_char,
                                   num_blocks_usize, ctypes.byref(h))
        if rc != 0:
            raise FhsError(f"fhs_prog_build({op}) failed ({rc}): {self.L.fhs_host_last_error().decode()}")
        self.h = h
        info = np.zeros(8, dtype=np.uint64)
        self.L.fhs_prog_info(self.h, info.ctypes.data)
        (self.n_nodes, self.n_inputs, self.n_outputs, self.n_terms, self.n_luts, self.depth, self.pbs_count,
         self.max_noise) = [int(x) for x in info]
        self.desc = self.L.fhs_prog_output_desc(self.h).decode()
        self.output_noise = float(self.L.fhs_prog_output_noise(self.h))
        self.nb_char, self.nb_usize = num_blocks_char, num_blocks_usize

    def __del__(self):
        try:
            if getattr(self, "h", None):
                self.L.fhs_prog_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def export(self):
        n, t = self.n_nodes, self.n_terms
        d = dict(kind=np.zeros(n, np.uint8), lut=np.zeros(n, np.int32), t0=np.zeros(n, np.uint32),
                 tn=np.zeros(n, np.uint32), cst=np.zeros(n, np.int64), level=np.zeros(n, np.int32),
                 term_src=np.zeros(max(t, 1), np.int32), term_coef=np.zeros(max(t, 1), np.int32),
                 luts=np.zeros((max(self.n_luts, 1), 16), np.uint64), outputs=np.zeros(self.n_outputs, np.int64))
        rc = self.L.fhs_prog_export(self.h, *[d[k].ctypes.data for k in
                                              ("kind", "lut", "t0", "tn", "cst", "level", "term_src", "term_coef",
                                               "luts", "outputs")])
        if rc != 0:
            raise FhsError("fhs_prog_export failed")
        r = dict(in_lo=np.zeros(n, np.int32), in_hi=np.zeros(n, np.int32), out_lo=np.zeros(n, np.int32),
                 out_hi=np.zeros(n, np.int32), signed_in=np.zeros(n, np.uint8))
        if self.L.fhs_prog_export_ranges(self.h, *[r[k].ctypes.data for k in
                                                   ("in_lo", "in_hi", "out_lo", "out_hi", "signed_in")]) != 0:
            raise FhsError("fhs_prog_export_ranges failed")
        d.update(r)
        return d

    def run(self, engine, in_blocks):
        """in_blocks: u64 [n_inputs][2049] -> (out_blocks [n_outputs][2049], stats dict)."""
        inb = np.ascontiguousarray(in_blocks, dtype=np.uint64).reshape(self.n_inputs, -1)
        out = np.empty((self.n_outputs, inb.shape[1] if self.n_inputs else engine.big), dtype=np.uint64)
        stats = np.zeros(4, dtype=np.float64)
        rc = self.L.fhs_prog_run(self.h, engine.h, inb.ctypes.data, out.ctypes.data, stats.ctypes.data)
        if rc != 0:
            raise FhsError(f"fhs_prog_run({self.op}) failed ({rc}): {engine.L.fhs_last_error().decode()}")
        return out, dict(levels=int(stats[0]), pbs=int(stats[1]), launches=int(stats[2]), ms=float(stats[3]))

    def run_ex(self, engine, operands, out_host=None, out_dev=0, rank=0, world=1, shard_min=74):
        """fhs_prog_run_ex.  operands: list of ("host", ndarray[n][big]) / ("dev", device_ptr, n_blocks);
        out_host: ndarray [n_outputs][big] or None; out_dev: device pointer or 0.  Cached execution state
        per (program, engine, rank, world, shard_min).  Multi-rank runs need Engine.comm_init*."""
        ops = (Operand * max(1, len(operands)))()
        keep = []
        for i, o in enumerate(operands):
            if o[0] == "dev":
                ops[i].host, ops[i].dev, ops[i].n_blocks = None, o[1], int(o[2])
            else:
                a = o[1]
                if not (a.flags["C_CONTIGUOUS"] and a.dtype == np.uint64):
                    a = np.ascontiguousarray(a, dtype=np.uint64)
                keep.append(a)
                ops[i].host, ops[i].dev, ops[i].n_blocks = a.ctypes.data, None, a.size // engine.big
        stats = np.zeros(8, dtype=np.float64)
        rc = self.L.fhs_prog_run_ex(self.h, engine.h, ctypes.addressof(ops), len(operands),
                                    out_host.ctypes.data if out_host is not None else None,
                                    ctypes.c_void_p(out_dev) if out_dev else None, stats.ctypes.data,
                                    int(rank), int(world), int(shard_min), None, None)
        if rc != 0:
            raise FhsError(f"fhs_prog_run_ex({self.op}) failed ({rc}): {engine.L.fhs_last_error().decode()} "
                           f"{self.L.fhs_host_last_error().decode()}")
        return _stats(stats)

    def run_many(self, engine, in_blocks):
        """M independent instances: in_blocks u64 [M][n_inputs][2049] -> ([M][n_outputs][2049], stats).
        Every level runs for all instances at once (fhs_prog_run_many)."""
        inb = np.ascontiguousarray(in_blocks, dtype=np.uint64)
        m = inb.shape[0]
        inb = inb.reshape(m, self.n_inputs, -1)
        out = np.empty((m, self.n_outputs, inb.shape[2]), dtype=np.uint64)
        stats = np.zeros(4, dtype=np.float64)
        rc = self.L.fhs_prog_run_many(self.h, engine.h, m, inb.ctypes.data, out.ctypes.data, stats.ctypes.data)
        if rc != 0:
            raise FhsError(f"fhs_prog_run_many({self.op}) failed ({rc}): {engine.L.fhs_last_error().decode()}")
        return out, dict(levels=int(stats[0]), pbs=int(stats[1]), launches=int(stats[2]), ms=float(stats[3]))

    def plan(self, rank, world, shard_min=74):
        """Per level: (B, blocks per rank or 0 if replicated, lo, hi) for `rank` of `world`."""
        n = ctypes.c_size_t()
        buf = np.zeros(4 * max(self.depth, 1), dtype=np.uint64)
        rc = self.L.fhs_prog_plan(self.h, rank, world, shard_min, buf.ctypes.data, self.depth, ctypes.byref(n))
        if rc != 0:
            raise FhsError("fhs_prog_plan failed")
        return buf[: 4 * n.value].reshape(-1, 4)

    def schedule(self, world=1, shard_min=74):
        """(level per node as executed on `world` GPUs, modelled ms of the ASAP schedule, of the balanced one)."""
        lvl = np.zeros(self.n_nodes, dtype=np.int32)
        ms = np.zeros(2, dtype=np.float64)
        if self.L.fhs_prog_export_schedule(self.h, int(world), int(shard_min), lvl.ctypes.data, ms.ctypes.data) != 0:
            raise FhsError("fhs_prog_export_schedule failed")
        return lvl, float(ms[0]), float(ms[1])

    def run_sharded(self, engine, in_blocks, rank, world, shard_min=74):
        """One process per GPU; every rank passes the same inputs.  Levels are split by rank and
        re-assembled in place on the device pool: by the library's own ncclAllGather on the context's
        stream when the engine has a communicator (Engine.comm_init*; no host synchronisation between
        levels), else through a torch.distributed.all_gather_into_tensor callback."""
        inb = np.ascontiguousarray(in_blocks, dtype=np.uint64).reshape(self.n_inputs, -1)
        out = np.empty((self.n_outputs, inb.shape[1] if self.n_inputs else engine.big), dtype=np.uint64)
        stats = np.zeros(4, dtype=np.float64)
        if getattr(engine, "comm", None) == (rank, world):
            rc = self.L.fhs_prog_run_sharded(self.h, engine.h, inb.ctypes.data, out.ctypes.data, stats.ctypes.data,
                                             rank, world, shard_min, None, None)
            if rc != 0:
                raise FhsError(f"fhs_prog_run_sharded({self.op}) failed ({rc}): {engine.L.fhs_last_error().decode()}")
            return out, dict(levels=int(stats[0]), pbs=int(stats[1]), launches=int(stats[2]), ms=float(stats[3]))
        import torch
        import torch.distributed as dist

        class _Dev:  # raw device pointer -> torch tensor via __cuda_array_interface__
            def __init__(self, ptr, n):
                self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<i8", "data": (ptr, False), "version": 3}

        def _allgather(_user, base, bytes_per_rank):
            try:
                n = bytes_per_rank // 8
                whole = torch.as_tensor(_Dev(base, n * world), device=torch.device("cuda", engine.device))
                dist.all_gather_into_tensor(whole, whole[rank * n:(rank + 1) * n].clone())
                torch.cuda.synchronize()
                return 0
            except Exception as e:  # never unwind into C
                print("allgather failed:", e)
                return 1

        cb_t = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t)
        cb = cb_t(_allgather)
        rc = self.L.fhs_prog_run_sharded(self.h, engine.h, inb.ctypes.data, out.ctypes.data, stats.ctypes.data,
                                         rank, world, shard_min, ctypes.cast(cb, ctypes.c_void_p), None)
        if rc != 0:
            raise FhsError(f"fhs_prog_run_sharded({self.op}) failed ({rc}): {engine.L.fhs_last_error().decode()}")
        return out, dict(levels=int(stats[0]), pbs=int(stats[1]), launches=int(stats[2]), ms=float(stats[3]))
