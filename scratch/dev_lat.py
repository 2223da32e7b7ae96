"""dev: per-level overhead: wall vs kernel time for small batches via the level API."""
import sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
from oracle import oracle as O
import fhe_string_b200 as F
from fhe_string_b200.program import Program
from test_gpu_strings import run_gpu
ks = O.Keyset(seed=1); eng = F.Engine(0); eng.upload_keys(ks.p, ks.ksk(), ks.bsk())
from conftest import bench_tables
tabs = bench_tables(); rng = np.random.default_rng(0)
for B in (16, 148, 296):
    m = rng.integers(0, 16, B); c = ks.encrypt(m, seed=B); l = (np.arange(B) % 8).astype(np.uint32)
    for rep in range(3):
        t = time.perf_counter(); o = eng.ks_pbs(c, l, tabs); dt = (time.perf_counter() - t) * 1e3
    print(f"B={B} host-buffer call wall {dt:.3f} ms, kernels ks/br {eng.last_kernel_ms()}", flush=True)
s = "the quick brown fox jumps over the lazy dog and runs away"[:40]
for op, args, pads in (("split", [s, "fox"], [64, None]), ("find", [s, "lazy"], [64, None])):
    for rep in range(2):
        t = time.perf_counter(); v, st = run_gpu(eng, ks, op, args, pads); dt = (time.perf_counter() - t) * 1e3
    print(op, st, f"wall incl. encrypt/decrypt {dt:.1f} ms -> {st['ms']/st['levels']:.3f} ms/level", flush=True)
