"""dev: back-to-back small KS-PBS launches, wall per launch vs kernel times (launch gaps)."""
import sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np, torch
from oracle import oracle as O
import fhe_string_b200 as F
from conftest import bench_tables
ks = O.Keyset(seed=1); eng = F.Engine(0); eng.upload_keys(ks.p, ks.ksk(), ks.bsk())
tabs = bench_tables(); rng = np.random.default_rng(0)
dev = torch.device("cuda", 0)
for B in (16, 100):
    m = rng.integers(0, 16, B); c = ks.encrypt(m, seed=B); l = (np.arange(B) % 8).astype(np.int32)
    d_in = torch.from_numpy(c.view(np.int64)).to(dev); d_lut = torch.from_numpy(l).to(dev)
    d_tab = torch.from_numpy(tabs.view(np.int64).copy()).to(dev); d_out = torch.empty_like(d_in)
    for reps in (1, 50):
        torch.cuda.synchronize(); eng.sync() if hasattr(eng, "sync") else None
        t = time.perf_counter()
        for _ in range(reps):
            eng.ks_pbs_dev(d_in.data_ptr(), d_lut.data_ptr(), d_tab.data_ptr(), len(tabs), B, d_out.data_ptr())
        t_issue = (time.perf_counter() - t) * 1e3
        eng.L.fhs_ctx_synchronize(eng.h)
        dt = (time.perf_counter() - t) * 1e3
        print(f"B={B} reps={reps}: issue {t_issue/reps:.3f} ms/call, total {dt/reps:.3f} ms/call, kernels {eng.last_kernel_ms()}", flush=True)
