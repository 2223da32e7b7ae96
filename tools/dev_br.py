"""dev tool: time + check the blind rotation (device-buffer entry point) at the batch sizes given on the
command line; env FHS_BR_* select kernel variants (read once per context).  Test infrastructure: uses oracle keys."""
import sys, time, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..')); sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', 'tests'))
import numpy as np, torch
from oracle import oracle as O
import fhe_string_b200 as F
from conftest import bench_tables
ks = O.Keyset(seed=1); eng = F.Engine(0); eng.upload_keys(ks.p, ks.ksk(), ks.bsk())
tabs = bench_tables(); rng = np.random.default_rng(0)
tag = " ".join(f"{k}={v}" for k, v in os.environ.items() if k.startswith("FHS_"))
d_tabs = torch.from_numpy(tabs.view(np.int64)).cuda()
for B in [int(x) for x in (sys.argv[1:] or ["4096", "148", "592"])]:
    m = rng.integers(0, 16, B); c = ks.encrypt(m, seed=10 + B); l = (np.arange(B) % 8).astype(np.uint32)
    d_in = torch.from_numpy(c.view(np.int64)).cuda(); d_l = torch.from_numpy(l.view(np.int32)).cuda()
    d_out = torch.zeros_like(d_in)
    best = 1e9
    for rep in range(4):
        eng.ks_pbs_dev(d_in.data_ptr(), d_l.data_ptr(), d_tabs.data_ptr(), len(tabs), B, d_out.data_ptr())
        eng.synchronize(); km = eng.last_kernel_ms(); best = min(best, km[1]); bks = min(km[0], bks) if rep else km[0]
    o = d_out.cpu().numpy().view(np.uint64)
    e = tabs[l, m]; d = ks.decrypt(o)
    print(f"[{tag}] B={B} ks_ms={bks:.4f} br_ms={best:.3f} PBS/s={B/best*1e3:.0f} correct={np.array_equal(d, e)} "
          f"noise={np.std(ks.phase_error(o, e)):.3e}", flush=True)
