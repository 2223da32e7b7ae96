"""dev: many back-to-back pair-kernel launches of random sizes (a hang here would be a protocol bug)."""
import sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np, torch
from oracle import oracle as O
import fhe_string_b200 as F
from conftest import bench_tables
ks = O.Keyset(seed=1); eng = F.Engine(0); eng.upload_keys(ks.p, ks.ksk(), ks.bsk())
tabs = bench_tables(); rng = np.random.default_rng(5)
d_tabs = torch.from_numpy(tabs.view(np.int64)).cuda()
B = 74
m = rng.integers(0, 16, B); c = ks.encrypt(m, seed=77); l = (np.arange(B) % 8).astype(np.uint32)
d_in = torch.from_numpy(c.view(np.int64)).cuda(); d_l = torch.from_numpy(l.view(np.int32)).cuda()
d_out = torch.zeros_like(d_in)
t0 = time.time(); bad = 0
for it in range(400):
    b = int(rng.integers(1, 75))
    eng.ks_pbs_dev(d_in.data_ptr(), d_l.data_ptr(), d_tabs.data_ptr(), len(tabs), b, d_out.data_ptr())
    if it % 50 == 49:
        eng.synchronize()
        o = d_out.cpu().numpy().view(np.uint64)[:b]
        bad += int(not np.array_equal(ks.decrypt(o), tabs[l[:b], m[:b]]))
eng.synchronize()
print(f"stress ok: 400 launches in {time.time() - t0:.2f} s, wrong checks {bad}")
