"""PBS output noise over reps x 4096 samples (env FHS_BSK_XF=0 -> f64 BSK transform)."""
import sys, os
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
from oracle import oracle as O
import fhe_string_b200 as F
from conftest import bench_tables
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 8
B = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
ks = O.Keyset(seed=1); eng = F.Engine(0); eng.upload_keys(ks.p, ks.ksk(), ks.bsk())
tabs = bench_tables()
l = (np.arange(B) % 8).astype(np.uint32)
s1 = s2 = 0.0; ok = True
for r in range(reps):
    rng = np.random.default_rng(1000 + r); m = rng.integers(0, 16, B); c = ks.encrypt(m, seed=1000 + r)
    o = eng.ks_pbs(c, l, tabs); e = tabs[l, m]
    ok &= bool(np.array_equal(ks.decrypt(o), e))
    err = ks.phase_error(o, e); s1 += err.sum(); s2 += (err * err).sum()
n = reps * B; var = s2 / n - (s1 / n) ** 2
print(f"[FHS_BSK_XF={os.environ.get('FHS_BSK_XF','1')}] B={B} n={n} correct={ok} sigma={var**0.5:.4e} var={var:.4e}", flush=True)
