"""dev: time + check the blind rotation at B=4096 (and small B); env FHS_BR_* select variants."""
import sys, time, os
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
from oracle import oracle as O
import fhe_string_b200 as F
from conftest import bench_tables
ks = O.Keyset(seed=1); eng = F.Engine(0); eng.upload_keys(ks.p, ks.ksk(), ks.bsk())
tabs = bench_tables(); rng = np.random.default_rng(0)
tag = " ".join(f"{k}={v}" for k, v in os.environ.items() if k.startswith("FHS_"))
for B in [int(x) for x in (sys.argv[1:] or ["4096", "148", "592"])]:
    m = rng.integers(0, 16, B); c = ks.encrypt(m, seed=10 + B); l = (np.arange(B) % 8).astype(np.uint32)
    best = 1e9
    for rep in range(3):
        o = eng.ks_pbs(c, l, tabs); best = min(best, eng.last_kernel_ms()[1])
    e = tabs[l, m]; d = ks.decrypt(o)
    print(f"[{tag}] B={B} br_ms={best:.3f} PBS/s={B/best*1e3:.0f} correct={np.array_equal(d, e)} "
          f"noise={np.std(ks.phase_error(o, e)):.3e}", flush=True)
