import sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
from oracle import oracle as O
import fhe_string_b200 as F
from conftest import bench_tables
ks=O.Keyset(seed=1); eng=F.Engine(0); eng.upload_keys(ks.p, ks.ksk(), ks.bsk())
tabs=bench_tables(); rng=np.random.default_rng(0)
for B in [int(x) for x in sys.argv[1:]] or [444, 4096]:
    m=rng.integers(0,16,B); c=ks.encrypt(m, seed=10+B); l=(np.arange(B)%8).astype(np.uint32)
    for rep in range(3):
        o=eng.ks_pbs(c,l,tabs)
    kms=eng.last_kernel_ms()
    e=tabs[l,m]; d=ks.decrypt(o)
    print('B',B,'kernel ms',kms,'PBS/s', round(B/(sum(kms)/1e3)), 'correct', np.array_equal(d,e), 'noise', np.std(ks.phase_error(o,e)), flush=True)
