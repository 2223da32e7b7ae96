import sys
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
from oracle import oracle as O
import fhe_string_b200 as F
from conftest import bench_tables
B=int(sys.argv[1]) if len(sys.argv)>1 else 444
ks=O.Keyset(seed=1); eng=F.Engine(0); eng.upload_keys(ks.p, ks.ksk(), ks.bsk())
tabs=bench_tables(); rng=np.random.default_rng(0)
m=rng.integers(0,16,B); c=ks.trivial(m); c[:, :-1] = rng.integers(0,2**64,size=(B,ks.big-1),dtype=np.uint64)
l=(np.arange(B)%8).astype(np.uint32)
for rep in range(2):
    o=eng.ks_pbs(c,l,tabs); print(eng.last_kernel_ms())
