import sys, time
sys.path.insert(0, '.')
import numpy as np
from oracle import oracle as O
import fhe_string_b200 as F
sys.path.insert(0, 'tests')
from conftest import bench_tables
t=time.time(); ks=O.Keyset(seed=1); print('keygen', time.time()-t, flush=True)
eng=F.Engine(0)
t=time.time(); eng.upload_keys(ks.p, ks.ksk(), ks.bsk()); print('upload', time.time()-t, flush=True)
tabs=bench_tables()
# LUT
print('lut exact', all(np.array_equal(eng.build_lut(t), ks.build_lut(t)) for t in tabs), flush=True)
# KS
rng=np.random.default_rng(0)
msgs=rng.integers(0,16,64); cts=ks.encrypt(msgs, seed=3)
got=eng.keyswitch(cts); want=ks.keyswitch(cts)
print('ks exact', np.array_equal(got,want), 'mismatch frac', np.mean(got!=want), flush=True)
print('ks ms', eng.last_kernel_ms())
# PBS
lut=(np.arange(64)%8).astype(np.uint32)
out=eng.ks_pbs(cts,lut,tabs)
dec=ks.decrypt(out); exp=tabs[lut,msgs]
print('pbs decrypt ok', np.array_equal(dec,exp), 'n wrong', int(np.sum(dec!=exp)))
print('dec', dec[:16]); print('exp', exp[:16])
err=ks.phase_error(out,exp); print('noise std', np.std(err), 'max', np.abs(err).max())
print('ms', eng.last_kernel_ms())
for B in (148*3, 4096):
    m=rng.integers(0,16,B); c=ks.encrypt(m, seed=10+B); l=(np.arange(B)%8).astype(np.uint32)
    for rep in range(2):
        t=time.time(); o=eng.ks_pbs(c,l,tabs); dt=time.time()-t
        print('B',B,'wall',dt,'kernel ms',eng.last_kernel_ms(), 'PBS/s (kernel)', B/ (sum(eng.last_kernel_ms())/1e3), flush=True)
    e=tabs[l,m]; d=ks.decrypt(o); print(' correct', np.array_equal(d,e), 'noise', np.std(ks.phase_error(o,e)))
