import sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
from oracle import oracle as O
import fhe_string_b200 as F
from test_gpu_strings import run_gpu
ks=O.Keyset(seed=1); eng=F.Engine(0); eng.upload_keys(ks.p, ks.ksk(), ks.bsk())
rng = np.random.default_rng(7)
s = "".join(chr(int(c)) for c in rng.integers(33, 127, 40)); pat=s[13:21]
for op,args,pads in [("eq",[s,s],[64,64]),("contains",[s,pat],[64,None]),("find",[s,pat],[64,None]),("lt",[s,s[:20]+"~"],[64,64]),("to_uppercase",[s],[64]),("trim",[s],[64]),("strip_prefix",[s,pat],[64,None])]:
    for rep in range(2):
        t=time.time(); v,st=run_gpu(eng,ks,op,args,pads); dt=time.time()-t
    print(op, repr(v)[:40], st, 'wall %.1f ms'%(dt*1e3), flush=True)
