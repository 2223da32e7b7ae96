"""dev: per-call wall times of the API string ops (as bench.py's op_latency), optionally after a big batch."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fhe_string_b200.api import ClientKey, ServerKey
from fhe_string_b200.engine import Engine
ck = ClientKey(l=255, seed=1)
eng = Engine(0)
eng.upload_keys(ck.p, ck.ksk(), ck.bsk())
if "--big" in sys.argv:
    rng = np.random.default_rng(0)
    cts = rng.integers(0, 2**64, size=(4096, eng.big), dtype=np.uint64)
    tabs = np.tile(np.arange(16, dtype=np.uint64), (8, 1))
    for _ in range(2):
        eng.ks_pbs(cts, (np.arange(4096) % 8).astype(np.uint32), tabs)
sk = ServerKey(ck, engine=eng)
rng = np.random.default_rng(7)
s = "".join(chr(int(c)) for c in rng.integers(33, 127, 40))
pat = s[13:21]
es, ep = ck.encrypt_str(s, 64), ck.encrypt_str(pat)
for name, call in (("eq", lambda: es.eq(sk, es)), ("contains", lambda: es.contains(sk, ep)),
                   ("find", lambda: es.find(sk, ep)), ("split", lambda: es.split(sk, ep)), ("eq", lambda: es.eq(sk, es))):
    ts = []
    for _ in range(5):
        t0 = time.perf_counter(); r = call(); ts.append((time.perf_counter() - t0) * 1e3)
    print(name, [round(t, 1) for t in ts], sk.last_stats, flush=True)
