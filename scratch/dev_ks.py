"""dev: time keyswitch variants at several batch sizes and check them against each other."""
import sys, os, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fhe_string_b200.engine import Engine, default_params
from fhe_string_b200.api import ClientKey
ck = ClientKey(l=8, seed=1)
eng = Engine(0)
eng.upload_keys(ck.p, ck.ksk(), ck.bsk())
rng = np.random.default_rng(1)
for B in [int(x) for x in sys.argv[1:]] or [512, 1184, 4096]:
    cts = rng.integers(0, 2**64, size=(B, eng.big), dtype=np.uint64)
    d_in = torch.from_numpy(cts.view(np.int64)).cuda()
    d_out = torch.zeros((B, eng.small), dtype=torch.int64, device="cuda")
    st = torch.cuda.Stream()
    torch.cuda.set_stream(st)
    eng.set_stream(st.cuda_stream)
    for _ in range(3):
        eng.keyswitch_dev(d_in.data_ptr(), B, d_out.data_ptr())
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        eng.keyswitch_dev(d_in.data_ptr(), B, d_out.data_ptr())
    e1.record(); torch.cuda.synchronize()
    h = d_out.cpu().numpy().view(np.uint64)
    print(f"B={B} variant={os.environ.get('FHS_KS_VARIANT','3')} ks {e0.elapsed_time(e1)/10:.3f} ms  checksum {int(np.bitwise_xor.reduce(h.ravel())):016x}", flush=True)
