"""dev tool (2+ GPUs, under torch.distributed.run): split levels exchanged by the peer-window kernel vs
ncclAllGather, device ms per op (CUDA events around operand copies + levels), same process, two contexts.
usage: torchrun ... tools/dev_peer.py [op:chars:pattern_chars ...]"""
import os, sys, time, json
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..'))
import numpy as np, torch, torch.distributed as dist
from fhe_string_b200.api import ClientKey
from fhe_string_b200.engine import Engine, default_params
from fhe_string_b200.program import Program
rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
import faulthandler
faulthandler.dump_traceback_later(int(os.environ.get("FHS_WORKER_DUMP_S", "150")), exit=True)
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
eng = Engine(local); eng.comm_init_from_torch()
os.environ["FHS_COMM_P2P"] = "0"
eng_n = Engine(local); eng_n.comm_init_from_torch()
del os.environ["FHS_COMM_P2P"]
ck = ClientKey(l=255, seed=1) if rank == 0 else None
prm = ck.p if rank == 0 else default_params()
big = prm.glwe_dim * prm.poly_size
ksk = np.ascontiguousarray(ck.ksk(), dtype=np.uint64).reshape(-1).copy() if rank == 0 else np.empty(big * prm.ks_level * (prm.lwe_dim + 1), np.uint64)
bsk = np.ascontiguousarray(ck.bsk(), dtype=np.uint64).reshape(-1).copy() if rank == 0 else np.empty(prm.lwe_dim * prm.pbs_level * (prm.glwe_dim + 1) ** 2 * prm.poly_size, np.uint64)
eng.comm_broadcast(ksk); eng.comm_broadcast(bsk)
eng.upload_keys(prm, ksk, bsk); eng_n.upload_keys(prm, ksk, bsk)
res = {}
for spec in sys.argv[1:] or ["to_uppercase:257", "contains:65:9", "split:65:9"]:
    op, *nc = spec.split(":")
    n_chars = [int(x) for x in nc]
    n_in = sum(n_chars) * 4
    blocks = np.zeros((n_in, big + 1), dtype=np.uint64)     # trivial zero ciphertexts: timing only
    prog = Program(op, n_chars, 0, (), 4, 5)
    out = np.empty((prog.n_outputs, big + 1), dtype=np.uint64)
    r = {}
    for name, e in (("p2p", eng), ("nccl", eng_n), ("p2p_again", eng)):
        best = 1e9
        for it in range(6):
            dist.barrier()
            st = prog.run_ex(e, [("host", blocks)], out_host=out, rank=rank, world=world)
            if it:
                best = min(best, st["device_ms"])
        t = torch.tensor([best], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        r[name] = round(float(t.item()), 3)
    r.update(levels=st["levels"], sharded=st["sharded_levels"], pushes=eng.peer_pushes(), nccl_pushes=eng_n.peer_pushes())
    res[spec] = r
if rank == 0:
    print("PEER_AB " + json.dumps({"world": world, "device_ms_max_over_ranks": res}), flush=True)
eng.comm_destroy(); eng.close(); eng_n.comm_destroy(); eng_n.close()
dist.barrier(); dist.destroy_process_group()
