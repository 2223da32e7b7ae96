"""Worker of tests/test_gpu_multi.py, one process per GPU (launched by torch.distributed.run):
contains / find / split on real ciphertexts with the levels split by rank, exchanged (a) by the
library's own communicator (fhs_comm_init + fhs_comm_allgather, NULL callback) and (b) by the
torch.distributed callback; both must give bit-identical result blocks on every rank and decrypt to
the Python str answer.  Keys travel with fhs_comm_broadcast_host."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from fhe_string_b200.api import ClientKey, _wrap
    from fhe_string_b200.engine import Engine, default_params
    from fhe_string_b200.program import Program

    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = Engine(local)
    eng.comm_init_from_torch()
    ck = ClientKey(l=255, seed=1) if rank == 0 else None
    prm = ck.p if rank == 0 else default_params()
    big = prm.glwe_dim * prm.poly_size
    n_ksk = big * prm.ks_level * (prm.lwe_dim + 1)
    n_bsk = prm.lwe_dim * prm.pbs_level * (prm.glwe_dim + 1) ** 2 * prm.poly_size
    ksk = np.ascontiguousarray(ck.ksk(), dtype=np.uint64).reshape(-1).copy() if rank == 0 else np.empty(n_ksk, np.uint64)
    bsk = np.ascontiguousarray(ck.bsk(), dtype=np.uint64).reshape(-1).copy() if rank == 0 else np.empty(n_bsk, np.uint64)
    eng.comm_broadcast(ksk)
    eng.comm_broadcast(bsk)
    eng.upload_keys(prm, ksk, bsk)

    s, pat = "the quick brown fox jumps over the dog", "fox jump"
    cases = [("contains", True), ("find", s.find(pat)), ("split", s.split(pat))]
    ok = True
    report = {}
    for op, want in cases:
        n_chars = [65, len(pat) + 1]
        n_in = sum(n_chars) * 4
        blocks = np.empty((n_in, big + 1), dtype=np.uint64)
        if rank == 0:
            blocks[:] = np.concatenate([ck.encrypt_str(s, 64).blocks, ck.encrypt_str(pat, None).blocks])
        eng.comm_broadcast(blocks)
        prog = Program(op, n_chars, 0, (), 4, 4 if ck is None else ck.num_blocks_usize)
        out_native, st = prog.run_sharded(eng, blocks, rank, world)
        saved, eng.comm = eng.comm, None          # force the torch callback path
        out_torch, _ = prog.run_sharded(eng, blocks, rank, world)
        eng.comm = saved
        same = np.array_equal(out_native, out_torch)
        # every rank must hold the same result blocks
        t = torch.from_numpy(out_native.view(np.int64).copy()).cuda()
        ref = t.clone()
        dist.broadcast(ref, 0)
        same_ranks = bool(torch.equal(t, ref))
        dec_ok = True
        if rank == 0:
            got = _wrap(prog.desc, out_native, ck.num_blocks_usize).decrypt(ck)
            dec_ok = got == want
            report[op] = {"levels": st["levels"], "pbs": st["pbs"], "ms": round(st["ms"], 2), "got": repr(got)[:80]}
        flags = torch.tensor([float(same), float(same_ranks), float(dec_ok)], device="cuda")
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
        ok &= bool(flags.min().item() == 1.0)
        if rank == 0:
            report[op]["flags"] = flags.tolist()
    eng.comm_destroy()
    eng.close()
    if rank == 0:
        print("MULTI_GPU_RESULT " + json.dumps({"ok": ok, "world": world, "ops": report}))
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
