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
    from fhe_string_b200.api import ClientKey, _Out, _wrap
    from fhe_string_b200.engine import Engine, default_params
    from fhe_string_b200.program import Program

    if os.environ.get("FHS_WORKER_DUMP_S"):       # debugging aid: where is a stuck rank?
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ["FHS_WORKER_DUMP_S"]), exit=True)
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

    # second context with the peer-window exchange switched off: same ops through ncclAllGather
    os.environ["FHS_COMM_P2P"] = "0"
    eng_nccl = Engine(local)
    del os.environ["FHS_COMM_P2P"]
    eng_nccl.comm_init_from_torch()
    eng_nccl.upload_keys(prm, ksk, bsk)

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
        pushes0 = eng.peer_pushes()
        out_native, st = prog.run_sharded(eng, blocks, rank, world)
        # the same cached execution again (run-to-run handshake of the peer window), and NCCL for reference
        repeat_same = True
        best = {"p2p": 1e9, "nccl": 1e9}
        for _ in range(4):
            again, st2 = prog.run_sharded(eng, blocks, rank, world)
            repeat_same &= bool(np.array_equal(again, out_native))
            best["p2p"] = min(best["p2p"], st2["ms"])
        pushes = eng.peer_pushes() - pushes0
        out_nccl, _ = prog.run_sharded(eng_nccl, blocks, rank, world)
        for _ in range(4):
            _, st3 = prog.run_sharded(eng_nccl, blocks, rank, world)
            best["nccl"] = min(best["nccl"], st3["ms"])
        repeat_same &= bool(np.array_equal(out_nccl, out_native)) and eng_nccl.peer_pushes() == 0
        saved, eng.comm = eng.comm, None          # force the torch callback path
        out_torch, _ = prog.run_sharded(eng, blocks, rank, world)
        eng.comm = saved
        same = np.array_equal(out_native, out_torch) and repeat_same
        # every rank must hold the same result blocks
        t = torch.from_numpy(out_native.view(np.int64).copy()).cuda()
        ref = t.clone()
        dist.broadcast(ref, 0)
        same_ranks = bool(torch.equal(t, ref))
        dec_ok = True
        if rank == 0:
            got = _wrap(prog.desc, _Out(host=out_native), ck.num_blocks_usize).decrypt(ck)
            dec_ok = got == want
            report[op] = {"levels": st["levels"], "pbs": st["pbs"], "ms": round(st["ms"], 2), "got": repr(got)[:80],
                          "sharded_levels": st.get("sharded_levels"), "peer_pushes_5_runs": pushes,
                          "wall_ms_p2p": round(best["p2p"], 3), "wall_ms_nccl": round(best["nccl"], 3)}
        flags = torch.tensor([float(same), float(same_ranks), float(dec_ok)], device="cuda")
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
        ok &= bool(flags.min().item() == 1.0)
        if rank == 0:
            report[op]["flags"] = flags.tolist()
    eng.comm_destroy()
    eng.close()
    eng_nccl.comm_destroy()
    eng_nccl.close()
    if rank == 0:
        print("MULTI_GPU_RESULT " + json.dumps({"ok": ok, "world": world, "ops": report}))
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
