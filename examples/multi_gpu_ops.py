#!/usr/bin/env python
"""String ops sharded over the GPUs of one box: one process per GPU (torchrun), key replica per GPU
(broadcast over NCCL), every level of at least 148*world LUT evaluations split by rank and
re-assembled with all_gather_into_tensor.  Run:
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 examples/multi_gpu_ops.py
"""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fhe_string_b200.api import ClientKey  # noqa: E402
from fhe_string_b200.engine import Engine  # noqa: E402
from fhe_string_b200.program import Program  # noqa: E402


def main():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    ck = ClientKey(l=1023, seed=5) if rank == 0 else None
    # rank 0 owns the client key; server key material and ciphertexts are broadcast
    if world > 1:
        meta = [None]
        if rank == 0:
            meta = [(ck.p.lwe_dim, ck.p.glwe_dim, ck.p.poly_size, ck.p.ks_level, ck.num_blocks_usize)]
        dist.broadcast_object_list(meta, 0)
        n, k, N, kl, nbu = meta[0]
        ksk = torch.from_numpy(ck.ksk().view(np.int64).copy()).to(dev) if rank == 0 else torch.empty(k * N * kl * (n + 1), dtype=torch.int64, device=dev)
        bsk = torch.from_numpy(ck.bsk().view(np.int64).copy()).to(dev) if rank == 0 else torch.empty(n * 4 * N, dtype=torch.int64, device=dev)
        dist.broadcast(ksk, 0)
        dist.broadcast(bsk, 0)
        ksk_np, bsk_np = ksk.cpu().numpy().view(np.uint64), bsk.cpu().numpy().view(np.uint64)
    else:
        nbu = ck.num_blocks_usize
        ksk_np, bsk_np = ck.ksk(), ck.bsk()
    eng = Engine(local)
    from fhe_string_b200.engine import default_params
    eng.upload_keys(ck.p if rank == 0 else default_params(), ksk_np, bsk_np)
    rng = np.random.default_rng(11)
    words = ["".join(chr(int(c)) for c in rng.integers(97, 123, rng.integers(2, 7))) for _ in range(60)]
    s = " ".join(words)[:200]
    results = {}
    for op, strs, pads in (("to_uppercase", [s], [256]), ("contains", [s, s[100:108]], [256, None]),
                           ("find", [s, s[100:108]], [256, None])):
        if rank == 0:
            blocks = np.concatenate([ck.encrypt_str(x, p).blocks for x, p in zip(strs, pads)])
            shape = torch.tensor(list(blocks.shape), device=dev)
        else:
            shape = torch.zeros(2, dtype=torch.int64, device=dev)
        if world > 1:
            dist.broadcast(shape, 0)
            t = torch.from_numpy(blocks.view(np.int64).copy()).to(dev) if rank == 0 else torch.empty(tuple(shape.tolist()), dtype=torch.int64, device=dev)
            dist.broadcast(t, 0)
            blocks = t.cpu().numpy().view(np.uint64)
        n_chars = [len(x) + 1 if p is None else p + 1 for x, p in zip(strs, pads)]
        prog = Program(op, n_chars, 0, (), 4, nbu)
        for rep in range(2):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            out, st = prog.run_sharded(eng, blocks, rank, world) if world > 1 else prog.run(eng, blocks)
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) * 1e3
        if rank == 0:
            from fhe_string_b200.api import _Out, _wrap
            got = _wrap(prog.desc, _Out(host=out), nbu).decrypt(ck)
            want = {"to_uppercase": s.upper(), "contains": True, "find": 100}[op]
            results[op] = {"ms": dt, "lut_evals": prog.pbs_count, "levels": prog.depth, "correct": got == want}
    if rank == 0:
        print(json.dumps({"n_gpus": world, "string": "200 chars padded to 256 (5-block FheUsize)", "ops": results}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
