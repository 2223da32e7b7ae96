#!/usr/bin/env python
"""bench_ops.py -- string-operation benches of BASELINE.json configs[2..4] (bench.py is configs[1]).

  config 3: eq / contains / find on a 64-char padded string, 8-char pattern           (1 GPU)
  config 4: split / replace / to_uppercase on a 256-char padded string (l=256, U=5)    (N GPUs,
            levels of more than 74 LUT evaluations split by rank, all-gather of result blocks)
  config 5: 1024 independent comparisons (lt, eq) of 64-char padded strings, instances sharded
            over the ranks with no exchange (fhs_prog_run_many)

Run single GPU:  python bench_ops.py
Run N GPUs:      python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 bench_ops.py
Wall-clock per call through the public API (host buffers in, host buffers out), max over ranks.
Every result is decrypted with the client key and compared with Python str semantics.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=1024)
    ap.add_argument("--skip", default="", help="comma list of configs to skip: 3,4,5")
    args = ap.parse_args()
    skip = set(args.skip.split(","))

    import torch
    from fhe_string_b200.api import ClientKey, _wrap
    from fhe_string_b200.engine import Engine, default_params
    from fhe_string_b200.program import Program

    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def bcast_np(arr, shape=None, root=0):
        """numpy u64 array from rank 0 to every rank over NCCL."""
        if world == 1:
            return arr
        if rank == root:
            meta = [tuple(arr.shape)]
        else:
            meta = [None]
        dist.broadcast_object_list(meta, root)
        t = (torch.from_numpy(arr.view(np.int64).copy()).to(dev) if rank == root
             else torch.empty(meta[0], dtype=torch.int64, device=dev))
        dist.broadcast(t, root)
        return t.cpu().numpy().view(np.uint64)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # FHS_TORCH_ALLGATHER=1: A/B against the torch.distributed callback exchange
    native_comm = os.environ.get("FHS_TORCH_ALLGATHER", "0") != "1"
    result = {"n_gpus": world, "unit": "ms wall-clock per call, max over ranks"}
    if world > 1:
        result["exchange"] = ("fhs_comm_allgather (ncclAllGather in place on the context's stream)" if native_comm
                              else "torch.distributed.all_gather_into_tensor callback")

    def setup(l, seed):
        """keys for max length l: rank 0 generates, server keys broadcast over NCCL, uploaded per GPU."""
        ck = ClientKey(l=l, seed=seed) if rank == 0 else None
        nbu = [ck.num_blocks_usize] if rank == 0 else [None]
        if dist is not None:
            dist.broadcast_object_list(nbu, 0)
        eng = Engine(local)
        prm = ck.p if rank == 0 else default_params()
        if world > 1 and native_comm:
            # the library's own communicator: keys by fhs_comm_broadcast_host, levels by fhs_comm_allgather
            eng.comm_init_from_torch()
            big = prm.glwe_dim * prm.poly_size
            shapes = (big * prm.ks_level * (prm.lwe_dim + 1), prm.lwe_dim * prm.pbs_level * (prm.glwe_dim + 1) ** 2 * prm.poly_size)
            ksk, bsk = (np.ascontiguousarray(src(), dtype=np.uint64).reshape(-1).copy() if rank == 0
                        else np.empty(n, dtype=np.uint64)
                        for src, n in zip((ck.ksk, ck.bsk) if rank == 0 else (None, None), shapes))
            eng.comm_broadcast(ksk)
            eng.comm_broadcast(bsk)
        else:
            ksk = bcast_np(ck.ksk() if rank == 0 else None)
            bsk = bcast_np(ck.bsk() if rank == 0 else None)
        eng.upload_keys(prm, ksk, bsk)
        return ck, eng, nbu[0]

    def run_op(ck, eng, nbu, op, strs, pads, scalars=(), reps=2, want=None):
        blocks = np.concatenate([ck.encrypt_str(x, p).blocks for x, p in zip(strs, pads)]) if rank == 0 else None
        blocks = bcast_np(blocks)
        if os.environ.get("FHS_BENCH_PAGEABLE", "0") != "1":   # pinned host inputs, as in bench.py's e2e leg
            pin = torch.empty(blocks.shape, dtype=torch.int64, pin_memory=True)
            pin.numpy().view(np.uint64)[...] = blocks
            blocks = pin.numpy().view(np.uint64)
        n_chars = [len(x) + 1 if p is None else p + 1 for x, p in zip(strs, pads)]
        prog = Program(op, n_chars, 0, scalars, 4, nbu)
        best, all_ms = None, []
        for _ in range(reps):
            barrier()
            t0 = time.perf_counter()
            out, st = prog.run_sharded(eng, blocks, rank, world) if world > 1 else prog.run(eng, blocks)
            torch.cuda.synchronize()
            dt = max_over_ranks((time.perf_counter() - t0) * 1e3)
            best = dt if best is None else min(best, dt)
            all_ms.append(round(dt, 1))
        rec = {"ms": best, "all_ms": all_ms, "lut_evals": prog.pbs_count, "levels": prog.depth}
        if rank == 0:
            rec["correct"] = _wrap(prog.desc, out, nbu).decrypt(ck) == want
        return rec

    # ---------------- config 3 ----------------
    if "3" not in skip or "5" not in skip:
        ck, eng, nbu = setup(255, 1)
    if "3" not in skip:
        rng = np.random.default_rng(7)
        s = "".join(chr(int(c)) for c in rng.integers(33, 127, 40))
        pat, miss = s[13:21], "~" + s[14:21]
        ops = {}
        for name, strs, pads, want in (("eq", [s, s], [64, 64], True),
                                       ("contains_hit", [s, pat], [64, None], True),
                                       ("contains_miss", [s, miss], [64, None], miss in s),
                                       ("find", [s, pat], [64, None], 13),
                                       ("split", [s, pat], [64, None], s.split(pat))):
            ops[name] = run_op(ck, eng, nbu, name.split("_")[0], strs, pads, want=want)
        result["config3_64char"] = ops

    # ---------------- config 5 ----------------
    if "5" not in skip:
        M = args.pairs
        per = (M + world - 1) // world
        lo, hi = min(M, rank * per), min(M, (rank + 1) * per)
        rng = np.random.default_rng(13 + rank)
        sweep = {}
        for op in ("eq", "lt"):
            prog = Program(op, [65, 65], 0, (), 4, nbu)
            pairs, inb = [], []
            for i in range(lo, hi):
                a = "".join(chr(int(c)) for c in rng.integers(33, 127, 40))
                b = a if i % 4 == 0 else (a[:20] + "".join(chr(int(c)) for c in rng.integers(33, 127, 20)) if i % 2 else
                                          "".join(chr(int(c)) for c in rng.integers(33, 127, 40)))
                pairs.append((a, b))
            # every rank encrypts its own slice with the (deterministic, seed 1) client key
            ckr = ck if rank == 0 else ClientKey(l=255, seed=1)
            for a, b in pairs:
                inb.append(np.concatenate([ckr.encrypt_str(a, 64).blocks, ckr.encrypt_str(b, 64).blocks]))
            inb = np.stack(inb) if inb else np.zeros((0, prog.n_inputs, eng.big), np.uint64)
            # pinned host memory, as for bench.py's e2e leg: 8.7 GB of input blocks cross PCIe
            pin = torch.empty(inb.shape, dtype=torch.int64, pin_memory=True)
            pin.numpy().view(np.uint64)[...] = inb
            inb = pin.numpy().view(np.uint64)
            all_ms = []
            for _ in range(2):      # the first call also grows the device memory pool (GBs of ciphertext pools)
                barrier()
                t0 = time.perf_counter()
                out, st = prog.run_many(eng, inb)
                torch.cuda.synchronize()
                all_ms.append(round(max_over_ranks((time.perf_counter() - t0) * 1e3), 1))
            dt = min(all_ms)
            got = [bool(ckr.decrypt_blocks(out[m, :1])[0]) for m in range(len(pairs))]
            want = [(a == b) if op == "eq" else (a < b) for a, b in pairs]
            ok = max_over_ranks(0.0 if got == want else 1.0) == 0.0
            sweep[op] = {"ms": dt, "all_ms": all_ms, "pairs": M, "lut_evals": prog.pbs_count * M, "levels": prog.depth,
                         "comparisons_per_s": M / dt * 1e3, "pbs_per_s": prog.pbs_count * M / dt * 1e3,
                         "correct": ok}
        result["config5_1024_comparisons"] = sweep
    if "3" not in skip or "5" not in skip:
        eng.close()

    # ---------------- config 4 ----------------
    if "4" not in skip:
        ck, eng, nbu = setup(256, 5)
        rng = np.random.default_rng(11)
        words = ["".join(chr(int(c)) for c in rng.integers(97, 123, rng.integers(2, 7))) for _ in range(80)]
        s = " ".join(words)[:200]
        ops = {}
        ops["to_uppercase"] = run_op(ck, eng, nbu, "to_uppercase", [s], [256], want=s.upper())
        ops["split"] = run_op(ck, eng, nbu, "split", [s, " "], [256, None], want=s.split(" "))
        pat = s[50:52]
        ops["replace"] = run_op(ck, eng, nbu, "replace", [s, pat, "c"], [256, None, None], scalars=(256,), reps=3,
                                want=s.replace(pat, "c"))
        result["config4_256char"] = ops
        eng.close()

    if rank == 0:
        print(json.dumps(result), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
