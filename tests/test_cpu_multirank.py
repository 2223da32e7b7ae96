"""CPU suite, world_size 2 over gloo: the level-sharding plan of a string-op program is consistent
across ranks (slices are disjoint, cover every level, sharded levels are padded identically)."""
import os
import socket
import sys

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from fhe_string_b200.program import Program
    import torch
    p = Program("contains", [65, 9])
    plan = p.plan(rank, world, shard_min=148).astype(np.int64)
    gathered = [torch.zeros_like(torch.from_numpy(plan)) for _ in range(world)]
    dist.all_gather(gathered, torch.from_numpy(plan))
    ok = True
    plans = [g.numpy() for g in gathered]
    for lvl in range(plan.shape[0]):
        B, per = plans[0][lvl][0], plans[0][lvl][1]
        ok &= all(pl[lvl][0] == B and pl[lvl][1] == per for pl in plans)
        if per == 0:   # replicated: everyone computes [0, B)
            ok &= all(pl[lvl][2] == 0 and pl[lvl][3] == B for pl in plans)
        else:          # sharded: contiguous, disjoint, covering, chunk r starts at r * per
            edges = [(pl[lvl][2], pl[lvl][3]) for pl in plans]
            ok &= edges[0][0] == 0 and edges[-1][1] == B
            ok &= all(edges[r][1] == edges[r + 1][0] for r in range(world - 1))
            ok &= all(edges[r][0] == min(B, r * per) for r in range(world))
    sharded = int((plans[0][:, 1] > 0).sum())
    if rank == 0:
        q.put((bool(ok), plan.shape[0], sharded, int(p.depth)))
    dist.destroy_process_group()


def test_level_plan_two_ranks_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for pr in procs:
        pr.start()
    ok, n_levels, sharded, depth = q.get(timeout=120)
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    assert ok and n_levels == depth
    assert sharded >= 2   # the big first levels of contains(65 x 9) are split, the reductions are not


def test_default_shard_threshold_is_one_pair_wave():
    """Levels above 74 evaluations (one wave of the pair blind-rotation kernel) are split by rank; at or
    below it they are replicated.  eq(65 x 65) has levels [130, 9, 1]: on 8 ranks the first one becomes
    17 per rank, the reductions stay whole."""
    sys.path.insert(0, ROOT)
    from fhe_string_b200.program import Program
    p = Program("eq", [65, 65])
    plans = [p.plan(r, 8).astype(np.int64) for r in range(8)]
    assert [int(x) for x in plans[0][:, 0]] == [130, 9, 1]
    assert int(plans[0][0, 1]) == 17 and all(int(pl[1, 1]) == 0 and int(pl[2, 1]) == 0 for pl in plans)
    covered = sorted((int(pl[0, 2]), int(pl[0, 3])) for pl in plans)
    assert covered[0][0] == 0 and covered[-1][1] == 130
    assert all(covered[i][1] == covered[i + 1][0] for i in range(7))
    one = Program("contains", [65, 9]).plan(0, 1)
    assert all(int(x) == 0 for x in one[:, 1])           # a single rank never splits
