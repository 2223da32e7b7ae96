"""GPU suite: the library's own NCCL plumbing (fhs_comm_*, include/fhs_cuda.h).
One GPU: communicator of world size 1 (NCCL resolved with dlopen, in-place all-gather and host
broadcast are identities).  Two or more GPUs: tests/multi_gpu_worker.py under torch.distributed.run."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_comm_world_of_one(keyset):
    import torch
    import fhe_string_b200 as F
    eng = F.Engine(0)
    try:
        with pytest.raises(F.FhsError):
            eng.comm_allgather_dev(1, 8)            # no communicator yet
        eng.comm_init(eng.comm_unique_id(), 0, 1)
        rank, world = (__import__("ctypes").c_int(), __import__("ctypes").c_int())
        assert eng.L.fhs_comm_info(eng.h, rank, world) == 0 and (rank.value, world.value) == (0, 1)
        a = np.arange(1000, dtype=np.uint64)
        assert np.array_equal(eng.comm_broadcast(a.copy()), a)
        t = torch.arange(4096, dtype=torch.int64, device="cuda")
        eng.comm_allgather_dev(t.data_ptr(), t.numel() * 8)
        eng.synchronize()
        assert torch.equal(t.cpu(), torch.arange(4096, dtype=torch.int64))
        with pytest.raises(F.FhsError):
            eng.comm_init(eng.comm_unique_id(), 1, 1)   # rank out of range
        eng.comm_destroy()
    finally:
        eng.close()


@pytest.mark.gpu
def test_shared_pool_without_peers_is_an_ordinary_pool(engine):
    """fhs_pool_create_shared is collective over the context's communicator; with no communicator (or a world of
    one) it must hand back a plain pool, and the peer entry points must refuse it instead of launching."""
    import ctypes
    L = engine.L
    vp, sz = ctypes.c_void_p, ctypes.c_size_t
    L.fhs_pool_create_shared.argtypes = [vp, sz, ctypes.POINTER(vp)]
    L.fhs_pool_is_shared.argtypes = [vp]
    L.fhs_pool_peer_allgather.argtypes = [vp, sz, sz, sz]
    L.fhs_pool_peer_run_done.argtypes = [vp]
    L.fhs_pool_peer_status.argtypes = [vp]
    L.fhs_pool_destroy.argtypes = [vp]
    L.fhs_pool_destroy.restype = None
    L.fhs_pool_capacity.argtypes = [vp]
    L.fhs_pool_capacity.restype = sz
    pool = vp()
    assert L.fhs_pool_create_shared(engine.h, 16, ctypes.byref(pool)) == 0
    assert L.fhs_pool_capacity(pool) == 16 and L.fhs_pool_is_shared(pool) == 0
    assert L.fhs_pool_peer_allgather(pool, 0, 2, 2) == -5          # FHS_ERR_COMM: no peer window
    assert L.fhs_pool_peer_run_done(pool) == 0 and L.fhs_pool_peer_status(pool) == 0
    assert engine.peer_pushes() == 0
    L.fhs_pool_destroy(pool)


@pytest.mark.gpu
def test_sharded_string_ops_native_exchange_two_gpus():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(ROOT, "tests", "multi_gpu_worker.py")],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "MULTI_GPU_RESULT" in r.stdout, (r.stdout[-2000:], r.stderr[-3000:])
    assert '"ok": true' in r.stdout
