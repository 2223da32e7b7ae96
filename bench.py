#!/usr/bin/env python
"""bench.py -- batched KS-PBS microbench (BASELINE.json configs[1]): 4096 LUT evaluations,
PARAM_MESSAGE_2_CARRY_2_KS_PBS, per GPU.

A "step" is one pass of the hot path (keyswitch -> blind rotation -> sample extract) over one batch
of 4096 synthetic encrypted blocks per GPU.  `value` is device-timed with inputs resident in HBM;
`e2e` goes through the host-buffer C-ABI call (fhs_ks_pbs_batch) with pinned host memory, H2D and
D2H inside the timed region.  `--impl reference` times the CPU restatement of the reference's path
(oracle/, OpenMP over ciphertexts = the rayon stand-in) on the box's host cores.

Multi-GPU: one process per GPU (torchrun), full key replica per GPU (broadcast over NCCL), the
batch is sharded by ciphertext with no data-path collective (weak scaling: 4096 per GPU).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BATCH = 4096
FLOP_PER_PBS = 194_510_848          # 742 * 262144
METRIC = "KS-PBS throughput (PBS/s), 4096 LUT evals per GPU, PARAM_MESSAGE_2_CARRY_2_KS_PBS"


def bench_tables():
    t = np.zeros((8, 16), dtype=np.uint64)
    for x in range(16):
        t[0, x] = x
        t[1, x] = x % 4
        t[2, x] = x // 4
        t[3, x] = 1 if x == 5 else 0
        t[4, x] = 1 if x >= 9 else 0
        t[5, x] = 1 if (x // 4) == (x % 4) else 0
        t[6, x] = ((x // 4) * (x % 4)) % 4
        t[7, x] = 2 if x < 8 else 0
    return t


def synth_inputs(ks, batch, seed):
    """Synthetic encrypted blocks: valid big-key LWE encryptions of uniform 4-bit messages."""
    rng = np.random.default_rng(seed)
    msgs = rng.integers(0, 16, batch)
    s = ks.glwe_sk().astype(bool)
    mask = rng.integers(0, 2**64, size=(batch, ks.big - 1), dtype=np.uint64)
    noise = np.rint(rng.normal(0.0, ks.p.glwe_std * 2.0**64, batch)).astype(np.int64).astype(np.uint64)
    with np.errstate(over="ignore"):
        body = mask[:, s].sum(axis=1, dtype=np.uint64) + msgs.astype(np.uint64) * np.uint64(ks.delta) + noise
    cts = np.concatenate([mask, body[:, None]], axis=1)
    lut_id = (np.arange(batch) % 8).astype(np.uint32)
    return msgs, cts, lut_id


def synth_inputs_product(ck, batch, seed):
    """Synthetic encrypted blocks made with the product's client key (fhs_encrypt_blocks)."""
    rng = np.random.default_rng(seed)
    msgs = rng.integers(0, 16, batch)
    cts = ck.encrypt_blocks(msgs.astype(np.uint64))
    lut_id = (np.arange(batch) % 8).astype(np.uint32)
    return msgs, cts, lut_id


class ClockSampler:
    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.samples = []
        self.reasons = set()
        self.max_mhz = None
        self._stop = threading.Event()
        self._t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                     timeout=5).stdout.strip().splitlines()
                if out:
                    f = [x.strip() for x in out[0].split(",")]
                    self.samples.append(float(f[0]))
                    self.max_mhz = float(f[1])
                    for nm, v in zip(names, f[2:]):
                        if v.lower().startswith("active"):
                            self.reasons.add(nm)
            except Exception:
                pass
            self._stop.wait(0.1)

    def start(self):
        self._t.start()

    def stop(self):
        self._stop.set()
        self._t.join(timeout=6)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def op_suite(eng, rank, world, dist, dev, quick=False):
    """BASELINE configs[2..4] through the public API at the CURRENT world size, every rank in lock step:
      config 3: eq / contains / find / split on a 64-char padded string, 8-char pattern (keys l = 255);
      config 4: to_uppercase / split(" ") / replace(2-char, "c", l=256) on a 256-char padded string
                (l = 256: 5-block FheUsize);
      config 5: 1024 independent 64-char comparisons (eq, lt) sliced by rank (fhs_prog_run_many).
    At N > 1 the ops run through fhs_prog_run_ex sharded: levels above 74 evaluations are split by rank and
    re-assembled by the library's own peer-window kernel (fhs_pool_peer_allgather: stores into the peers' pools
    over NVLink), or by ncclAllGather (fhs_comm_allgather) where peer windows are unavailable.
    Per op: `ms` = wall clock per call with HOST operands and a HOST result (H2D + levels + D2H inside, the
    reference's calling convention), `resident_ms` = operands and result device-resident (what a chain of
    ops pays), `device_ms` = CUDA events around operand copies + levels of the resident call; medians of 5
    calls after one warm-up call, max over ranks; results decrypted and compared with Python str."""
    import torch
    from fhe_string_b200.api import ClientKey, ServerKey

    def vmax(x):
        if dist is None:
            return float(x)
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(call, reps):
        ts, dv, st = [], [], None
        r = call()
        for _ in range(reps):
            barrier()
            t0 = time.perf_counter()
            r = call()
            ts.append((time.perf_counter() - t0) * 1e3)
            st = dict(sk_cur[0].last_stats)
            dv.append(st["device_ms"])
        return r, vmax(float(np.median(ts))), vmax(float(np.median(dv))), st

    sk_cur = [None]

    def run_cfg(l, rows, reps):
        # same seed => same key material as the uploaded one (keys do not depend on l); every rank makes
        # the same ciphertexts (counter-based generator), so no operand broadcast is needed
        ck = ClientKey(l=l, seed=1)
        res = {}
        for name, make, call, want in rows:
            host_sk, res_sk = ServerKey(ck, engine=eng, resident=False), ServerKey(ck, engine=eng, resident=True)

            args_host, args_res = make(ck), make(ck)     # client-side encryption stays outside the timing

            def host_call():
                sk_cur[0] = host_sk
                return call(host_sk, *args_host)         # host operands cross PCIe on every call, host result

            def res_call():
                sk_cur[0] = res_sk
                return call(res_sk, *args_res)           # uploaded at first use, result stays in HBM

            r1, ms, _, st = timed(host_call, reps)
            r2, rms, dms, st2 = timed(res_call, reps)
            ok = (r1.decrypt(ck) == want) and (r2.decrypt(ck) == want)
            res[name] = {"ms": ms, "resident_ms": rms, "device_ms": dms, "lut_evals": st["pbs"],
                         "levels": st["levels"], "sharded_levels": st2["sharded_levels"], "correct": bool(ok)}
        return res

    out = {"n_gpus": world, "unit": "ms per call: median of 5 after a warm-up call, max over ranks",
           "exchange": ("fhs_comm_allgather: ncclAllGather in place on the context's stream, levels > 74 LUT "
                        "evaluations split by rank") if world > 1 else "none (1 GPU)"}
    reps = 3 if quick else 5
    # ---- config 3 ----
    rng = np.random.default_rng(7)
    s = "".join(chr(int(c)) for c in rng.integers(33, 127, 40))
    pat = s[13:21]
    mk2 = lambda a, pa, b, pb: (lambda ck: (ck.encrypt_str(a, pa), ck.encrypt_str(b, pb)))
    phase("op suite: config 3")
    out["config3_64char"] = run_cfg(255, [
        ("eq", mk2(s, 64, s, 64), lambda sk, a, b: a.eq(sk, b), True),
        ("contains", mk2(s, 64, pat, None), lambda sk, a, b: a.contains(sk, b), True),
        ("find", mk2(s, 64, pat, None), lambda sk, a, b: a.find(sk, b), 13),
        ("split", mk2(s, 64, pat, None), lambda sk, a, b: a.split(sk, b), s.split(pat)),
    ], reps)
    # ---- config 4 ----
    rng = np.random.default_rng(11)
    words = ["".join(chr(int(c)) for c in rng.integers(97, 123, rng.integers(2, 7))) for _ in range(80)]
    s4 = " ".join(words)[:200]
    p4 = s4[50:52]
    phase("op suite: config 4")
    out["config4_256char"] = run_cfg(256, [
        ("to_uppercase", lambda ck: (ck.encrypt_str(s4, 256),), lambda sk, a: a.to_uppercase(sk), s4.upper()),
        ("split", mk2(s4, 256, " ", None), lambda sk, a, b: a.split(sk, b), s4.split(" ")),
        ("replace", lambda ck: (ck.encrypt_str(s4, 256), ck.encrypt_str(p4), ck.encrypt_str("c")),
         lambda sk, a, b, c: a.replace(sk, b, c, 256), s4.replace(p4, "c")),
    ], 2 if quick else 3)
    # ---- config 5: independent instances, no exchange ----
    phase("op suite: config 5")
    from fhe_string_b200.program import Program
    ck = ClientKey(l=255, seed=1)
    M = 1024
    per = (M + world - 1) // world
    lo, hi = min(M, rank * per), min(M, (rank + 1) * per)
    rng = np.random.default_rng(13)
    distinct = []
    for i in range(64):      # 64 distinct pairs (encryption is client-side CPU work), tiled to 1024 instances
        a = "".join(chr(int(c)) for c in rng.integers(33, 127, 40))
        b = a if i % 4 == 0 else (a[:20] + "".join(chr(int(c)) for c in rng.integers(33, 127, 20)) if i % 2 else
                                  "".join(chr(int(c)) for c in rng.integers(33, 127, 40)))
        distinct.append((a, b, np.concatenate([ck.encrypt_str(a, 64).blocks, ck.encrypt_str(b, 64).blocks])))
    mine = [distinct[i % 64] for i in range(lo, hi)]
    inb = np.stack([d[2] for d in mine]) if mine else np.zeros((0, 520, eng.big), np.uint64)
    pin = torch.empty(inb.shape, dtype=torch.int64, pin_memory=True)     # 8.7 GB / N of operand blocks
    pin.numpy().view(np.uint64)[...] = inb
    inb = pin.numpy().view(np.uint64)
    sweep = {}
    for op in ("eq", "lt"):
        prog = Program(op, [65, 65], 0, (), 4, ck.num_blocks_usize)
        ms_all = []
        for _ in range(2):          # the first call also grows the device memory pool
            barrier()
            t0 = time.perf_counter()
            res, st = prog.run_many(eng, inb)
            torch.cuda.synchronize()
            ms_all.append(vmax((time.perf_counter() - t0) * 1e3))
        got = [bool(ck.decrypt_blocks(res[m, :1])[0]) for m in range(len(mine))]
        want = [(a == b) if op == "eq" else (a < b) for a, b, _ in mine]
        ok = vmax(0.0 if got == want else 1.0) == 0.0
        dt = min(ms_all)
        sweep[op] = {"ms": dt, "pairs": M, "lut_evals": prog.pbs_count * M, "levels": prog.depth,
                     "comparisons_per_s": M / dt * 1e3, "pbs_per_s": prog.pbs_count * M / dt * 1e3, "correct": bool(ok),
                     "h2d_bytes": int(inb.nbytes) * world}
    out["config5_1024_comparisons"] = sweep
    if world > 1:
        pushes = eng.peer_pushes()
        out["exchange"] = (("fhs_pool_peer_allgather: one kernel of the library stores a rank's result blocks into every "
                            "peer's pool over NVLink peer memory (CUDA IPC) and waits on sequence numbers; "
                            f"{pushes} such exchanges in this suite") if pushes else
                           "fhs_comm_allgather: ncclAllGather in place on the context's stream (peer windows unavailable or FHS_COMM_P2P=0)") \
            + "; levels > 74 LUT evaluations split by rank"
    return out


def host_threads():
    """All host cores this process may run on.  torchrun exports OMP_NUM_THREADS=1 for every rank, which
    would make the CPU arm single-threaded under N > 1; the oracle takes an explicit thread count."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_arm_keyset():
    """The CPU arm: the oracle port compiled ON THIS HOST with -march=native (oracle/Makefile `fast`: SoA
    radix-4 FFT, per-thread scratch).  Falls back to the portable build if the host has no compiler."""
    from oracle import oracle as O
    try:
        return O, O.Keyset(seed=1, fast=True), "fast (-O3 -march=native -ffp-contract=fast, radix-4 SoA FFT)"
    except Exception as e:      # no gcc on the box: the prebuilt portable library still runs
        return O, O.Keyset(seed=1), f"portable (-march=x86-64-v2; fast build failed: {type(e).__name__})"


def cpu_baseline(tables, threads=0, target_seconds=12.0):
    """Oracle port of the reference's CPU path on a bounded sample of the same workload, all host cores."""
    O, ks, build = cpu_arm_keyset()
    threads = threads or max(ks.L.orc_max_threads(), host_threads())
    msgs, cts, lut_id = synth_inputs(ks, 2 * threads, seed=4242)
    t0 = time.perf_counter()
    ks.ks_pbs_batch(cts, lut_id, tables, threads=threads)
    dt = time.perf_counter() - t0
    per = dt / 2.0                                  # seconds per "round" of `threads` PBS
    rounds = max(2, int(target_seconds / max(per, 1e-3)))
    n = min(BATCH, rounds * threads)
    msgs, cts, lut_id = synth_inputs(ks, n, seed=4243)
    t0 = time.perf_counter()
    out = ks.ks_pbs_batch(cts, lut_id, tables, threads=threads)
    dt = time.perf_counter() - t0
    ok = bool(np.array_equal(ks.decrypt(out), tables[lut_id, msgs]))
    return {"value": n / dt, "unit": "PBS/s", "cores": int(threads), "kind": "port", "per_core": n / dt / threads,
            "isa": O.host_isa(), "build": build,
            "note": "port of the tfhe-rs 0.5 algorithm, not tfhe-rs: concrete-fft with AVX-512 is expected to be "
                    "about 1.5-2x faster per core than this build",
            "sample": f"{n} of the 4096 KS-PBS of one step, OpenMP over ciphertexts, {dt:.1f} s, "
                      f"decrypt-correct={ok}"}


def run_reference(args):
    """CPU arm: the oracle port of the reference's tfhe-rs path (the Rust reference cannot be built
    here: no cargo, tfhe crate not vendored), compiled on this host with -march=native.  Rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    O, ks, build = cpu_arm_keyset()
    tables = bench_tables()
    threads = max(ks.L.orc_max_threads(), host_threads())   # not torchrun's OMP_NUM_THREADS=1
    n = 8 * threads                                   # bounded sample of one 4096-block step
    msgs, cts, lut_id = synth_inputs(ks, n, seed=7)
    for _ in range(args.warmup):
        ks.ks_pbs_batch(cts[:threads], lut_id[:threads], tables, threads=threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = ks.ks_pbs_batch(cts, lut_id, tables, threads=threads)
    dt = time.perf_counter() - t0
    ok = bool(np.array_equal(ks.decrypt(out), tables[lut_id, msgs]))
    v = args.steps * n / dt
    sample = (f"{n} KS-PBS per step (bounded sample of the 4096-block batch), {threads} OpenMP threads, "
              f"decrypt-correct={ok}")
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "PBS/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64+u64",
        "data": "synthetic",
        "config": {"workload": "ks_pbs_microbench_4096_PARAM_MESSAGE_2_CARRY_2_KS_PBS",
                   "batch_per_gpu": BATCH, "reference_arm": "CPU port of the tfhe-rs 0.5 KS-PBS (oracle/), " + build},
        "cpu_baseline": {"value": v, "unit": "PBS/s", "cores": threads, "kind": "port", "per_core": v / threads,
                         "isa": O.host_isa(), "build": build, "sample": sample},
        "e2e": {"value": v, "unit": "PBS/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


_T0 = time.perf_counter()


def phase(msg):
    """Progress line on stderr (rank 0): where a run's wall-clock goes."""
    if int(os.environ.get("RANK", "0")) == 0:
        print(f"[bench +{time.perf_counter() - _T0:6.1f}s] {msg}", file=sys.stderr, flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-op-latency", action="store_true", help="skip the string-op latency block (ncu runs)")
    ap.add_argument("--quick-ops", action="store_true", help="fewer repetitions in the string-op block")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return

    import torch
    import fhe_string_b200 as F
    from fhe_string_b200.api import ClientKey   # product client: keygen / encrypt / decrypt (CPU)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    warm = max(args.warmup, 3)
    batch = args.batch

    # ---- keys: rank 0 generates (client side, CPU), broadcast over NCCL, upload per GPU ----
    ck = ClientKey(l=255, seed=1)    # deterministic: every rank could regenerate, but the product
    p = ck.p                         # path is "broadcast keys over NVLink" (north_star)
    if world > 1:
        if rank == 0:
            ksk_t = torch.from_numpy(ck.ksk().view(np.int64).copy()).to(dev)
            bsk_t = torch.from_numpy(ck.bsk().view(np.int64).copy()).to(dev)
        else:
            ksk_t = torch.empty(p.glwe_dim * p.poly_size * p.ks_level * (p.lwe_dim + 1), dtype=torch.int64, device=dev)
            bsk_t = torch.empty(p.lwe_dim * 4 * p.poly_size, dtype=torch.int64, device=dev)
        dist.broadcast(ksk_t, 0)
        dist.broadcast(bsk_t, 0)
        ksk = ksk_t.cpu().numpy().view(np.uint64)
        bsk = bsk_t.cpu().numpy().view(np.uint64)
        del ksk_t, bsk_t
    else:
        ksk, bsk = ck.ksk(), ck.bsk()
    phase("keys generated / broadcast")
    eng = F.Engine(local)
    eng.upload_keys(p, ksk, bsk, ksk_level_order=0)
    if world > 1:
        eng.comm_init_from_torch()      # the library's own NCCL communicator (sharded string ops below)
    # a non-default torch stream: its handle is non-NULL, so the library launches on exactly the
    # stream the torch.cuda.Event timers below are recorded on
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    eng.set_stream(stream.cuda_stream)
    tables = bench_tables()

    # ---- synthetic inputs: this rank's shard; 3 rotating input buffers (201 MB > L2) ----
    nbuf = 3
    host_in, dev_in, msgs_l = [], [], []
    for b in range(nbuf):
        msgs, cts, lut_id = synth_inputs_product(ck, batch, seed=1000 * (rank + 1) + b)
        hp = torch.from_numpy(cts.view(np.int64)).pin_memory()
        host_in.append(hp)
        dev_in.append(hp.to(dev, non_blocking=True))
        msgs_l.append(msgs)
    lut_host = torch.from_numpy(lut_id.astype(np.int32)).pin_memory()
    tab_host = torch.from_numpy(tables.view(np.int64).copy()).pin_memory()
    d_lut = lut_host.to(dev)
    d_tab = tab_host.to(dev)
    d_out = torch.empty((batch, eng.big), dtype=torch.int64, device=dev)
    host_out = torch.empty((batch, eng.big), dtype=torch.int64).pin_memory()
    torch.cuda.synchronize()

    def step_dev(i):
        eng.ks_pbs_dev(dev_in[i % nbuf].data_ptr(), d_lut.data_ptr(), d_tab.data_ptr(), len(tables),
                       batch, d_out.data_ptr())

    def step_e2e(i):
        eng.ks_pbs_ptr(host_in[i % nbuf].data_ptr(), lut_host.data_ptr(), tab_host.data_ptr(),
                       len(tables), batch, host_out.data_ptr())

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    phase("inputs encrypted and uploaded")
    # ---- device-resident timing ----
    for i in range(warm):
        step_dev(i)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = eng.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ks_ms = br_ms = 0.0
    e0.record(stream)
    for i in range(args.steps):
        step_dev(i)
    e1.record(stream)
    barrier()
    launches = eng.launch_count() - launches0
    ms_total = e0.elapsed_time(e1)
    # per-kernel durations (CUDA events recorded inside the library around each launch)
    for i in range(min(args.steps, 5)):
        step_dev(i)
        torch.cuda.synchronize()
        a, b = eng.last_kernel_ms()
        ks_ms += a
        br_ms += b
    ks_ms /= min(args.steps, 5)
    br_ms /= min(args.steps, 5)
    clocks = sampler.stop()

    # correctness of the last device step (decrypt with the client key; not timed)
    torch.cuda.synchronize()
    last = (min(args.steps, 5) - 1) % nbuf
    out_np = d_out.cpu().numpy().view(np.uint64)
    ok_dev = bool(np.array_equal(ck.decrypt_blocks(out_np), tables[lut_id, msgs_l[last]]))

    # ---- end-to-end timing through the host-buffer C ABI ----
    for i in range(warm):
        step_e2e(i)
    barrier()
    t_e2e0 = torch.cuda.Event(enable_timing=True)
    t_e2e1 = torch.cuda.Event(enable_timing=True)
    t_e2e0.record(stream)
    for i in range(args.steps):
        step_e2e(i)          # blocking call: H2D + kernels + D2H + stream sync inside
    t_e2e1.record(stream)
    barrier()
    e2e_ms = t_e2e0.elapsed_time(t_e2e1)
    out_np = host_out.numpy().view(np.uint64)
    ok_e2e = bool(np.array_equal(ck.decrypt_blocks(out_np), tables[lut_id, msgs_l[(args.steps - 1) % nbuf]]))

    # ---- max over ranks ----
    if dist is not None:
        t = torch.tensor([ms_total, e2e_ms, br_ms, ks_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total, e2e_ms, br_ms, ks_ms = [float(x) for x in t.tolist()]
        okt = torch.tensor([int(ok_dev and ok_e2e)], device=dev)
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
        ok_all = bool(okt.item())
    else:
        ok_all = ok_dev and ok_e2e

    phase("KS-PBS microbench timed (device + e2e)")
    # ---- string ops (configs 3-5) at this world size: every rank takes part ----
    ops = None
    if not args.no_op_latency:
        torch.cuda.set_stream(torch.cuda.default_stream(dev))
        eng.set_stream(0)               # the context's own stream from here on
        try:
            ops = op_suite(eng, rank, world, dist, dev, quick=args.quick_ops)
        except Exception as e:       # the headline line must survive a failure of the secondary block
            ops = {"error": f"{type(e).__name__}: {e}"}
        phase("string-op suite done")

    if rank == 0:
        total = batch * world
        value = total * args.steps / (ms_total * 1e-3)
        e2e_val = total * args.steps / (e2e_ms * 1e-3)
        fp64_peak = eng.measure_fp64_tflops()
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        hbm_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
        traffic = {}
        try:    # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu capture
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            traffic = {k: v["dram_bytes_read"] + v["dram_bytes_write"] for k, v in tj.items() if isinstance(v, dict)}
        except Exception:
            pass
        achieved_tf = batch * FLOP_PER_PBS / (br_ms * 1e-3) / 1e12
        ks_bytes = 60_866_560 + batch * (16_392 + 5_944)
        ks_gbs = ks_bytes / (ks_ms * 1e-3) / 1e9 if ks_ms > 0 else None
        # tensor-core view of the same kernel: [B x 10240] s8 digits x [10240 x 5944] u8 key bytes -> s32
        ks_ops = 2.0 * batch * 10240 * 5944
        ks_tops = ks_ops / (ks_ms * 1e-3) / 1e12 if ks_ms > 0 else None
        i8_peak = 2.0 * float(peaks.get("bf16_tflops", 2250.0))
        i8_src = ("2 x measured dense bf16 (MEASURED_PEAKS.json bf16_tflops, burst)" if "bf16_tflops" in peaks
                  else "2 x nominal dense bf16 (B200_PROFILING.md fallback)")
        line = {
            "metric": METRIC, "value": value, "unit": "PBS/s", "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64+u64", "data": "synthetic",
            "config": {"workload": "ks_pbs_microbench_4096_PARAM_MESSAGE_2_CARRY_2_KS_PBS",
                       "batch_per_gpu": batch, "luts": 8, "lwe_dim": 742, "poly_size": 2048,
                       "parallelism": f"replica x{world}, batch sharded by ciphertext, no data-path collective",
                       "l2": "3 rotating 67 MB input buffers + 67 MB output + 110 MB keys per step > 126 MB L2"},
            "correct": ok_all,
            "gpu_launches": int(launches),
            "kernels_per_step": {"keyswitch_ms": ks_ms, "blind_rotate_ms": br_ms},
            "roofline": {"bound": "fp64", "kernel": "blind_rotate_kernel", "achieved": achieved_tf,
                         "peak": fp64_peak, "unit": "TFLOP/s", "frac": achieved_tf / fp64_peak,
                         "traffic": traffic.get("blind_rotate") if batch == BATCH else None,
                         "peak_source": "measured in this run: DFMA chains (fhs_measure_fp64_tflops); "
                                        "MEASURED_PEAKS.json has no FP64 figure",
                         "algorithmic": f"{FLOP_PER_PBS} flop/PBS x {batch} PBS per launch"},
            "roofline_keyswitch": {"bound": "tensor", "kernel": "ks_digits_kernel + keyswitch_umma_kernel",
                                   "achieved": ks_tops, "peak": i8_peak, "unit": "TOP/s (int8)",
                                   "frac": (ks_tops / i8_peak) if ks_tops else None,
                                   "traffic": traffic.get("keyswitch") if batch == BATCH else None,
                                   "peak_source": i8_src,
                                   "algorithmic": f"2 x {batch} x 10240 x 5944 int8 op per launch",
                                   "hbm": {"achieved": ks_gbs, "peak": hbm_peak, "unit": "GB/s",
                                           "frac": (ks_gbs / hbm_peak) if ks_gbs else None, "peak_source": hbm_src,
                                           "algorithmic": f"{ks_bytes} B per launch (KSK once + B x (16392 + 5944))"}},
            "e2e": {"value": e2e_val, "unit": "PBS/s",
                    "h2d_bytes_per_step": int(batch * eng.big * 8 + batch * 4 + tables.size * 8) * world,
                    "d2h_bytes_per_step": int(batch * eng.big * 8) * world,
                    "ms_per_step": e2e_ms / args.steps},
            "clocks": clocks,
        }
        line["op_latency"] = ops
        line["cpu_baseline"] = None if args.no_cpu_baseline else cpu_baseline(tables)   # the only use of oracle/ here
        phase("cpu baseline done")
        print(json.dumps(line), flush=True)
    eng.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
