"""dev: fhs_prog_run latency of eq (65 x 65 chars) with pageable vs pinned host input blocks."""
import sys, time
sys.path.insert(0, '.')
import numpy as np, torch
from fhe_string_b200.api import generate_keys
from fhe_string_b200.program import Program
ck, sk = generate_keys(l=255, seed=3)
a = ck.encrypt_str("the quick brown fox jumps over the lazy dog", 64)
b = ck.encrypt_str("the quick brown fox jumps over the lazy dot", 64)
blocks = np.concatenate([a.blocks, b.blocks])
pin = torch.empty(blocks.shape, dtype=torch.int64).pin_memory()
pin.numpy().view(np.uint64)[:] = blocks
pinned = pin.numpy().view(np.uint64)
for op, nc in (("eq", [65, 65]), ("lt", [65, 65])):
    prog = Program(op, nc, 0, (), 4, ck.num_blocks_usize)
    for name, buf in (("pageable", blocks), ("pinned", pinned)):
        ts = []
        for _ in range(7):
            t0 = time.perf_counter(); out, st = prog.run(sk.engine, buf); ts.append((time.perf_counter() - t0) * 1e3)
        print(op, name, "median ms", round(sorted(ts)[3], 3), "levels", st["levels"], "MB in", round(buf.nbytes / 1e6, 1))
