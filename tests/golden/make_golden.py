"""Generates tests/golden/hotpath_golden.npz from the CPU oracle (seeded, deterministic).

The reference (Rust + tfhe-rs 0.5 from crates.io) cannot run in this container, and its own tests
hold no known-answer vectors at the PBS boundary -- "parity unpinned" at the tfhe-rs
boundary.  These vectors therefore pin the ORACLE against drift (a change to oracle/fhs_oracle.c that
alters any integer result fails tests/test_cpu_oracle.py) and give the GPU tests fixed inputs with
known keyswitch outputs.  Run:  python tests/golden/make_golden.py
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    ks = O.Keyset(seed=1)
    msgs = np.array([0, 5, 9, 15, 3, 12], dtype=np.uint64)
    cts = ks.encrypt(msgs, seed=77)
    small = ks.keyswitch(cts)
    table = (np.arange(16, dtype=np.uint64) * 7 + 3) % 16
    lut = ks.build_lut(table)
    pbs = ks.ks_pbs_batch(cts, np.zeros(len(cts), np.uint32), table[None, :])
    dec_x = np.array([0, 1, 0x8000000000000000, 0xFFFFFFFFFFFFFFFF, 0x0123456789ABCDEF, 0x7FFFC00000000000,
                      0x0008000000000000, 0xFFF7FFFFFFFFFFFF], dtype=np.uint64)
    ks_digits = np.stack([O.decompose(int(x), 3, 5) for x in dec_x])
    pbs_digits = np.stack([O.decompose(int(x), 23, 1) for x in dec_x])
    modsw = np.array([ks.modswitch(int(x)) for x in dec_x], dtype=np.uint64)
    np.savez_compressed(
        os.path.join(ROOT, "tests", "golden", "hotpath_golden.npz"),
        seed=np.array([1]), msgs=msgs, cts=cts, small=small, table=table,
        lut_body_sha=np.array([sha(lut)]), ksk_sha=np.array([sha(ks.ksk())]),
        bsk_sha=np.array([sha(ks.bsk())]), glwe_sk_sha=np.array([sha(ks.glwe_sk())]),
        pbs_decrypted=ks.decrypt(pbs), pbs_out_sha=np.array([sha(pbs)]),
        dec_x=dec_x, ks_digits=ks_digits, pbs_digits=pbs_digits, modswitch=modsw)
    print("written; pbs decrypted:", ks.decrypt(pbs), "expected", table[msgs])


if __name__ == "__main__":
    main()
