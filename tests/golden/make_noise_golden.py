"""Output-noise reference of the CPU oracle over a large sample -> tests/golden/noise_golden.json.

The GPU noise tier (tests/test_gpu_parity.py::test_noise_variance_1e5_samples) compares the
variance of 102 400 GPU KS-PBS outputs with the oracle's.  192 oracle samples computed inside the
test only fix that variance to +-10 %; this script runs the oracle (same algorithm class as
tfhe-rs 0.5: f64 negacyclic FFT, 64-bit torus accumulator) over N_SAMPLES fresh encryptions of the
SURVEY.md 8(d) config-2 workload (messages uniform in [0,16), lut_id = b mod 8, bench tables) once,
offline, and commits the moments.

    python tests/golden/make_noise_golden.py [n_samples]
"""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import oracle as O  # noqa: E402
from conftest import bench_tables  # noqa: E402


def main():
    n_samples = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
    chunk = 1024
    ks = O.Keyset(seed=1)
    tables = bench_tables()
    s1 = s2 = s4 = 0.0
    worst = 0.0
    done = 0
    t0 = time.time()
    r = 0
    while done < n_samples:
        n = min(chunk, n_samples - done)
        rng = np.random.default_rng(500_000 + r)
        msgs = rng.integers(0, 16, n)
        cts = ks.encrypt(msgs, seed=500_000 + r)
        lut_id = (np.arange(n) % 8).astype(np.uint32)
        out = ks.ks_pbs_batch(cts, lut_id, tables)
        want = tables[lut_id, msgs]
        assert np.array_equal(ks.decrypt(out), want)
        err = ks.phase_error(out, want)
        s1 += float(err.sum())
        s2 += float((err * err).sum())
        s4 += float((err ** 4).sum())
        worst = max(worst, float(np.abs(err).max()))
        done += n
        r += 1
        print(f"{done}/{n_samples} samples, sigma so far {np.sqrt(s2 / done - (s1 / done) ** 2):.4e}, "
              f"{time.time() - t0:.0f} s", flush=True)
    mean = s1 / done
    var = s2 / done - mean * mean
    # relative standard error of a sample variance: sqrt((kurtosis - 1) / n)
    kurt = (s4 / done) / (var * var)
    rel_se = float(np.sqrt(max(kurt - 1.0, 0.0) / done))
    out = {
        "what": "phase error of oracle KS-PBS outputs (torus units), keyset seed 1, bench tables, lut_id = b mod 8",
        "n_samples": done,
        "mean": mean,
        "variance": var,
        "sigma": float(np.sqrt(var)),
        "kurtosis": kurt,
        "variance_rel_std_error": rel_se,
        "worst_abs": worst,
    }
    with open(os.path.join(HERE, "noise_golden.json"), "w") as f:
        json.dump(out, f, indent=1)
        f.write("\n")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
