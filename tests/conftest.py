import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")


@pytest.fixture(scope="session")
def keyset():
    """Oracle keyset (CPU, test infrastructure): seed 1 as in SURVEY.md 8(d)."""
    from oracle import oracle as O
    return O.Keyset(seed=1)


@pytest.fixture(scope="session")
def engine(keyset):
    """The product: libfhs_cuda.so context on cuda:0 with the oracle-generated server keys."""
    import fhe_string_b200 as F
    eng = F.Engine(0)
    eng.upload_keys(keyset.p, keyset.ksk(), keyset.bsk(), ksk_level_order=0)
    yield eng
    eng.close()


# The eight LUTs of the KS-PBS microbench.
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
        t[7, x] = 2 if x < 8 else 0   # "sign" style table, also exercised with the padding bit set
    return t
