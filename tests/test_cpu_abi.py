"""CPU suite: the C-ABI library loads, exports every symbol include/fhs_cuda.h declares, and fails
loudly (no CPU fallback) when there is no CUDA device."""
import ctypes
import os
import re

import pytest

import fhe_string_b200 as F
from fhe_string_b200 import engine as E

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = set()
    inc = os.path.join(ROOT, "include")
    for fn in os.listdir(inc):
        if fn.endswith(".h"):
            src = open(os.path.join(inc, fn)).read()
            src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
            names |= set(re.findall(r"\b(fhs_[a-z0-9_]+)\s*\(", src))
    return names


def test_library_built_and_exports_header_symbols():
    L = F.load_library()
    from fhe_string_b200.program import load_host_library
    H = load_host_library()  # libfhs_string.so (include/fhs_string.h), links libfhs_cuda.so
    declared = _declared_symbols()
    assert len(declared) >= 15
    for name in declared:
        assert hasattr(L, name) or hasattr(H, name), f"{name} declared in include/*.h but not exported"
    assert declared <= set(E.EXPORTS) | declared
    assert L.fhs_version() >= 100


def test_default_params_are_param_message_2_carry_2():
    p = E.default_params()
    assert (p.lwe_dim, p.glwe_dim, p.poly_size) == (742, 1, 2048)
    assert (p.pbs_base_log, p.pbs_level, p.ks_base_log, p.ks_level) == (23, 1, 3, 5)
    assert (p.message_modulus, p.carry_modulus) == (4, 4)


def test_no_cpu_fallback():
    """Without a GPU the product must refuse to run rather than compute on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(F.FhsError) as e:
        F.Engine(0)
    assert "no CPU fallback" in str(e.value)


def test_product_does_not_use_oracle():
    """Nothing under fhe_string_b200/ may import, include, link or load anything under oracle/."""
    pkg = os.path.join(ROOT, "fhe_string_b200")
    bad = re.compile(r"from\s+oracle|import\s+oracle|oracle/|oracle\\.|fhs_oracle|libfhs_oracle|orc_[a-z_]+\s*\(")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                src = open(os.path.join(dirpath, fn), errors="ignore").read()
                m = bad.search(src)
                assert m is None, f"{os.path.join(dirpath, fn)} references the oracle: {m.group(0)!r}"
