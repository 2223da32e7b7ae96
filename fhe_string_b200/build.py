"""Builds libfhs_cuda.so in-tree with nvcc for sm_100a (no torch extension machinery: the product
is a plain C-ABI shared library, include/fhs_cuda.h)."""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libfhs_cuda.so")
SOURCES = ["fhs_api.cu", "keyswitch.cu", "keyswitch_mma.cu", "keyswitch_umma.cu", "blind_rotate.cu", "blind_rotate_wide.cu", "blind_rotate_pair.cu", "bsk_fourier.cu", "peaks.cu", "level.cu", "comm.cu", "resident.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-O2", "--fmad=true",
]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    nvcc = _nvcc()
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(HERE, "..", "include", "fhs_cuda.h"))
    objs = []
    for src in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(CSRC, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _stale(o, [s] + headers):
            cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", s, "-o", o]
            subprocess.check_call(cmd)
    if force or _stale(LIB, objs):
        subprocess.check_call([nvcc, "-shared", "-o", LIB] + objs + ["-gencode",
                              "arch=compute_100a,code=sm_100a", "-ldl"])
    build_host(force)
    return LIB


HOST_LIB = os.path.join(HERE, "libfhs_string.so")


def build_host(force=False):
    """g++: host-side level scheduler (include/fhs_string.h), linked against libfhs_cuda.so."""
    hdir = os.path.join(CSRC, "host")
    srcs = [os.path.join(hdir, f) for f in sorted(os.listdir(hdir)) if f.endswith(".cpp")]
    deps = [os.path.join(hdir, f) for f in os.listdir(hdir)] + [
        os.path.join(HERE, "..", "include", "fhs_string.h"), os.path.join(HERE, "..", "include", "fhs_cuda.h"), LIB]
    if force or _stale(HOST_LIB, deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-o", HOST_LIB] + srcs +
                              ["-L" + HERE, "-lfhs_cuda", "-Wl,-rpath,$ORIGIN", "-lpthread"])
    return HOST_LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
