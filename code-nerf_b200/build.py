"""Build code-nerf_b200/lib/libcnrf.so for sm_100a with nvcc (cross-compiles without a GPU).

    python code-nerf_b200/build.py [--force]

The library is self-contained C ABI (include/cnrf.h): no torch, no pybind.  It is built in-tree so that the
.so travels to the GPU box with the repo snapshot.  Every .cu is compiled to its own object file (in parallel, re-used
while the source, the headers and the flags are unchanged) and the objects are linked into one shared library.
"""
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
OBJDIR = os.path.join(LIBDIR, "obj")
LIB = os.path.join(LIBDIR, "libcnrf.so")
STAMP = os.path.join(LIBDIR, "libcnrf.stamp")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--expt-relaxed-constexpr",
              "-diag-suppress", "177", "-Xcompiler", "-fPIC"] + os.environ.get("CNRF_NVCC_EXTRA", "").split()   # e.g. -DCNRF_FWD_TIMERS


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _headers():
    hs = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h")))
    return hs + [os.path.join(HERE, "..", "include", "cnrf.h")]


def _digest(paths, extra=""):
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    h.update(extra.encode())
    for p in paths:
        with open(p, "rb") as fh:
            h.update(os.path.basename(p).encode())
            h.update(fh.read())
    return h.hexdigest()


def fingerprint():
    return _digest(sources() + _headers())


def _compile(nvcc, src, hdr_digest, verbose):
    obj = os.path.join(OBJDIR, os.path.basename(src)[:-3] + ".o")
    key = _digest([src], hdr_digest)
    stamp = obj + ".stamp"
    if os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read().strip() == key:
        return obj
    cmd = [nvcc] + NVCC_FLAGS + ["-c", src, "-o", obj]
    if verbose:
        print(" ".join(cmd), flush=True)
    subprocess.run(cmd, check=True)
    with open(stamp, "w") as fh:
        fh.write(key)
    return obj


def build(force=False, verbose=True):
    os.makedirs(OBJDIR, exist_ok=True)
    fp = fingerprint()
    if not force and os.path.exists(LIB) and os.path.exists(STAMP) and open(STAMP).read().strip() == fp:
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    if force:
        for f in os.listdir(OBJDIR):
            os.remove(os.path.join(OBJDIR, f))
    hdr = _digest(_headers())
    with ThreadPoolExecutor(max_workers=max(1, min(8, os.cpu_count() or 1))) as pool:
        objs = list(pool.map(lambda s: _compile(nvcc, s, hdr, verbose), sources()))
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a"] + objs + ["-o", LIB]
    if verbose:
        print(" ".join(cmd), flush=True)
    subprocess.run(cmd, check=True)
    with open(STAMP, "w") as fh:
        fh.write(fp)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
