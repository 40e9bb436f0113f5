"""Build the in-tree CUDA library ``crackfl/_lib/libcrackfl.so`` for sm_100a with nvcc.

The library is a plain C-ABI shared object (``include/crackfl.h``), compiled straight with
``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo`` (nvcc cross-compiles without a
GPU) and loaded with ctypes, so it travels to the GPU box with the repo snapshot and shows
up in the process's loaded-library list.  ``python -m crackfl.ops.build`` rebuilds it.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CSRC = os.path.join(ROOT, "csrc")
LIBDIR = os.path.join(ROOT, "crackfl", "_lib")
LIBPATH = os.path.join(LIBDIR, "libcrackfl.so")
# The same optimiser / elementwise sources compiled WITHOUT --use_fast_math (IEEE division, sqrtf, logf, expf): not used
# by the engine, only by tests/test_gpu_fastmath.py to bound what the fast-math build changes in Adam, BN finalize and BCE.
IEEE_LIBPATH = os.path.join(LIBDIR, "libcrackfl_ieee.so")
IEEE_SOURCES = ("optim.cu", "elementwise.cu")
ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--use_fast_math", "-Xptxas", "-v",
              "-I", os.path.join(ROOT, "include")]


def nvcc_path() -> str:
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found")
    return p


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _digest() -> str:
    h = hashlib.sha256()
    for fn in sources() + [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cuh", ".h"))] + \
            [os.path.join(ROOT, "include", "crackfl.h")]:
        with open(fn, "rb") as f:
            h.update(os.path.relpath(fn, ROOT).encode() + f.read())  # relative: the repo is snapshotted to other paths
    h.update(" ".join(ARCH_FLAGS + [f for f in NVCC_FLAGS if not f.startswith(ROOT)]).encode())
    return h.hexdigest()


def is_current() -> bool:
    stamp = os.path.join(LIBDIR, "build.stamp")
    return os.path.exists(LIBPATH) and os.path.exists(IEEE_LIBPATH) and os.path.exists(stamp) and \
        open(stamp).read().strip() == _digest()


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile + link under an exclusive file lock; the .so is replaced atomically, so concurrent ranks of a
    torchrun job can never observe a half-written library."""
    if not force and is_current():
        return LIBPATH
    os.makedirs(LIBDIR, exist_ok=True)
    import fcntl
    with open(os.path.join(LIBDIR, ".build.lock"), "w") as lk:
        fcntl.flock(lk, fcntl.LOCK_EX)
        try:
            if not force and is_current():  # another process built it while we waited
                return LIBPATH
            return _build_locked(verbose)
        finally:
            fcntl.flock(lk, fcntl.LOCK_UN)


def _build_locked(verbose: bool) -> str:
    nvcc = nvcc_path()
    objs, logs = [], {}

    def compile_one(src):
        obj = os.path.join(LIBDIR, os.path.basename(src)[:-3] + ".o")
        cmd = [nvcc, *ARCH_FLAGS, *NVCC_FLAGS, "-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        logs[src] = r.stdout + r.stderr
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{r.stdout}\n{r.stderr}")
        return obj

    def compile_ieee(_):
        tmp = IEEE_LIBPATH + f".tmp.{os.getpid()}"
        flags = [f for f in NVCC_FLAGS if f != "--use_fast_math"]
        cmd = [nvcc, *ARCH_FLAGS, *flags, "-shared", "-o", tmp, *[os.path.join(CSRC, f) for f in IEEE_SOURCES], "-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on the IEEE variant:\n{r.stdout}\n{r.stderr}")
        os.replace(tmp, IEEE_LIBPATH)

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 4)) as ex:
        ieee_job = ex.submit(compile_ieee, None)
        objs = list(ex.map(compile_one, sources()))
        ieee_job.result()
    with open(os.path.join(LIBDIR, "ptxas.log"), "w") as f:
        for s, l in logs.items():
            f.write(f"==== {os.path.basename(s)}\n{l}\n")
    tmp = LIBPATH + f".tmp.{os.getpid()}"
    cmd = [nvcc, *ARCH_FLAGS, "-shared", "-o", tmp, *objs, "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    os.replace(tmp, LIBPATH)
    with open(os.path.join(LIBDIR, "build.stamp"), "w") as f:
        f.write(_digest())
    if verbose:
        print(open(os.path.join(LIBDIR, "ptxas.log")).read())
    return LIBPATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
