"""Compile the CUDA library in-tree for sm_100a:  python -m py_mc_b200._build

nvcc cross-compiles without a GPU.  The output (py_mc_b200/_lib/libpymc_b200.so) is
git-ignored but travels to the GPU box with the repo snapshot.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")
LIBDIR = os.path.join(HERE, "_lib")
LIBPATH = os.path.join(LIBDIR, "libpymc_b200.so")

SOURCES = ["pymc_b200.cu"]
NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "-Xcompiler", "-fPIC", "-shared",
]


def _nvcc():
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found; cannot build libpymc_b200.so")
    return exe


def _deps():
    out = []
    for d in (CSRC, INCLUDE):
        for f in os.listdir(d):
            if f.endswith((".cu", ".cuh", ".h")):
                out.append(os.path.join(d, f))
    return out


def is_stale():
    if not os.path.exists(LIBPATH):
        return True
    t = os.path.getmtime(LIBPATH)
    return any(os.path.getmtime(f) > t for f in _deps())


def build(force=False, verbose=False, extra_flags=()):
    """Build libpymc_b200.so if it is missing or older than its sources.  Returns its path."""
    if not force and not is_stale():
        return LIBPATH
    os.makedirs(LIBDIR, exist_ok=True)
    import fcntl
    with open(os.path.join(LIBDIR, ".build.lock"), "w") as lock:  # ranks of one job may get here together
        fcntl.flock(lock, fcntl.LOCK_EX)
        if not force and not is_stale():  # somebody else built it while we waited
            return LIBPATH
        tmp = LIBPATH + ".tmp.%d" % os.getpid()
        cmd = [_nvcc(), *NVCC_FLAGS, *extra_flags, "-I", INCLUDE, "-o", tmp] + [os.path.join(CSRC, s) for s in SOURCES]
        if verbose:
            print(" ".join(cmd), flush=True)
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            if os.path.exists(tmp):
                os.remove(tmp)
            raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
        os.replace(tmp, LIBPATH)  # a process that has the old file mapped keeps its copy
        if verbose and (res.stdout or res.stderr):
            print(res.stdout + res.stderr)
    return LIBPATH


if __name__ == "__main__":
    flags = ["-Xptxas", "-v"] if "-v" in sys.argv else []
    print(build(force="--force" in sys.argv or "-f" in sys.argv, verbose=True, extra_flags=flags))
