"""Build and load the C part of the oracle (oracle/pymc_oracle.c).  TEST INFRASTRUCTURE ONLY.

``build()`` compiles with gcc into oracle/_build/libpymc_oracle.so (git-ignored, travels to the GPU box);
``load()`` returns a small ctypes wrapper.  Nothing in the product imports this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "pymc_oracle.c")
OUTDIR = os.path.join(HERE, "_build")
LIB = os.path.join(OUTDIR, "libpymc_oracle.so")


def build(force=False, verbose=False):
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(SRC):
        return LIB
    os.makedirs(OUTDIR, exist_ok=True)
    cmd = ["gcc", "-O2", "-std=c99", "-fPIC", "-shared", "-o", LIB, SRC]
    if verbose:
        print(" ".join(cmd), flush=True)
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode:
        raise RuntimeError("gcc failed:\n" + res.stdout + res.stderr)
    return LIB


class COracle:
    def __init__(self):
        self.lib = C.CDLL(build())
        self.lib.pymc_oracle_ising_sweep.restype = C.c_longlong
        self.lib.pymc_oracle_ising_sweep.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_uint32,
                                                    C.c_uint64, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]
        self.lib.pymc_oracle_ising_observables.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_longlong),
                                                          C.POINTER(C.c_longlong)]
        self.lib.pymc_oracle_ising_init.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_uint32]

    def init(self, lenX, lenY, seed, replica=0):
        cfg = np.empty((lenX, lenY), dtype=np.int8)
        self.lib.pymc_oracle_ising_init(cfg.ctypes.data, lenX, lenY, seed, replica)
        return cfg

    def sweep(self, cfg, thr5, seed, replica, sweep_index):
        """In place; thr5[a] for a = aligned neighbours 0..4.  Returns (nacc, dB, dM)."""
        assert cfg.dtype == np.int8 and cfg.flags.c_contiguous
        t = np.ascontiguousarray(thr5, dtype=np.uint32)
        dB, dM = C.c_longlong(0), C.c_longlong(0)
        n = self.lib.pymc_oracle_ising_sweep(cfg.ctypes.data, cfg.shape[0], cfg.shape[1], t.ctypes.data, seed, replica,
                                             sweep_index, C.byref(dB), C.byref(dM))
        return int(n), dB.value, dM.value

    def observables(self, cfg):
        B, M = C.c_longlong(0), C.c_longlong(0)
        self.lib.pymc_oracle_ising_observables(cfg.ctypes.data, cfg.shape[0], cfg.shape[1], C.byref(B), C.byref(M))
        return B.value, M.value


def load():
    return COracle()
