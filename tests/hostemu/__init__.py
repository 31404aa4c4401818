"""The kernels' core headers compiled for the host (g++) so that the CPU test-suite can run the source the GPU
executes.  TEST INFRASTRUCTURE ONLY -- nothing outside tests/ imports this; the product has no CPU path."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "py_mc_b200", "csrc")
SRC = os.path.join(HERE, "hostemu.cpp")
OUTDIR = os.path.join(HERE, "_build")
LIB = os.path.join(OUTDIR, "libhostemu.so")
DEPS = [SRC, os.path.join(HERE, "cuda_host_shim.h")] + [os.path.join(CSRC, f) for f in
                                                         ("pmc_stream.cuh", "pmc_core.cuh", "pmc_pair.cuh", "pmc_persist.cuh", "pmc_pbatch.cuh", "pmc_rowstream.cuh", "pmc_pack.cuh")]


def build(force=False):
    if not force and os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(d) for d in DEPS):
        return LIB
    os.makedirs(OUTDIR, exist_ok=True)
    tmp = LIB + ".tmp.%d" % os.getpid()
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-Wall", "-Wno-unknown-pragmas", "-o", tmp, SRC]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode:
        raise RuntimeError("g++ failed:\n" + res.stdout + res.stderr)
    os.replace(tmp, LIB)
    return LIB


class HostCore:
    MODEL_ISING, MODEL_LATGAS = 0, 1

    def __init__(self):
        self.lib = C.CDLL(build())
        LL = C.POINTER(C.c_longlong)
        self.lib.hostemu_flip_sweep.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_uint32,
                                                C.c_uint64, LL, LL]
        self.lib.hostemu_persist_sweeps.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_uint32,
                                                    C.c_uint64, C.c_int, C.c_int, C.c_int, LL, LL]
        self.lib.hostemu_pbatch_sweeps.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_uint64,
                                                   C.c_uint32, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_void_p]
        self.lib.hostemu_pbatch_slab_sweeps.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_uint32,
                                                        C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        self.lib.hostemu_rowstream_sweep.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_uint32,
                                                     C.c_uint64, C.c_int, LL, LL]
        self.lib.hostemu_pair_sweep.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_uint32,
                                                C.c_uint64, C.c_uint32, LL]

    def persist_sweeps(self, model, img, thr, seed, replica, sweep0, nsweeps, nblocks, nthreads):
        """nsweeps sweeps through the persistent kernel's slab logic with nblocks emulated blocks; returns (dB, dS)."""
        assert img.dtype == np.int8 and img.flags.c_contiguous
        t = np.ascontiguousarray(thr, dtype=np.uint32)
        dB, dS = C.c_longlong(0), C.c_longlong(0)
        rc = self.lib.hostemu_persist_sweeps(model, img.ctypes.data, img.shape[0], img.shape[1], t.ctypes.data, seed,
                                             replica, sweep0, nsweeps, nblocks, nthreads, C.byref(dB), C.byref(dS))
        assert rc == 0, rc
        return dB.value, dS.value

    def pbatch_slab_sweeps(self, model, img, thr, seed, replica, sweep0, nsweeps, nslabs, nblocks, nthreads):
        """nsweeps sweeps of ONE lattice img[lenX, lenY] (in place) cut into nslabs row slabs, each run by nblocks emulated blocks
        of the batch-resident kernel's slab instantiation with its own exchange buffer (the GPUs of a slab run); returns the
        (bond sum, second observable) added up over the slabs."""
        assert img.dtype == np.int8 and img.flags.c_contiguous and img.ndim == 2
        t = np.ascontiguousarray(thr, dtype=np.uint32)
        sums = np.zeros(2, dtype=np.int64)
        rc = self.lib.hostemu_pbatch_slab_sweeps(model, img.ctypes.data, img.shape[0], img.shape[1], t.ctypes.data, seed, replica,
                                                 sweep0, nsweeps, nslabs, nblocks, nthreads, sums.ctypes.data)
        assert rc == 0, rc
        return int(sums[0]), int(sums[1])

    def rowstream_sweep(self, model, img, thr, seed, replica, sweep, pieces):
        """One sweep (in place) through the row-streaming kernel's schedule and per-thread phases with `pieces` emulated
        blocks; returns (dB, dS).  A negative return code of the emulation (a hazard it checks for) raises."""
        assert img.dtype == np.int8 and img.flags.c_contiguous
        t = np.ascontiguousarray(thr, dtype=np.uint32)
        dB, dS = C.c_longlong(0), C.c_longlong(0)
        rc = self.lib.hostemu_rowstream_sweep(model, img.ctypes.data, img.shape[0], img.shape[1], t.ctypes.data, seed, replica,
                                              sweep, pieces, C.byref(dB), C.byref(dS))
        if rc:
            raise RuntimeError("rowstream emulation: code %d" % rc)
        return dB.value, dS.value

    def rowstream_threads(self):
        return self.lib.hostemu_rowstream_threads()

    def pbatch_sweeps(self, model, imgs, thr, seed, rep_offset, sweep0, nsweeps, nblocks, nthreads):
        """nsweeps sweeps of all replicas imgs[nrep, lenX, lenY] (in place) through the batch-resident kernel's logic with
        nblocks emulated blocks; thr[nrep, nclass] gives every replica its own thresholds.  Returns sums[nrep, 2]
        (bond sum, second observable) as the kernel's reduction computes them."""
        assert imgs.dtype == np.int8 and imgs.flags.c_contiguous and imgs.ndim == 3
        t = np.ascontiguousarray(thr, dtype=np.uint32)
        assert t.shape == (imgs.shape[0], 5 if model == self.MODEL_ISING else 10)
        sums = np.zeros((imgs.shape[0], 2), dtype=np.int64)
        rc = self.lib.hostemu_pbatch_sweeps(model, imgs.ctypes.data, imgs.shape[0], imgs.shape[1], imgs.shape[2], t.ctypes.data,
                                            seed, rep_offset, sweep0, nsweeps, nblocks, nthreads, sums.ctypes.data)
        assert rc == 0, rc
        return sums

    def flip_sweep(self, model, img, thr, seed, replica, sweep):
        """One checkerboard sweep of the int8 image, in place; returns (dB, dS)."""
        assert img.dtype == np.int8 and img.flags.c_contiguous
        t = np.ascontiguousarray(thr, dtype=np.uint32)
        assert t.size == (5 if model == self.MODEL_ISING else 10)
        dB, dS = C.c_longlong(0), C.c_longlong(0)
        rc = self.lib.hostemu_flip_sweep(model, img.ctypes.data, img.shape[0], img.shape[1], t.ctypes.data, seed, replica,
                                         sweep, C.byref(dB), C.byref(dS))
        assert rc == 0, rc
        return dB.value, dS.value

    def pair_sweep(self, img, thr7, seed, replica, sweep, class_mask=0xFFFF):
        assert img.dtype == np.int8 and img.flags.c_contiguous
        t = np.ascontiguousarray(thr7, dtype=np.uint32)
        assert t.size == 7
        dB = C.c_longlong(0)
        rc = self.lib.hostemu_pair_sweep(img.ctypes.data, img.shape[0], img.shape[1], t.ctypes.data, seed, replica, sweep,
                                         class_mask, C.byref(dB))
        assert rc == 0, rc
        return dB.value
