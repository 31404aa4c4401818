"""In-process stand-in for ``mpi4py`` so the UNMODIFIED ``py_mc/mc_mpi.py`` can run here.

TEST INFRASTRUCTURE ONLY.  mpi4py and mpirun are not installed in the build container; the
reference needs only a handful of calls (py_mc/mc_mpi.py:27-31,70-71,106-138,204): ranks are forked
processes, messages travel through one multiprocessing queue per destination rank.
"""
import multiprocessing as mp
import sys
import types

import numpy as np


class _Datatype:
    def __init__(self, name):
        self.name = name


class _Op:
    def __init__(self, name):
        self.name = name


def _payload(buf):
    if isinstance(buf, (list, tuple)):  # [data, count, datatype]
        buf = buf[0]
    return np.array(buf, copy=True)


def _target(buf):
    if isinstance(buf, (list, tuple)):
        buf = buf[0]
    return buf


def _rank_of(x):
    return int(np.asarray(x).ravel()[0])


class Comm:
    def __init__(self, rank, size, queues):
        self.rank, self.size, self.queues = rank, size, queues
        self.stash = []
        self.coll = 0

    def Get_rank(self):
        return self.rank

    def Get_size(self):
        return self.size

    def Split(self, color=0, key=0):
        return self

    def Free(self):
        pass

    def _put(self, dest, tag, data):
        self.queues[dest].put((self.rank, tag, data))

    def _get(self, source, tag):
        for n, (s, t, d) in enumerate(self.stash):
            if s == source and t == tag:
                del self.stash[n]
                return d
        while True:
            s, t, d = self.queues[self.rank].get(timeout=600)
            if s == source and t == tag:
                return d
            self.stash.append((s, t, d))

    def Send(self, buf, dest, tag=0):
        self._put(_rank_of(dest), ("p2p", tag), _payload(buf))

    def Recv(self, buf, source, tag=0):
        data = self._get(_rank_of(source), ("p2p", tag))
        tgt = _target(buf)
        tgt[...] = np.asarray(data).reshape(np.shape(tgt))

    def Allgather(self, sendbuf, recvbuf):
        self.coll += 1
        tag = ("ag", self.coll)
        data = _payload(sendbuf)
        for r in range(self.size):
            if r != self.rank:
                self._put(r, tag, data)
        out = _target(recvbuf)
        flat = np.asarray(out).reshape(self.size, -1)
        for r in range(self.size):
            flat[r] = (data if r == self.rank else self._get(r, tag)).reshape(-1)

    def Allreduce(self, sendbuf, recvbuf, op=None):
        self.coll += 1
        tag = ("ar", self.coll)
        data = _payload(sendbuf)
        for r in range(self.size):
            if r != self.rank:
                self._put(r, tag, data)
        total = np.zeros_like(data)
        for r in range(self.size):  # rank order: same summation order on every rank
            total = total + (data if r == self.rank else self._get(r, tag))
        _target(recvbuf)[...] = total


def _reduce(self, sendbuf, recvbuf, op=None, root=0):
    """Sum to root in rank order (examples/analyze_ising2D_mpi.py:34,44)."""
    self.coll += 1
    tag = ("rd", self.coll)
    data = _payload(sendbuf)
    if self.rank != root:
        self._put(root, tag, data)
        return
    total = np.zeros_like(data)
    for r in range(self.size):
        total = total + (data if r == root else self._get(r, tag))
    _target(recvbuf)[...] = total


Comm.Reduce = _reduce


def install(comm):
    """Register a fake ``mpi4py`` package whose COMM_WORLD is ``comm``."""
    MPI = types.SimpleNamespace(COMM_WORLD=comm, INT=_Datatype("INT"), DOUBLE=_Datatype("DOUBLE"), SUM=_Op("SUM"))
    pkg = types.ModuleType("mpi4py")
    pkg.MPI = MPI
    sys.modules["mpi4py"] = pkg
    return MPI


def run_ranks(nranks, target, args=()):
    """Fork ``nranks`` processes running ``target(comm, *args)``; returns their results by rank."""
    ctx = mp.get_context("fork")
    queues = [ctx.Queue() for _ in range(nranks)]
    results = ctx.Queue()

    def body(rank):
        try:
            comm = Comm(rank, nranks, queues)
            results.put((rank, "ok", target(comm, *args)))
        except BaseException as e:  # noqa: BLE001
            import traceback
            results.put((rank, "error", traceback.format_exc()))

    procs = [ctx.Process(target=body, args=(r,)) for r in range(nranks)]
    for p in procs:
        p.start()
    out = {}
    for _ in range(nranks):
        rank, status, val = results.get(timeout=900)
        if status != "ok":
            for p in procs:
                p.terminate()
            raise RuntimeError("rank %d failed:\n%s" % (rank, val))
        out[rank] = val
    for p in procs:
        p.join()
    return [out[r] for r in range(nranks)]
