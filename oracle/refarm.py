"""Timing harness for the UNMODIFIED reference.  TEST / BENCH INFRASTRUCTURE ONLY.

The reference is imported from ``oracle/_ref`` (a git-ignored copy made by ``oracle/make_ref.py``; it travels to
the GPU box) or, in the build container, from ``/root/reference``.  Nothing here is a restatement: every leg calls
the reference's own classes --

  leg 1/2  ``CanonicalMonteCarlo.MCstep`` (py_mc/mc.py:170-185) on ``ising2D`` (examples/ising2D.py:7-52),
           one replica per host core (how the reference uses cores: one MPI rank per replica, mc_mpi.py:47-56)
  leg 3    ``TemperatureRX_MPI.run`` (py_mc/mc_mpi.py:143-209) under an in-process stand-in for mpi4py
           (tests/golden/fake_mpi.py; mpi4py / mpirun are not installed), shipped driver parameters
           (examples/ising2D_mpi.py:18-33: 10 x 10, kTs = linspace(0.01, 5, R), exchange every 4 trials)
  leg 4    ``latticegas.trialstep`` (examples/latgas.py:31-57), the flip-or-move ``trialstep`` of
           examples/latgas_move.py:31-95, and ``HC_2D.trialstep`` (examples/2D_hardcore/model_setup.py:40-62),
           each through ``MCstep``

Only two things are done to make the code run: ``np.int`` / ``np.float`` are aliased (removed from numpy 1.24,
used at mc_mpi.py:79-81) and a stub ``matplotlib.pyplot`` is registered for the hard-disc module's import.
``calc.energy`` is set by hand instead of calling ``run()``, whose pure-Python ``energy()`` loop takes minutes on
a large lattice and is not part of a trial.
"""
import os
import sys
import time
import types

import numpy as np

from .make_ref import reference_root

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
TC = 2.269185314213022


def available():
    return reference_root() is not None


def _import_reference():
    root = reference_root()
    if root is None:
        raise RuntimeError("the reference is neither under oracle/_ref nor under /root/reference")
    for p in (os.path.join(root, "examples", "2D_hardcore"), os.path.join(root, "py_mc"),
              os.path.join(root, "examples"), root):
        if p not in sys.path:
            sys.path.insert(0, p)
    if not hasattr(np, "int"):
        np.int = int
        np.float = float
    if "matplotlib" not in sys.modules:  # model_setup.py imports pyplot for its window loop only
        mpl = types.ModuleType("matplotlib")
        mpl.pyplot = types.ModuleType("matplotlib.pyplot")
        sys.modules["matplotlib"] = mpl
        sys.modules["matplotlib.pyplot"] = mpl.pyplot
    import py_mc.mc as refmc
    return root, refmc


def _time_mcsteps(calc, ntrials):
    step = calc.MCstep
    for _ in range(min(ntrials, 2000)):  # warm the interpreter
        step()
    t0 = time.perf_counter()
    for _ in range(ntrials):
        step()
    return time.perf_counter() - t0


def _build_calc(what, L, kT, seed):
    import random
    random.seed(seed)
    np.random.seed(seed % (2 ** 32))
    _, refmc = _import_reference()
    rng = np.random.default_rng(seed)
    if what == "ising":
        import ising2D as m
        cfg = m.ising2D_config(L, L)
        cfg.config = np.where(rng.random((L, L)) >= 0.5, 1.0, -1.0)
        model = m.ising2D(-1.0)
    elif what in ("latgas", "latgas_move"):
        m = __import__(what)
        cfg = m.latgas_config(L, L)
        cfg.config = np.where(rng.random((L, L)) >= 0.5, 1.0, 0.0)
        model = m.latticegas(Eads=0.0, J=-1.0, mu=0.0)
    elif what == "hc2d":
        import model_setup as m
        cfg = m.HC_2D_config(1.0, 20, 1.0 / 14.0)
        cfg.prepare_ordered()
        model = m.HC_2D(1.0 / 28.0)
    else:
        raise ValueError(what)
    calc = refmc.CanonicalMonteCarlo(model, kT, cfg)
    calc.energy = 0.0  # see module docstring
    return calc


def _worker(args):
    what, L, kT, ntrials, seed = args
    return _time_mcsteps(_build_calc(what, L, kT, seed), ntrials)


def time_mcstep(what="ising", L=1024, kT=TC, ntrials_per_core=100000, ncores=1):
    """``ncores`` independent replicas, one process each, every one timing ``ntrials_per_core`` calls of the
    reference's ``MCstep``.  Returns (aggregate trials/s, wall seconds of the slowest process)."""
    import multiprocessing as mp
    jobs = [(what, L, kT, ntrials_per_core, 1000 + c) for c in range(ncores)]
    if ncores == 1:
        times = [_worker(jobs[0])]
    else:
        with mp.get_context("fork").Pool(ncores) as pool:
            times = pool.map(_worker, jobs)
    wall = max(times)
    return ncores * ntrials_per_core / wall, wall


def _rx_rank(comm, R, L, nsteps, every, workdir):
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import fake_mpi
    import random
    fake_mpi.install(comm)
    _, refmc = _import_reference()
    import py_mc.mc_mpi as refmpi
    import ising2D as m
    random.seed(17 + comm.Get_rank())
    os.chdir(workdir)
    kTs = np.linspace(0.01, 5.0, R)
    configs = []
    for _ in range(R):
        cfg = m.ising2D_config(L, L)
        cfg.prepare_random()
        configs.append(cfg)
    rx = refmpi.TemperatureRX_MPI(comm, refmc.CanonicalMonteCarlo, m.ising2D(-1.0), configs, kTs)
    t0 = time.perf_counter()
    rx.run(nsteps, every)
    return time.perf_counter() - t0


def time_rx(nreplicas=8, L=10, nsteps=20000, every=4):
    """The reference's replica exchange, one process per replica, shipped driver parameters.  Returns
    (replica-trials/s including the exchanges, wall seconds).  The figure includes the overhead of the mpi4py
    stand-in (queues between forked processes), which a real MPI would not have."""
    import tempfile
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import fake_mpi
    workdir = tempfile.mkdtemp(prefix="pymc_ref_rx_")
    times = fake_mpi.run_ranks(nreplicas, _rx_rank, (nreplicas, L, nsteps, every, workdir))
    wall = max(times)
    return nreplicas * nsteps / wall, wall


def all_legs(seconds=3.0, cores=None):
    """BASELINE.md section 4, every leg sized for about ``seconds`` of wall time.  Returns a dict for the bench line."""
    cores = cores or os.cpu_count() or 1
    out = {"cores": cores, "reference_root": os.path.relpath(reference_root(), ROOT)}

    def sized(what, L, kT, n0=20000):
        rate, _ = time_mcstep(what, L, kT, n0, 1)
        return int(max(rate, 1e3) * seconds)

    n = sized("ising", 10, TC)
    out["ising_10x10_1core"] = time_mcstep("ising", 10, TC, n, 1)[0]
    out["ising_1024_1core"] = time_mcstep("ising", 1024, TC, n, 1)[0]
    out["ising_1024_allcores"] = time_mcstep("ising", 1024, TC, n, cores)[0]
    out["latgas_flip_256_1core"] = time_mcstep("latgas", 256, 1.0, sized("latgas", 256, 1.0), 1)[0]
    # the move branch builds full occupied / empty index lists per trial (latgas_move.py:59-60): O(N) per trial
    out["latgas_flip_or_move_64_1core"] = time_mcstep("latgas_move", 64, 1.0, sized("latgas_move", 64, 1.0, 2000), 1)[0]
    out["hard_discs_20_1core"] = time_mcstep("hc2d", 0, 1.0, sized("hc2d", 0, 1.0, 5000), 1)[0]
    R = min(cores, 8)
    rate, _ = time_rx(R, 10, 4000, 4)
    out["rx_%d_replicas_10x10_every_4" % R] = time_rx(R, 10, int(max(rate / R, 1e3) * seconds), 4)[0]
    out["unit"] = "trials/s"
    return out
