"""The reference's CPU hot loop, restated for timing.  TEST/BENCH INFRASTRUCTURE ONLY.

``run_trials`` restates ``CanonicalMonteCarlo.MCstep`` (py_mc/mc.py:170-185) with ``ising2D.trialstep`` /
``newconfig`` (examples/ising2D.py:28-52) inlined into one loop: the same arithmetic and the same draws per trial (two
``randrange``, five numpy scalar reads, one float64 dE, ``math.exp`` and ``random()`` only when dE >= 0), but without
the reference's method calls, ``dconfig`` list and attribute look-ups -- it runs about 1.2x faster than the real loop.
``bench.py`` times the UNMODIFIED reference instead (oracle/refarm.py on oracle/_ref, ``kind: "reference"``); this
restatement is only the declared fall-back (``kind: "port"``) on a machine where the reference files are absent.
``tests/test_oracle_ising.py`` checks that, with injected draws, it follows the same trajectory as the reference.
"""
import math
import random

import numpy as np


def run_trials(config, J, kT, ntrials, energy, rng=random):
    """ntrials single-site Metropolis trials at uniformly random sites; returns the tracked energy."""
    lenX, lenY = config.shape
    for _ in range(ntrials):
        x = rng.randrange(lenX)
        y = rng.randrange(lenY)
        x1 = x + 1
        if x == lenX - 1:
            x1 = 0
        y1 = y + 1
        if y == lenY - 1:
            y1 = 0
        dE = -2.0 * J * config[x, y] * (config[x - 1, y] + config[x1, y] + config[x, y - 1] + config[x, y1])
        if dE < 0.0:
            config[x, y] *= -1
            energy += dE
        else:
            if rng.random() <= math.exp(-dE / kT):
                config[x, y] *= -1
                energy += dE
    return energy


def _worker(args):
    import time
    L, J, kT, ntrials, seed = args
    rng = random.Random(seed)
    cfg = np.where(np.random.default_rng(seed).random((L, L)) >= 0.5, 1.0, -1.0)
    run_trials(cfg, J, kT, min(ntrials, 2000), 0.0, rng)  # warm the interpreter
    t0 = time.perf_counter()
    run_trials(cfg, J, kT, ntrials, 0.0, rng)
    return time.perf_counter() - t0


def time_all_cores(L, J, kT, ntrials_per_core, ncores):
    """One independent replica per core (how the reference uses cores: one MPI rank per replica,
    py_mc/mc_mpi.py:47-56).  Returns (aggregate attempts/s, wall seconds)."""
    import multiprocessing as mp
    ctx = mp.get_context("fork")
    with ctx.Pool(ncores) as pool:
        times = pool.map(_worker, [(L, J, kT, ntrials_per_core, 1000 + c) for c in range(ncores)])
    wall = max(times)  # all workers run concurrently; the slowest one bounds the job
    return ncores * ntrials_per_core / wall, wall
