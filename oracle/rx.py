"""CPU restatement of temperature replica exchange.  TEST INFRASTRUCTURE ONLY.

Follows ``TemperatureRX_MPI.Xtrial`` (py_mc/mc_mpi.py:102-140): neighbouring temperatures
(t, t+1) with (t + XCscheme) even are paired; the holder of the lower temperature decides with
``delta = (betas[t+1] - betas[t]) * (E_t - E_{t+1})``, accepting iff ``delta < 0`` or
``random() <= exp(-delta)``; on acceptance the two replicas swap temperature labels
(``rank_to_T``), configurations stay.  The reference then toggles XCscheme (mc_mpi.py:169).
The uniform of pair t in exchange round n is stream domain 2 of ``oracle/philox.py``.
"""
import math

import numpy as np

from . import philox


def xtrial(rank_to_T, energies, betas, xcscheme, seed, rnd):
    """One exchange round over all pairs; rank_to_T (int array, modified in place).
    Returns the list of accepted lower-temperature indices."""
    R = len(rank_to_T)
    T_to_rank = np.empty(R, dtype=np.int64)
    T_to_rank[np.asarray(rank_to_T)] = np.arange(R)
    accepted = []
    for t in range(xcscheme & 1, R - 1, 2):
        lo, hi = T_to_rank[t], T_to_rank[t + 1]
        delta = (betas[t + 1] - betas[t]) * (energies[lo] - energies[hi])
        ok = delta < 0.0
        if not ok:
            u = float(philox.x32_to_uniform(philox.rx_x32(seed, rnd, t)))
            ok = u <= math.exp(-delta)
        if ok:
            rank_to_T[lo], rank_to_T[hi] = t + 1, t
            accepted.append(t)
    return accepted
