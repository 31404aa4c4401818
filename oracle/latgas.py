"""CPU restatement of the lattice-gas hot path (flip moves).  TEST INFRASTRUCTURE ONLY.

Follows ``examples/latgas.py``:

* ``energy``   17-25   E = J * sum c(i,j)(c(i-1,j)+c(i,j-1)) + sum(c) * (Eads - mu)
* ``density``  27-29   N = sum c
* ``trialstep`` 31-51  dE = -(2c-1) * (J*(up+down+left+right) + Eads - mu)
* ``newconfig`` 53-57  c -> 1 - c

and the same Metropolis rule / checkerboard replay as ``oracle/ising.py``.  Classes are indexed
``5*c + nn`` (c = occupancy, nn = occupied neighbours).
"""
import numpy as np

from . import ising as o_ising

ALWAYS = o_ising.ALWAYS


def delta_e(J, Eads, mu, c, nn):
    """The expression of latgas.py:48-50 with the same operation order."""
    return -(2 * float(c) - 1) * (J * float(nn) + Eads - mu)


def threshold_table(J, Eads, mu, kT):
    thr = np.zeros(10, dtype=np.uint32)
    for c in (0, 1):
        for nn in range(5):
            thr[5 * c + nn] = o_ising.threshold_from_de(delta_e(J, Eads, mu, c, nn), kT)
    return thr


def bond_sum(config):
    c = np.asarray(config).astype(np.int64)
    return int((c * (np.roll(c, 1, 0) + np.roll(c, 1, 1))).sum())


def energy(config, J, Eads, mu):
    """J*B + N*(Eads-mu): equals the reference's sequential sum for dyadic J (latgas.py:19-24)."""
    return J * float(bond_sum(config)) + float(np.asarray(config).astype(np.int64).sum()) * (Eads - mu)


def neighbour_sum(config):
    c = np.asarray(config).astype(np.int32)
    return np.roll(c, 1, 0) + np.roll(c, -1, 0) + np.roll(c, 1, 1) + np.roll(c, -1, 1)


def sweep(config, thr, seed, replica, sweep_index):
    """One checkerboard sweep of flip moves, vectorised; config int8 0/1 in place.
    Returns (n_accepted, dB, dN)."""
    lenX, lenY = config.shape
    tot = [0, 0, 0]
    for colour in (0, 1):
        x32 = o_ising.site_x32(seed, replica, sweep_index, colour, lenX, lenY)
        nn = neighbour_sum(config)
        ii, kk = np.meshgrid(np.arange(lenX), np.arange(lenY // 2), indexing="ij")
        jj = 2 * kk + ((ii + colour) & 1)
        c = config[ii, jj].astype(np.int32)
        n = nn[ii, jj]
        acc = x32 <= thr[5 * c + n]
        tot[0] += int(acc.sum())
        tot[1] += int(((1 - 2 * c) * n)[acc].sum())
        tot[2] += int((1 - 2 * c)[acc].sum())
        config[ii[acc], jj[acc]] = (1 - c[acc]).astype(config.dtype)
    return tuple(tot)


def init_config(seed, replica, lenX, lenY):
    """1 iff u >= 0.5, row-major (latgas.py:69-75), stream domain 1."""
    return (o_ising.init_config(seed, replica, lenX, lenY) > 0).astype(np.int8)
