"""CPU restatement of the nearest-neighbour particle-hole exchange (Kawasaki) sweep.
TEST INFRASTRUCTURE ONLY.

Why this exists.  ``examples/latgas_move.py:56-83`` moves a particle from a random occupied site to
a random empty site anywhere on the lattice; that needs O(N) work per trial (``np.where``, lines
59-60) and its dE is defective (the removal term is overwritten and
``y1`` is stale).  The GPU path instead exchanges *nearest-neighbour* particle-hole pairs on a
conflict-free schedule of bond classes with the physically correct energy difference, using the
energy convention of ``latticegas.energy`` (latgas_move.py:17-25):

    dE = J * (occupied neighbours of the destination other than the source
              - occupied neighbours of the source other than the destination)

(the (Eads - mu) term cancels because the particle number is conserved).  Acceptance is the
Metropolis rule of ``py_mc/mc.py:175-184``.  There is no reference code to pin this against, so
it is validated by conservation laws and exact enumeration (tests/test_gpu_pairmove.py) and this
file is the bit-exact specification the kernel is compared with.

Schedule: 16 bond classes per sweep; every nearest-neighbour bond belongs to exactly one.
  class 0..7   horizontal bonds (i, j)-(i, j+1) with i % 2 == c >> 2 and j % 4 == c & 3
  class 8..15  vertical bonds   (i, j)-(i+1, j) with i % 4 == (c-8) >> 1 and j % 2 == (c-8) & 1
Two bonds of one class share no site and no neighbour, so all of them can be updated at once.
Needs lenX % 4 == 0 and lenY % 4 == 0.  The uniform of a bond is the stream value (domain 3,
step = 16*sweep + class) of its first site (i, j) in the checkerboard word layout.
"""
import numpy as np

from . import ising as o_ising
from . import philox

ALWAYS = o_ising.ALWAYS


def delta_e(J, d):
    """Energy change for d = (neighbours gained) - (neighbours lost)."""
    return J * float(d)


def threshold_table(J, kT):
    """thr[d + 3] for d = -3..3."""
    return np.array([o_ising.threshold_from_de(delta_e(J, d), kT) for d in range(-3, 4)], dtype=np.uint32)


def class_bonds(c, lenX, lenY):
    """(ia, ja, ib, jb) index arrays of all bonds of class c."""
    if c < 8:
        ri, cj = c >> 2, c & 3
        ii, jj = np.meshgrid(np.arange(ri, lenX, 2), np.arange(cj, lenY, 4), indexing="ij")
        return ii.ravel(), jj.ravel(), ii.ravel(), (jj.ravel() + 1) % lenY
    ri, cj = (c - 8) >> 1, (c - 8) & 1
    ii, jj = np.meshgrid(np.arange(ri, lenX, 4), np.arange(cj, lenY, 2), indexing="ij")
    return ii.ravel(), jj.ravel(), (ii.ravel() + 1) % lenX, jj.ravel()


def site_x32_natural(seed, replica, step, lenX, lenY):
    """x32 of every site (natural [lenX, lenY] indexing) for one step of domain 3."""
    out = np.zeros((lenX, lenY), dtype=np.uint32)
    for colour in (0, 1):
        x = o_ising.site_x32(seed, replica, 0, colour, lenX, lenY, domain=philox.DOMAIN_PAIR, step=step)
        ii, kk = np.meshgrid(np.arange(lenX), np.arange(lenY // 2), indexing="ij")
        jj = 2 * kk + ((ii + colour) & 1)
        out[ii, jj] = x
    return out


def sweep_classes(config, thr, seed, replica, sweep_index, classes=range(16)):
    """Apply the given bond classes of one sweep.  config int8 0/1, in place.
    Returns (attempted moves with a != b, accepted, change of the bond count)."""
    lenX, lenY = config.shape
    assert lenX % 4 == 0 and lenY % 4 == 0
    ntry = nacc = dB = 0
    for c in classes:
        x32 = site_x32_natural(seed, replica, 16 * sweep_index + c, lenX, lenY)
        ia, ja, ib, jb = class_bonds(c, lenX, lenY)
        ca = config[ia, ja].astype(np.int32)
        cb = config[ib, jb].astype(np.int32)
        nn = (np.roll(config, 1, 0).astype(np.int32) + np.roll(config, -1, 0) + np.roll(config, 1, 1) +
              np.roll(config, -1, 1))
        na = nn[ia, ja] - cb  # occupied neighbours of a other than b
        nb = nn[ib, jb] - ca
        movable = ca != cb
        d = np.where(ca == 1, nb - na, na - nb)  # destination minus source
        acc = movable & (x32[ia, ja] <= thr[d + 3])
        ntry += int(movable.sum())
        nacc += int(acc.sum())
        dB += int(d[acc].sum())
        config[ia[acc], ja[acc]] = cb[acc].astype(config.dtype)
        config[ib[acc], jb[acc]] = ca[acc].astype(config.dtype)
    return ntry, nacc, dB
