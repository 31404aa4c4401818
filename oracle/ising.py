"""CPU restatement of the 2D Ising hot path.  TEST INFRASTRUCTURE ONLY.

Follows (reference tree):

* ``examples/ising2D.py:15-22``  ``energy``   E = J * sum_ij s(i,j) * (s(i-1,j) + s(i,j-1))
* ``examples/ising2D.py:24-26``  ``magnetization``  M = sum s
* ``examples/ising2D.py:28-47``  ``trialstep``  dE = -2.0*J*s*(up + down + left + right)
* ``examples/ising2D.py:49-52``  ``newconfig``  s(x,y) *= -1
* ``py_mc/mc.py:170-185``        ``MCstep``     accept iff dE < 0.0 or random() <= exp(-dE/kT)

The reference visits uniformly random sites; the product visits sites in
checkerboard order (colour = (x + y) & 1, colour 0 first).  Sites of one colour
do not interact, so a sequential pass over them in any order equals a
simultaneous update, and the reference reproduces the same trajectory when its
``randrange`` is made to return the checkerboard order and its ``random()`` the
per-site uniform of ``oracle/philox.py`` (that is what ``tests/golden/make_golden.py``
does with the real reference).

Two restatements are kept on purpose:

* ``mcstep_site`` / ``sweep_sites_python`` -- scalar float64, the reference's own
  expression order; slow, used for small cases and as the timed CPU port.
* ``sweep_colour`` -- vectorised numpy with integer thresholds
  (``x32 <= floor(p * 2**32)``), which is what the kernels compute.

``tests/test_oracle_ising.py`` proves the two agree with each other and with the
golden fixtures made from the reference.
"""
import math

import numpy as np

from . import philox

ALWAYS = 0xFFFFFFFF


# ----------------------------------------------------------------------------
# layout helpers (shared definition with the kernels)
# ----------------------------------------------------------------------------
def words_per_row(lenY):
    """Number of 64-site words that hold one colour of one row (lenY/2 sites)."""
    return (lenY // 2 + 63) // 64


def colour_columns(i, colour, lenY):
    """Columns j of row i that have the given colour, in increasing order (k = j >> 1)."""
    q = (i + colour) & 1
    return np.arange(q, lenY, 2)


def site_x32(seed, replica, sweep, colour, lenX, lenY, domain=philox.DOMAIN_SWEEP, step=None):
    """x32 of every site of ``colour`` for this sweep: uint32 [lenX, lenY//2] indexed [i, k]."""
    nw = words_per_row(lenY)
    if step is None:
        step = 2 * sweep + colour
    widx = np.arange(lenX * nw, dtype=np.uint64)
    x = philox.word_lane_x32(seed, replica, step, widx, domain)  # [lenX*nw, 64]
    x = x.reshape(lenX, nw * 64)
    return x[:, : lenY // 2]


# ----------------------------------------------------------------------------
# energies
# ----------------------------------------------------------------------------
def bond_sum(config):
    """B = sum_ij s(i,j) * (s(i-1,j) + s(i,j-1)) as an exact integer (ising2D.py:19-21)."""
    s = np.asarray(config).astype(np.int64)
    return int((s * (np.roll(s, 1, axis=0) + np.roll(s, 1, axis=1))).sum())


def magnetization(config):
    """M = sum s (ising2D.py:24-26)."""
    return int(np.asarray(config).astype(np.int64).sum())


def energy_reference_order(config, J):
    """The reference's sequential float64 accumulation, loop order of ising2D.py:17-22."""
    c = np.asarray(config, dtype=np.float64)
    lenX, lenY = c.shape
    e = 0.0
    for i in range(lenX):
        for j in range(lenY):
            e += J * c[i, j] * (c[i - 1, j] + c[i, j - 1])
    return e


def neighbour_sum(config):
    """Sum of the four periodic neighbours of every site (ising2D.py:39-45), int8-safe."""
    s = np.asarray(config).astype(np.int32)
    return (np.roll(s, 1, 0) + np.roll(s, -1, 0) + np.roll(s, 1, 1) + np.roll(s, -1, 1))


def delta_e(J, s, nnsum):
    """dE of flipping a spin s whose neighbours sum to nnsum: the expression of ising2D.py:45."""
    return -2.0 * J * float(s) * float(nnsum)


# ----------------------------------------------------------------------------
# acceptance
# ----------------------------------------------------------------------------
def threshold_from_de(dE, kT):
    """Integer threshold equivalent to mc.py:175-184 for u = x32 * 2**-32.

    dE < 0           -> always accept (no draw consumed by the reference)
    else p = exp(-dE/kT); accept iff u <= p  <=>  x32 <= floor(p * 2**32)
    (scaling by 2**32 is exact in float64), clamped to 2**32-1 when p >= 1.
    """
    if dE < 0.0:
        return ALWAYS
    p = math.exp(-dE / kT)
    t = math.floor(p * 4294967296.0)
    return min(int(t), ALWAYS)


def threshold_table(J, kT):
    """thr[si][ni], si = (s+1)//2, ni = (nnsum+4)//2 -- 10 entries, uint32."""
    thr = np.zeros((2, 5), dtype=np.uint32)
    for si, s in enumerate((-1, 1)):
        for ni, nn in enumerate((-4, -2, 0, 2, 4)):
            thr[si, ni] = threshold_from_de(delta_e(J, s, nn), kT)
    return thr


def mcstep_site(config, J, kT, x, y, u):
    """One reference trial at site (x, y) with uniform u; config is float64, updated in place.

    Returns (accepted, dE).  Restates ising2D.trialstep + CanonicalMonteCarlo.MCstep.
    """
    lenX, lenY = config.shape
    x1 = x + 1
    if x == lenX - 1:
        x1 = 0
    y1 = y + 1
    if y == lenY - 1:
        y1 = 0
    dE = -2.0 * J * config[x, y] * (config[x - 1, y] + config[x1, y] + config[x, y - 1] + config[x, y1])
    if dE < 0.0:
        config[x, y] *= -1
        return True, dE
    if u <= math.exp(-dE / kT):
        config[x, y] *= -1
        return True, dE
    return False, dE


def sweep_sites_python(config, J, kT, seed, replica, sweep, energy):
    """One full checkerboard sweep, site by site, in float64 (the slow, reference-order path).

    config: float64 [lenX, lenY] updated in place.  Returns the tracked energy
    (``energy += dE`` on accept, mc.py:177,184).
    """
    lenX, lenY = config.shape
    for colour in (0, 1):
        u = philox.x32_to_uniform(site_x32(seed, replica, sweep, colour, lenX, lenY))
        for i in range(lenX):
            q = (i + colour) & 1
            for k in range(lenY // 2):
                acc, dE = mcstep_site(config, J, kT, i, 2 * k + q, u[i, k])
                if acc:
                    energy += dE
    return energy


def sweep_colour(config, thr, x32, colour):
    """Vectorised update of one colour.  config int8 [lenX, lenY] (updated in place).

    thr: uint32 [2,5] from ``threshold_table``; x32: uint32 [lenX, lenY//2].
    Returns (n_accepted, dB, dM) with dB/dM the exact integer changes of the bond
    sum and the magnetisation.
    """
    lenX, lenY = config.shape
    nn = neighbour_sum(config)
    ii, kk = np.meshgrid(np.arange(lenX), np.arange(lenY // 2), indexing="ij")
    jj = 2 * kk + ((ii + colour) & 1)
    s = config[ii, jj].astype(np.int32)
    n = nn[ii, jj]
    t = thr[(s + 1) // 2, (n + 4) // 2]
    acc = x32 <= t
    dB = int((-2 * s * n)[acc].sum())
    dM = int((-2 * s)[acc].sum())
    config[ii[acc], jj[acc]] = (-s[acc]).astype(config.dtype)
    return int(acc.sum()), dB, dM


def sweep(config, J, kT, seed, replica, sweep_index, thr=None):
    """One full checkerboard sweep (colour 0 then colour 1), vectorised.  Returns (nacc, dB, dM)."""
    lenX, lenY = config.shape
    if thr is None:
        thr = threshold_table(J, kT)
    tot = [0, 0, 0]
    for colour in (0, 1):
        x = site_x32(seed, replica, sweep_index, colour, lenX, lenY)
        r = sweep_colour(config, thr, x, colour)
        for a in range(3):
            tot[a] += r[a]
    return tuple(tot)


def init_config(seed, replica, lenX, lenY):
    """Hot start: +1 iff u >= 0.5 in row-major order (ising2D.py:64-70), from stream domain 1."""
    x = philox.init_x32(seed, replica, lenX * lenY)
    return np.where(x >= np.uint32(0x80000000), 1, -1).astype(np.int8).reshape(lenX, lenY)
