"""CPU restatement of the hard-disc model.  TEST INFRASTRUCTURE ONLY.

Follows ``examples/2D_hardcore/model_setup.py``: ``HC_2D.energy`` (24-37), ``HC_2D.trialstep`` (40-62),
``HC_2D.newconfig`` (64-67), ``HC_2D_config.prepare_ordered`` (80-104), ``g_r`` (106-125), with the
Metropolis rule of ``py_mc/mc.py:170-185`` (dE = 0 is always accepted and consumes a draw; dE = inf is
accepted only if the uniform is exactly 0).  Random numbers: stream domain 4 of ``oracle/philox.py``.
"""
import numpy as np

from . import philox


def trial_draws(seed, replica, trial, ndisc, call=0):
    """(id, u53, x32) of one trial: id = (o0*ndisc)>>32, u53 from (o1,o2) as CPython's random(), x32 = o3."""
    ctr = np.array([trial & 0xFFFFFFFF, replica, (trial >> 32) & 0xFFFFFFFF,
                    (philox.DOMAIN_DISC << 28) | (call << 24)], dtype=np.uint32)
    o = [int(v) for v in philox.philox4x32_10(ctr, philox.seed_key(seed))]
    u53 = ((o[1] >> 5) * 67108864.0 + (o[2] >> 6)) * (1.0 / 9007199254740992.0)
    return (o[0] * ndisc) >> 32, u53, o[3]


def min_image(sij, L):
    return sij - np.around(sij / L) * L


def overlaps(coors, L, rdisc):
    """True iff HC_2D.energy would return inf."""
    n = len(coors)
    for a in range(n):
        for b in range(a + 1, n):
            s = min_image(coors[a] - coors[b], L)
            if not (np.sqrt((s * s).sum()) - rdisc * 2 > 0):
                return True
    return False


def trial(coors, L, rdisc, maxstep, seed, replica, t, isotropic=False):
    """One MCstep; coors float64 [N,2] in place.  Returns (id, accepted)."""
    n = len(coors)
    i, u, x32 = trial_draws(seed, replica, t, n)
    d = -1.0 + 2.0 * u
    new = coors[i].copy()
    if isotropic:
        _, u2, _ = trial_draws(seed, replica, t, n, call=1)
        new = new + maxstep * np.array([d, -1.0 + 2.0 * u2])
    else:
        new = new + maxstep * np.array(d)
    others = np.delete(coors, i, 0)
    s = min_image(others - new, L)
    free = bool(np.all(np.sqrt((s * s).sum(axis=1)) - rdisc * 2 > 0))
    acc = free or x32 == 0
    if acc:
        coors[i] = new
    return i, acc


def prepare_ordered(L, ndisc, rdisc):
    """Triangular start of HC_2D_config.prepare_ordered (model_setup.py:80-104)."""
    sx = 1.05 * rdisc * 2.0
    sy = sx * 3.0 ** 0.5 * 0.5
    maxX, maxY = int(L / sx), int(L / sy)
    maxX2, rem = divmod(ndisc, maxY)
    assert maxX2 < maxX
    c = np.zeros((ndisc, 2))
    k = 0
    for i in range(maxX2):
        for j in range(maxY):
            c[k] = (i + 0.5 * j % 1, j)
            k += 1
    for i in range(rem):
        c[k] = (maxX2 + 0.5 * i % 1, i)
        k += 1
    c[:, 0] *= sx
    c[:, 1] *= sy
    return c


def gr_counts(coors, L, dr, nbins):
    """2 * number of pairs with around(dist/dr) == b+1 (model_setup.py:116-122), b = 0..nbins-1."""
    n = len(coors)
    out = np.zeros(nbins, dtype=np.int64)
    for a in range(n):
        for b in range(a + 1, n):
            s = min_image(coors[a] - coors[b], L)
            k = np.around(np.sqrt((s * s).sum()) / dr)
            if 1 <= k <= nbins:
                out[int(k) - 1] += 2
    return out
