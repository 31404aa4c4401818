"""Table-driven generic lattice model.

The reference's plug-in point is "any object with energy / trialstep / newconfig" (py_mc/mc.py:13-40).
The GPU engine cannot call Python per site, but for the family all shipped lattice examples belong to
-- two site states, square lattice, periodic, energy built from nearest-neighbour pairs and single
sites -- everything the kernels need is a small table: dE of toggling a site as a function of
(its state, how many of its four neighbours are in the upper state).  ``nn_lattice_model`` obtains
that table by *probing the user's own ``energy``* on a tiny lattice, checks that the model really is
of this family (dE must not depend on anything else, the total energy must be affine in the pair
count and the site count), and then runs on the same sweep kernels as the built-in models.

    class my_model(model):                       # any user model, plain Python
        def energy(self, cfg): ...
    gm = nn_lattice_model(my_model(...), values=(-1, 1))
    cfg = gm.make_config(64, 64); cfg.prepare_random()
    CanonicalMonteCarlo(gm, kT, cfg).run(1000)

A model outside the family (longer range, more states, non-uniform couplings) is refused with an
error: there is no CPU fallback.
"""
import random as rand

import numpy as np

from . import _cabi, accept
from ._device import LatticeConfig
from .mc import model


class _HostConfig:
    """Minimal stand-in handed to the user's ``energy`` while probing: the attributes of the
    reference's config holders (examples/ising2D.py:56-62)."""

    def __init__(self, array):
        self.config = array
        self.lenX, self.lenY = array.shape


class generic_config(LatticeConfig):
    """Two-state configuration with arbitrary host values (lo, hi); stored as occupancy bits on the device."""

    def __init__(self, lenX, lenY, values, layout="auto", dtype=np.float64, device=0):
        super(generic_config, self).__init__(lenX, lenY, layout=layout, dtype=dtype, device=device)
        self._values = (values[0], values[1])

    def _to_int8(self, a):
        lo, hi = self._values
        return (np.asarray(a) > (lo + hi) / 2.0).astype(np.int8)

    def _from_int8(self, img):
        lo, hi = self._values
        return np.where(img > 0, hi, lo).astype(self._host.dtype)

    def prepare_random(self):
        lo, hi = self._values
        c = self.config
        for i in range(self.lenX):
            for j in range(self.lenY):
                c[i, j] = hi if rand.random() >= 0.5 else lo


class nn_lattice_model(model):
    _pmc_model = _cabi.MODEL_LATGAS  # ten classes: (state, upper-state neighbours)

    def __init__(self, user_model, values=(-1, 1), probe_size=6, rtol=1e-9):
        self.user = user_model
        self.values = (values[0], values[1])
        self.model_name = getattr(user_model, "model_name", None)
        self._probe(probe_size, rtol)

    # -- probing ---------------------------------------------------------------------------
    def _E(self, occ):
        lo, hi = self.values
        return float(self.user.energy(_HostConfig(np.where(occ > 0, hi, lo).astype(np.float64))))

    def _probe(self, n, rtol):
        if n < 5:
            raise ValueError("probe lattice must be at least 5 x 5 so that the neighbourhoods do not wrap onto themselves")
        rng = np.random.default_rng(12345)
        nbrs = [(-1, 0), (1, 0), (0, -1), (0, 1)]
        table = np.full((2, 5), np.nan)
        scale = 0.0
        for trial in range(6):  # different backgrounds and neighbour subsets must give the same dE
            bg = (rng.random((n, n)) >= 0.5).astype(np.int8)
            x, y = int(rng.integers(n)), int(rng.integers(n))
            for c in (0, 1):
                for k in range(5):
                    occ = bg.copy()
                    chosen = rng.permutation(4)[:k]
                    for m, (dx, dy) in enumerate(nbrs):
                        occ[(x + dx) % n, (y + dy) % n] = 1 if m in chosen else 0
                    occ[x, y] = c
                    e0 = self._E(occ)
                    occ[x, y] = 1 - c
                    dE = self._E(occ) - e0
                    scale = max(scale, abs(dE), abs(e0) / (n * n))
                    if np.isnan(table[c, k]):
                        table[c, k] = dE
                    elif abs(table[c, k] - dE) > rtol * max(scale, 1.0) * n * n:
                        raise TypeError(
                            "model %r is not a two-state nearest-neighbour lattice model: toggling a site in state %d "
                            "with %d upper-state neighbours changed the energy by %r in one surrounding and %r in "
                            "another. There is no CPU fallback." % (type(self.user).__name__, c, k, table[c, k], dE))
        # toggling back must undo the change
        if np.max(np.abs(table[0] + table[1])) > rtol * max(scale, 1.0) * n * n:
            raise TypeError("model %r: dE of a toggle and of its reverse do not cancel" % type(self.user).__name__)
        # snap to the exact antisymmetric values (removes the rounding of the energy differences)
        self._dE = np.stack([0.5 * (table[0] - table[1]), -0.5 * (table[0] - table[1])])
        # total energy = e0 + a*B + b*N per site count: three probes determine a, b, const/site
        empty = np.zeros((n, n), dtype=np.int8)
        full = np.ones((n, n), dtype=np.int8)
        one = empty.copy()
        one[0, 0] = 1
        Ee, Ef, E1 = self._E(empty), self._E(full), self._E(one)
        self._const_per_site = Ee / (n * n)
        self._b = E1 - Ee
        self._a = (Ef - Ee - self._b * n * n) / (2.0 * n * n)
        for _ in range(4):
            occ = (rng.random((n, n)) >= 0.5).astype(np.int8)
            B = int((occ.astype(np.int64) * (np.roll(occ, 1, 0) + np.roll(occ, 1, 1))).sum())
            want = self._E(occ)
            got = self._const_per_site * n * n + self._a * B + self._b * int(occ.sum())
            if abs(want - got) > rtol * max(abs(want), scale, 1.0) * n:
                raise TypeError("model %r: the total energy is not affine in the nearest-neighbour pair count and the "
                                "site count (%r vs %r)" % (type(self.user).__name__, want, got))

    # -- device protocol ---------------------------------------------------------------------
    def _pmc_table_key(self):
        return ("generic", id(self.user), tuple(self._dE.ravel()))

    def _pmc_thresholds(self, kT):
        return accept.table([self._dE[c, k] for c in (0, 1) for k in range(5)], kT)

    def _pmc_energy_for(self, config, bond, second):
        return self._const_per_site * (config.lenX * config.lenY) + self._a * float(bond) + self._b * float(second)

    def _pmc_energy_coeffs(self):
        return float(self._a), float(self._b)

    def _pmc_second(self, config):
        return self.upper_count(config)

    # -- model API ---------------------------------------------------------------------------
    def make_config(self, lenX, lenY, **kw):
        return generic_config(lenX, lenY, self.values, **kw)

    def energy(self, config):
        lat = config._attach(self._pmc_model)
        b, n = lat.reduce_observables()
        return self._pmc_energy_for(config, int(b[config._rep]), int(n[config._rep]))

    def upper_count(self, config):
        """Number of sites in the upper state (N for a lattice gas; (M + Nsites)/2 for spins)."""
        obs = config._device_observables()
        if obs is None:
            lat = config._attach(self._pmc_model)
            _, n = lat.reduce_observables()
            return float(n[config._rep])
        return float(obs[1])

    def trialstep(self, config, energy):
        x, y = rand.randrange(config.lenX), rand.randrange(config.lenY)
        lat = config._attach(self._pmc_model)
        c, nn = lat.propose(config._rep, [x], [y])
        return [x, y], float(self._dE[int(c[0]), int(nn[0])])

    def newconfig(self, config, dconfig):
        lo, hi = self.values
        a = config.config
        a[dconfig[0], dconfig[1]] = lo if a[dconfig[0], dconfig[1]] > (lo + hi) / 2.0 else hi
        return config
