"""Drop-in for ``examples/latgas.py`` (and the flip branch of ``examples/latgas_move.py``): the 2D
lattice gas with nearest-neighbour pair interaction, adsorption energy and chemical potential.

* ``latticegas``     (latgas.py:7-57)   energy, density, trialstep -> ([x, y], dE), newconfig
* ``latgas_config``  (latgas.py:61-86)  lenX, lenY, config (0/1), prepare_random, __str__
* ``observer``       (latgas.py:88-97)  logs (energy, N) and keeps the density series

Reference defects not reproduced: the shipped ``observer.__init__`` forgets
``super().__init__()`` and so crashes on first use -- this one initialises its base; the shipped
``__main__`` passes the observer in the ``print_frequency`` slot -- ``run`` here accepts that form.
"""
import random as rand

import numpy as np

from . import _cabi, accept
from ._device import LatticeConfig
from .mc import model, observer_base


class latticegas(model):
    """E = J * sum c(i,j)(c(i-1,j)+c(i,j-1)) + (Eads - mu) * sum c  (latgas.py:17-25)."""

    model_name = "latticegas"
    _pmc_model = _cabi.MODEL_LATGAS

    def __init__(self, Eads, J=0, mu=0):
        self.J = J
        self.Eads = Eads
        self.mu = mu

    # -- device protocol -----------------------------------------------------------------
    def _pmc_table_key(self):
        return ("latticegas", self.J, self.Eads, self.mu)

    def _pmc_dE(self, c, nnsum):
        """dE of toggling a site with occupancy c and nnsum occupied neighbours (latgas.py:48-50)."""
        return -(2 * float(c) - 1) * (self.J * float(nnsum) + self.Eads - self.mu)

    def _pmc_thresholds(self, kT):
        """Acceptance thresholds indexed 5*c + nn."""
        return accept.table([self._pmc_dE(c, nn) for c in (0, 1) for nn in range(5)], kT)

    def _pmc_energy(self, bond, second):
        return self.J * float(bond) + float(second) * (self.Eads - self.mu)

    def _pmc_second(self, config):
        return self.density(config)

    def _pmc_energy_coeffs(self):
        return float(self.J), float(self.Eads - self.mu)

    # -- reference API -------------------------------------------------------------------
    def energy(self, latgas_config):
        lat = latgas_config._attach(self._pmc_model)
        b, n = lat.reduce_observables()
        return self._pmc_energy(int(b[latgas_config._rep]), int(n[latgas_config._rep]))

    def density(self, latgas_config):
        """Number of particles (latgas.py:27-29)."""
        obs = latgas_config._device_observables()
        if obs is None:
            lat = latgas_config._attach(self._pmc_model)
            _, n = lat.reduce_observables()
            return float(n[latgas_config._rep])
        return float(obs[1])

    def trialstep(self, latgas_config, energy):
        x = rand.randrange(latgas_config.lenX)
        y = rand.randrange(latgas_config.lenY)
        lat = latgas_config._attach(self._pmc_model)
        c, nn = lat.propose(latgas_config._rep, [x], [y])
        return [x, y], self._pmc_dE(c[0], nn[0])

    def newconfig(self, latgas_config, dconfig):
        c = latgas_config.config
        c[dconfig[0], dconfig[1]] = 1 - c[dconfig[0], dconfig[1]]
        if len(dconfig) == 4:  # particle move of latgas_move.py:91-93
            c[dconfig[2], dconfig[3]] = 1 - c[dconfig[2], dconfig[3]]
        return latgas_config


HARD_J = 1.0e30
"""Pair energy standing in for "infinite" nearest-neighbour repulsion: exp(-HARD_J/kT) underflows to exactly 0.0
for any kT < 1e27, while HARD_J * 0 neighbours stays 0.0 (a true ``inf`` would give ``inf * 0 = nan`` in the
reference's dE formula, latgas.py:48-50)."""


def hardcore_latticegas(mu, Eads=0.0):
    """Nearest-neighbour-exclusion ("hard-core", hard-square) lattice gas: BASELINE config 5 read as a lattice
    model (note on C5).  It is the reference's own ``latticegas`` (latgas.py:7-57) with a
    pair energy so large that an insertion next to a particle is accepted only if the uniform draw is exactly
    0.0 -- the reference's ``random() <= exp(-dE/kT)`` has the same 2**-32 corner.  On the exclusion manifold
    E = (Eads - mu) * N, so temperature replica exchange (``TemperatureRX_MPI``) over a kT ladder is replica
    exchange over the activity z = exp(-(Eads - mu)/kT).  Start from the empty lattice (``config[:] = 0``) to stay
    on the manifold from the first sweep; a random start leaves it within a few sweeps, because removing a particle
    that has a neighbour is always accepted.

    Do not simulate at exactly Eads == mu (z = 1): then every acceptance probability of the reference's Metropolis
    rule is 0 or 1, and a sweep in fixed site order becomes a deterministic map (from the empty lattice it
    alternates between the two fully occupied sub-lattices).  The reference avoids this only through its random
    site selection; any z != 1 is fine (tests/test_gpu_hardcore_lattice.py checks densities against the exact
    transfer-matrix values)."""
    return latticegas(Eads=Eads, J=HARD_J, mu=mu)


class latgas_config(LatticeConfig):
    """Occupancy holder; ``config`` is float64 0/1 like the reference (latgas.py:61-75)."""

    _pmc_values = (0, 1)

    def prepare_random(self):
        c = self.config
        for i in range(self.lenX):
            for j in range(self.lenY):
                c[i, j] = 1 if rand.random() >= 0.5 else 0

    def __str__(self):
        c = self.peek()
        return "".join("".join("-" if v < 0 else "+" for v in row) + "\n" for row in c)


class observer(observer_base):
    def __init__(self):
        super(observer, self).__init__()
        self.energy_obs = []
        self.density_obs = []

    def logfunc(self, calc_state):
        energy = calc_state.energy
        num = calc_state.model.density(calc_state.config)
        self.density_obs.append(num)
        return energy, num
