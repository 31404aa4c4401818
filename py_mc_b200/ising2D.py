"""Drop-in for ``examples/ising2D.py``: the 2D Ising model, its config holder and observer.

* ``ising2D``        (ising2D.py:7-52)   energy, magnetization, trialstep -> ([x, y], dE), newconfig
* ``ising2D_config`` (ising2D.py:56-81)  lenX, lenY, config, prepare_random, __str__
* ``observer``       (ising2D.py:83-95)  logs (energy, |M|) and keeps both time series

``energy`` / ``magnetization`` / ``trialstep`` answer from the device (exact integer reductions and
the neighbour-sum probe kernel) with the reference's float64 expressions applied on the host, so
their results are bit-identical to the reference on the same configuration.  A ``CanonicalMonteCarlo``
from ``py_mc_b200.mc`` uses the batched sweep kernels instead of calling ``trialstep`` per site.
"""
import random as rand

import numpy as np

from . import _cabi, accept
from ._device import LatticeConfig
from .mc import model, observer_base


class ising2D(model):
    """E = J * sum_ij s(i,j) (s(i-1,j) + s(i,j-1)), periodic (ising2D.py:15-22)."""

    model_name = "ising2D"
    _pmc_model = _cabi.MODEL_ISING

    def __init__(self, J):
        self.J = J

    # -- device protocol -----------------------------------------------------------------
    def _pmc_table_key(self):
        return ("ising2D", self.J)

    def _pmc_dE(self, s, nnsum):
        """dE of flipping spin s with neighbour sum nnsum -- the expression of ising2D.py:45."""
        return -2.0 * self.J * float(s) * float(nnsum)

    def _pmc_thresholds(self, kT):
        """Acceptance thresholds indexed by the number of aligned neighbours a = 0..4
        (s * nnsum = 2a - 4).  The 10-entry (s, nnsum) table of the reference expression is
        symmetric under (s, nnsum) -> (-s, -nnsum) in IEEE arithmetic; that is asserted here."""
        up = accept.table([self._pmc_dE(1.0, 2 * a - 4) for a in range(5)], kT)
        down = accept.table([self._pmc_dE(-1.0, 4 - 2 * a) for a in range(5)], kT)
        if not np.array_equal(up, down):
            raise ArithmeticError("Ising acceptance table is not spin-flip symmetric for J=%r kT=%r" % (self.J, kT))
        return up

    def _pmc_energy(self, bond, second):
        return self.J * float(bond)

    def _pmc_second(self, config):
        return self.magnetization(config)

    def _pmc_energy_coeffs(self):
        """(c_bond, c_second) with E = c_bond*bond + c_second*second, for the on-device exchange test."""
        return float(self.J), 0.0

    # -- reference API -------------------------------------------------------------------
    def energy(self, ising2D_config):
        lat = ising2D_config._attach(self._pmc_model)
        bond, _ = lat.reduce_observables()
        return self._pmc_energy(int(bond[ising2D_config._rep]), 0)

    def magnetization(self, ising2D_config):
        obs = ising2D_config._device_observables()
        if obs is None:
            lat = ising2D_config._attach(self._pmc_model)
            _, m = lat.reduce_observables()
            return float(m[ising2D_config._rep])
        return float(obs[1])

    def trialstep(self, ising2D_config, energy):
        x = rand.randrange(ising2D_config.lenX)
        y = rand.randrange(ising2D_config.lenY)
        lat = ising2D_config._attach(self._pmc_model)
        s, nn = lat.propose(ising2D_config._rep, [x], [y])
        return [x, y], self._pmc_dE(s[0], nn[0])

    def newconfig(self, ising2D_config, dconfig):
        ising2D_config.config[dconfig[0], dconfig[1]] *= -1
        return ising2D_config


class ising2D_config(LatticeConfig):
    """Configuration holder with the reference's attributes; ``config`` is float64 +1/-1 by default
    (``dtype=np.int8`` keeps the host copy small for very large lattices).  ``layout`` selects the
    device storage: ``'int8'``, ``'bits'`` (needs lenY % 128 == 0) or ``'auto'``."""

    _pmc_values = (-1, 1)

    def prepare_random(self):
        """+1 where random() >= 0.5 else -1, row-major draw order (ising2D.py:64-70)."""
        c = self.config
        for i in range(self.lenX):
            for j in range(self.lenY):
                c[i, j] = 1 if rand.random() >= 0.5 else -1

    def __str__(self):
        c = self.peek()
        return "".join("".join("-" if v < 0 else "+" for v in row) + "\n" for row in c)


class observer(observer_base):
    def __init__(self):
        super(observer, self).__init__()
        self.energy_obs = []
        self.magnet_obs = []

    def logfunc(self, calc_state):
        energy = calc_state.energy
        absmag = abs(calc_state.model.magnetization(calc_state.config))
        self.energy_obs.append(energy)
        self.magnet_obs.append(absmag)
        return energy, absmag
