"""Drop-in for ``examples/latgas_move.py``: the lattice gas with add/delete *and* particle moves.

Same classes as ``py_mc_b200.latgas`` (``latticegas``, ``latgas_config``, ``observer``); the model
additionally proposes particle moves:

* ``trialstep`` (latgas_move.py:31-85) keeps the reference's flow -- with probability 1/2 (or always
  when the lattice is full or empty) a flip as in ``latgas.py``, otherwise one occupied and one
  empty site drawn uniformly from the whole lattice, ``dconfig = [xocc, yocc, xemp, yemp]``.
  ``move_dE='compat'`` (default, for parity with the reference) reproduces the shipped energy
  difference *including its defects* (SURVEY.md 2.3 item 3): the removal term is computed and then
  overwritten, and the wrapped ``y1`` of the insertion term is refreshed only when ``xemp`` is in the
  last row.  ``move_dE='physical'`` returns the true difference of ``energy()``.
* ``CanonicalMonteCarlo`` does not call ``trialstep``: one ``MCstep`` is a checkerboard sweep of flip
  moves followed by half a sweep of nearest-neighbour particle-hole exchanges (horizontal bonds on
  even steps, vertical on odd ones) with the physical dE -- ``lenX*lenY`` attempts of each kind,
  the reference's 1:1 mix.  Exchanges need ``lenX % 4 == 0`` and ``lenY % 4 == 0``.
"""
import random as rand

import numpy as np

from . import accept
from .latgas import latgas_config, observer  # noqa: F401  (re-exported, as in the reference module)
from .latgas import latticegas as _flip_latticegas


class latticegas(_flip_latticegas):
    _pmc_pair_moves = True

    def __init__(self, Eads, J=0, mu=0, move_dE="compat"):
        super(latticegas, self).__init__(Eads, J, mu)
        if move_dE not in ("compat", "physical"):
            raise ValueError("move_dE must be 'compat' or 'physical'")
        self.move_dE = move_dE

    def _pmc_table_key(self):
        return ("latticegas_move", self.J, self.Eads, self.mu)

    def _pmc_pair_thresholds(self, kT):
        """thr[d+3], d = (occupied neighbours gained) - (lost) = -3..3, dE = J*d."""
        return accept.table([self.J * float(d) for d in range(-3, 4)], kT)

    def trialstep(self, latgas_config, energy):
        lenX, lenY = latgas_config.lenX, latgas_config.lenY
        nocc = self.density(latgas_config)
        if rand.random() < 0.5 or nocc == lenX * lenY or nocc == 0:
            return super(latticegas, self).trialstep(latgas_config, energy)
        c = latgas_config.peek()
        xocc, yocc = rand.choice(np.transpose(np.where(c == 1)))
        xemp, yemp = rand.choice(np.transpose(np.where(c == 0)))
        lat = latgas_config._attach(self._pmc_model)

        def occ(x, y):
            v, _ = lat.propose(latgas_config._rep, [int(x) % lenX], [int(y) % lenY])
            return float(v[0])

        if self.move_dE == "physical":
            _, nn_occ = lat.propose(latgas_config._rep, [int(xocc)], [int(yocc)])
            _, nn_emp = lat.propose(latgas_config._rep, [int(xemp)], [int(yemp)])
            dx, dy = abs(int(xocc) - int(xemp)), abs(int(yocc) - int(yemp))
            adjacent = (min(dx, lenX - dx) + min(dy, lenY - dy)) == 1
            dE = self.J * float(int(nn_emp[0]) - int(nn_occ[0]) - (1 if adjacent else 0))
        else:
            # insertion term of the reference; its y1 is the wrapped yocc+1 left over from the removal
            # block unless xemp is in the last row (latgas_move.py:63-82)
            y1 = 0 if yocc == lenY - 1 else yocc + 1
            x1 = xemp + 1
            if xemp == lenX - 1:
                x1 = 0
                y1 = yemp + 1
            if yemp == lenY - 1:
                y1 = 0
            dE = (self.J * (occ(xemp - 1, yemp) + occ(x1, yemp) + occ(xemp, yemp - 1) + occ(xemp, y1))
                  + self.Eads - self.mu)
        return [xocc, yocc, xemp, yemp], dE
