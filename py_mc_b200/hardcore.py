"""Drop-in for ``examples/2D_hardcore/model_setup.py``: continuum hard discs in a periodic box.

* ``HC_2D``         (model_setup.py:16-67)   energy (0 or inf), trialstep -> ([id, new_coor], dE), newconfig
* ``HC_2D_config``  (model_setup.py:70-104)  Lcell, Ndisc, rdisc, coors, g_pref, prepare_ordered
* ``g_r``           (model_setup.py:106-125) radial distribution function on a ``grid_1D``
* ``observables``   (model_setup.py:129-130)

``plot_fig`` (model_setup.py:134-144, an interactive matplotlib window) is not provided: there is no
graphics stack on this box; read ``config.coors`` / ``g_r`` instead.

The reference moves a disc by ``maxstep * uniform(-1, 1)`` applied to BOTH coordinates (a diagonal
move, model_setup.py:48).  ``HC_2D(maxstep)`` reproduces that; ``HC_2D(maxstep, isotropic=True)``
draws the two displacements independently.  With ``CanonicalMonteCarlo`` one ``MCstep`` is one
single-disc trial exactly as in the reference; ``run(nsteps)`` executes the trials between two
sampling points in one kernel launch.  ``DiscBatch`` runs many independent replicas per launch.
"""
import numpy as np

from . import _cabi
from .mc import model
from .seeding import get_seed


class HC_2D(model):
    model_name = "HC_2D"
    _pmc_kind = "discs"

    def __init__(self, maxstep, isotropic=False):
        self.maxstep = maxstep
        self.isotropic = isotropic

    def _batch(self, cfg):
        return cfg._attach(self)

    def energy(self, HC_2D_config):
        """0 if no two discs overlap, else inf (model_setup.py:24-37)."""
        overlap, _ = self._batch(HC_2D_config).observe()
        return float("Infinity") if overlap[0] else 0

    def trialstep(self, HC_2D_config, energy_now):
        """The trial the engine would make next: ``dconfig = [id, new_coor]``, dE in {0, inf}."""
        i, new, overlap = self._batch(HC_2D_config).propose(0)
        return [i, new], (float("Infinity") if overlap else 0)

    def newconfig(self, HC_2D_config, dconfig):
        HC_2D_config.coors[dconfig[0]] = dconfig[1]
        return HC_2D_config


class HC_2D_config:
    """Disc coordinates with the reference's attributes; the live copy is on the device while an engine
    runs and comes back when ``coors`` is read."""

    def __init__(self, Lcell, Ndisc, rdisc, device=0):
        self.Lcell = Lcell
        self.Ndisc = Ndisc
        self.rdisc = rdisc
        self.device = device
        self._coors = np.zeros((Ndisc, 2))
        self.g_pref = Lcell ** 2.0 / (2.0 * np.pi * Ndisc * (Ndisc - 1))
        self._batch = None
        self._batch_key = None
        self._host_valid, self._dev_valid = True, False

    @property
    def coors(self):
        if not self._host_valid:
            self._coors[...] = self._batch.download(0)
            self._host_valid = True
        self._dev_valid = False  # the caller may write into the array
        return self._coors

    @coors.setter
    def coors(self, value):
        self._coors = np.asarray(value, dtype=np.float64)
        self._host_valid, self._dev_valid = True, False

    def _attach(self, mdl):
        key = (float(mdl.maxstep), bool(mdl.isotropic))
        if self._batch is None or self._batch_key != key:
            if self._batch is not None:
                self.coors  # pull
                self._batch.close()
            self._batch = _cabi.Discs(1, self.Ndisc, self.Lcell, self.rdisc, mdl.maxstep, seed=get_seed(),
                                      isotropic=mdl.isotropic, device=self.device)
            self._batch_key = key
            self._dev_valid = False
        if not self._dev_valid:
            self._batch.upload(0, self._coors)
            self._dev_valid = True
        return self._batch

    def _device_changed(self):
        self._host_valid, self._dev_valid = False, True

    def prepare_ordered(self):
        """Triangular lattice with a little space, filled along Y first (model_setup.py:80-104)."""
        sx = 1.05 * self.rdisc * 2.0
        sy = sx * 3.0 ** 0.5 * 0.5
        maxX, maxY = int(self.Lcell / sx), int(self.Lcell / sy)
        maxX2, remainder = divmod(self.Ndisc, maxY)
        assert maxX2 < maxX
        c = self.coors
        k = 0
        for i in range(maxX2):
            for j in range(maxY):
                c[k, 0] = (i + 0.5 * j % 1)
                c[k, 1] = j
                k += 1
        for i in range(remainder):
            c[k, 0] = maxX2 + 0.5 * i % 1
            c[k, 1] = i
            k += 1
        assert k == self.Ndisc
        c[:, 0] *= sx
        c[:, 1] *= sy

    def __getstate__(self):
        self.coors
        st = dict(self.__dict__)
        st.update(_batch=None, _batch_key=None, _host_valid=True, _dev_valid=False)
        return st


def g_r(HC_2D_config, grid_1D):
    """Pair histogram times ``g_pref/(X*dr)`` (model_setup.py:106-125); the counting runs on the device."""
    X, dr = grid_1D.x, grid_1D.dx
    cfg = HC_2D_config
    if cfg._batch is None or not cfg._dev_valid:
        batch = _cabi.Discs(1, cfg.Ndisc, cfg.Lcell, cfg.rdisc, 0.0, device=cfg.device)
        batch.upload(0, cfg._coors if cfg._host_valid else cfg.coors)
        _, counts = batch.observe(dr, len(X))
        batch.close()
    else:
        _, counts = cfg._batch.observe(dr, len(X))
    g = counts[0].astype(np.float64)
    g *= cfg.g_pref / (X * dr)
    return g


def observables(MCcalc, outputfi):
    return g_r(MCcalc.config, MCcalc.grid)


class DiscBatch:
    """Many independent hard-disc replicas per launch (the batched form of the model; the 64-"temperature"
    configuration of BASELINE.json is athermal -- E is 0 for every valid state, so every exchange would be
    accepted -- hence plain independent replicas)."""

    def __init__(self, nreplicas, Lcell, Ndisc, rdisc, maxstep, isotropic=False, seed=None, device=0, replica_offset=0):
        self.nreplicas, self.Lcell, self.Ndisc, self.rdisc = nreplicas, Lcell, Ndisc, rdisc
        self._d = _cabi.Discs(nreplicas, Ndisc, Lcell, rdisc, maxstep, seed=get_seed() if seed is None else seed,
                              isotropic=isotropic, device=device, replica_offset=replica_offset)
        start = HC_2D_config(Lcell, Ndisc, rdisc)
        start.prepare_ordered()
        for r in range(nreplicas):
            self._d.upload(r, start._coors)

    def run(self, ntrials):
        self._d.run(ntrials)

    def coors(self, replica):
        return self._d.download(replica)

    def g_r(self, grid_1D):
        """[nreplicas, len(grid.x)] radial distribution functions."""
        X, dr = grid_1D.x, grid_1D.dx
        _, counts = self._d.observe(dr, len(X))
        pref = self.Lcell ** 2.0 / (2.0 * np.pi * self.Ndisc * (self.Ndisc - 1))
        return counts.astype(np.float64) * (pref / (X * dr))

    def any_overlap(self):
        ov, _ = self._d.observe()
        return ov

    def acceptance(self):
        t, acc = self._d.counters()
        return acc.astype(np.float64) / max(t, 1)
