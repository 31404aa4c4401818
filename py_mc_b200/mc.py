"""Drop-in for ``py_mc/mc.py``: same names, signatures and return values; the Metropolis loop runs
on the GPU.

Reference items mirrored here (``py_mc/mc.py`` line ranges):

* ``model``                13-40    plug-in protocol: energy / trialstep -> (dconfig, dE) / newconfig
* ``grid_1D``              42-45
* ``binning``              47-61    blocking error analysis
* ``obs_encode`` / ``args_info`` / ``obs_decode``  65-104
* ``make_observefunc``     106-118  (legacy helper)
* ``observer_base``        121-160
* ``CanonicalMonteCarlo``  162-214  constructor, ``MCstep``, ``run``

Step accounting.  The reference's ``MCstep`` is ONE single-site trial at a random site; here
``MCstep`` is one full checkerboard sweep = ``lenX*lenY`` trials (``attempts_per_step``), because a
GPU launch that touches one site is pointless.  ``nsteps``, ``sample_frequency`` and
``print_frequency`` therefore count sweeps.  Means of observables are unaffected; autocorrelation
times (and so binning levels) are measured in sweeps.

Known reference quirks kept on purpose: ``run`` probes the observer once before the loop with
``lprint=True`` (mc.py:193), so a user observer that appends in ``logfunc`` records one extra
pre-run sample; the observer may be passed as the third positional argument because the reference's
own ``ParallelMC.run`` does so (py_mc/mc_mpi.py:66).  The leftover debug ``print`` of mc.py:157 is
not reproduced.
"""
import os
import sys

import numpy as np

verylargeint = sys.maxsize


class model:
    """Base template (mc.py:13-40): zero energy, configuration never changes."""

    model_name = None

    def energy(self, config):
        return 0.0

    def trialstep(self, config, energy):
        """Return (dconfig, dE) for a trial move; must leave ``config`` unchanged."""
        return None, 0.0

    def newconfig(self, config, dconfig):
        return config


class grid_1D:
    def __init__(self, dx, minx, maxx):
        self.dx = dx
        self.x = np.arange(minx, maxx, dx)


def binning(x, nlevels):
    """Standard-error estimates at blocking levels 0..nlevels (mc.py:47-61).

    Leading samples are dropped until the length is a multiple of 2**nlevels; each level averages
    neighbouring pairs of the previous one.
    """
    series = np.array(x, dtype=np.float64)
    drop = len(series) % (2 ** nlevels)
    if drop:
        series = series[drop:]
    errors = [np.sqrt(np.var(series, ddof=1) / len(series))]
    for _ in range(nlevels):
        series = 0.5 * (series[0::2] + series[1::2])
        errors.append(np.sqrt(np.var(series, ddof=1) / len(series)))
    return errors


def _flat(arg):
    return np.array([arg]).ravel()


def obs_encode(*args):
    """Flatten scalars/arrays into one 1-D float array (mc.py:65-77)."""
    out = np.array([])
    for arg in args:
        out = np.concatenate((out, _flat(arg)))
    return out


def args_info(*args):
    """[nargs, len(arg0), len(arg1), ...] -- the key ``obs_decode`` needs (mc.py:79-89)."""
    return np.concatenate((np.array([len(args)]), np.array([len(_flat(a)) for a in args], dtype=int)))


def obs_decode(args_info, obs_array):
    """Inverse of ``obs_encode``: scalars stay scalars, longer pieces come back as slices (mc.py:92-104)."""
    nargs = args_info[0]
    lengths = args_info[1:nargs + 1]
    out, pos = [], 0
    for n in lengths:
        out.append(obs_array[pos] if n == 1 else obs_array[pos:pos + n])
        pos += n
    return out


def make_observefunc(logfunc, *multiDfuncs):
    """Legacy observer factory (mc.py:106-118)."""

    def observefunc(calc_state, outputfi):
        log = logfunc(calc_state)
        outputfi.write(str(calc_state.kT) + "\t")
        if hasattr(log, "__getitem__"):
            outputfi.write("\t".join(str(v) for v in log) + "\n")
        else:
            outputfi.write(str(log) + "\n")
        extra = [f(calc_state) for f in multiDfuncs]
        return obs_encode(*log, *extra)

    return observefunc


class observer_base:
    """Observer protocol (mc.py:121-160).  Subclasses override ``logfunc`` (scalars written to
    obs.dat), ``savefuncs`` (arrays averaged only) and ``writefile``."""

    def __init__(self):
        self.lprintcount = 0

    def obs_info(self, calc_state):
        log = self.logfunc(calc_state)
        if not isinstance(log, tuple):
            log = (log,)
        saved = self.savefuncs(calc_state)
        extra = []
        if saved is not None:
            extra = list(saved) if isinstance(saved, tuple) else [saved]
        return args_info(*log, *extra)

    def logfunc(self, calc_state):
        return calc_state.energy,

    def savefuncs(self, calc_state):
        return None

    def writefile(self, calc_state):
        return None

    def observe(self, calc_state, outputfi, lprint=True):
        log = np.atleast_1d(self.logfunc(calc_state))
        if lprint:
            outputfi.write(str(self.lprintcount) + "\t" + str(calc_state.kT) + "\t" +
                           "\t".join(str(v) for v in log) + "\n")
            outputfi.flush()
            self.writefile(calc_state)
            self.lprintcount += 1
        saved = self.savefuncs(calc_state)
        if saved is None:
            return log
        return np.concatenate((log, np.atleast_1d(saved).ravel()))


def _is_observer(obj):
    return hasattr(obj, "observe") and hasattr(obj, "logfunc")


class CanonicalMonteCarlo:
    """GPU Metropolis engine with the reference's constructor, attributes and ``run`` signature
    (mc.py:162-214).  ``model`` must be one of the device-backed models of this package
    (``py_mc_b200.ising2D.ising2D``, ``py_mc_b200.latgas.latticegas``): there is no CPU fallback."""

    def __init__(self, model, kT, config, grid=0):
        self.model = model
        self.config = config
        self.kT = kT
        self.grid = grid
        self.obs_save = []
        self._lat = None
        self._table_kT = None
        self._energy = None

    # -- device plumbing ---------------------------------------------------------------
    def _is_discs(self):
        return getattr(self.model, "_pmc_kind", None) == "discs" and hasattr(self.config, "_attach")

    def _device(self):
        if self._is_discs():
            return self.config._attach(self.model)
        code = getattr(self.model, "_pmc_model", None)
        if code is None or not hasattr(self.config, "_attach"):
            raise TypeError(
                "py_mc_b200.CanonicalMonteCarlo runs device-backed models only "
                "(py_mc_b200.ising2D.ising2D, py_mc_b200.latgas.latticegas with their config classes); "
                "got model=%r config=%r. There is no CPU fallback." % (type(self.model).__name__,
                                                                        type(self.config).__name__))
        lat = self.config._attach(code)
        key = (self.kT, self.model._pmc_table_key())
        if lat is not self._lat or self._table_kT != key:
            lat.set_accept_tables(self.model._pmc_thresholds(self.kT)[None, :])
            if getattr(self.model, "_pmc_pair_moves", False):
                lat.set_pair_tables(self.model._pmc_pair_thresholds(self.kT)[None, :])
            self._lat = lat
            self._table_kT = key
        return lat

    @property
    def attempts_per_step(self):
        """Metropolis trials performed by one ``MCstep``: one sweep of single-site moves, plus as
        many particle-hole exchange attempts for models with particle moves (latgas_move); one
        single-disc trial for the hard-disc model."""
        if self._is_discs():
            return 1
        n = self.config.lenX * self.config.lenY
        return 2 * n if getattr(self.model, "_pmc_pair_moves", False) else n

    @property
    def energy(self):
        if self._energy is None and self._is_discs():
            self._energy = self.model.energy(self.config)
        if self._energy is None:
            obs = self.config._device_observables()
            if obs is None:  # the host copy is the newest (someone took a writable view): recompute
                self._energy = self.model.energy(self.config)
            elif hasattr(self.model, "_pmc_energy_for"):
                self._energy = self.model._pmc_energy_for(self.config, *obs)
            else:
                self._energy = self.model._pmc_energy(*obs)
        return self._energy

    @energy.setter
    def energy(self, value):
        self._energy = value

    # -- the reference API -------------------------------------------------------------
    def MCstep(self):
        """One full checkerboard sweep (see module docstring for the step accounting)."""
        self._sweeps(1)

    def _sweeps(self, n):
        lat = self._device()
        if self._is_discs():  # n single-disc trials in one launch; a valid state keeps E = 0
            lat.run(n)
            self.config._device_changed()
            return
        if getattr(self.model, "_pmc_pair_moves", False):
            # per step: flip sweep, then half of the bond classes (horizontal / vertical in turn)
            lat.mixed_sweep(n, 0x00FF, 0xFF00)
        else:
            lat.sweep(n)
        self.config._device_changed()
        self._energy = None

    def _run_device_stats(self, lat, nsteps, sample_frequency, observer):
        """Sampling on the device (py_mc_b200.observers.device_observer): no host round trip per sample."""
        lat.stats_reset(*self.model._pmc_energy_coeffs())
        done = 0
        while done < nsteps:
            chunk = min(nsteps - done, sample_frequency - done % sample_frequency)
            self._sweeps(chunk)
            done += chunk
            if done % sample_frequency == 0:
                lat.stats_accumulate()
        sums, counts = lat.stats_get()
        observer._finish(sums, counts, [self.kT])
        self.energy
        if counts[0]:
            return [observer.mean_E[0], observer.mean_absM[0]]

    def run(self, nsteps, sample_frequency=verylargeint, print_frequency=verylargeint,
            observer=None, save_obs=False):
        if _is_observer(print_frequency):  # run(nsteps, sample_frequency, observer): mc_mpi.py:66
            observer, print_frequency = print_frequency, verylargeint
        if observer is None:
            observer = _default_observer
        observables = 0.0
        nsample = 0
        lat = self._device()
        if self._is_discs():
            self._energy = self.model.energy(self.config)
        else:
            lat.reduce_observables()  # model.energy(config) at the top of run (mc.py:191)
            self._energy = None
        if getattr(observer, "_pmc_device", False) and not self._is_discs() and sample_frequency <= nsteps:
            return self._run_device_stats(lat, nsteps, sample_frequency, observer)
        output = open("obs.dat", "a")
        try:
            with open(os.devnull, "w") as devnull:
                probe = observer.observe(self, devnull)
            observe = hasattr(probe, "__add__")
            if not observe or sample_frequency > nsteps:
                self._sweeps(nsteps)
            else:
                done = 0
                while done < nsteps:
                    chunk = min(sample_frequency - (done % sample_frequency), nsteps - done)
                    self._sweeps(chunk)
                    done += chunk
                    if done % sample_frequency == 0:
                        obs_step = observer.observe(self, output, done % print_frequency == 0)
                        observables += obs_step
                        nsample += 1
                        if save_obs:
                            self.obs_save.append(obs_step)
        finally:
            output.close()
        self.energy  # cache the tracked energy while the device copy is the newest
        if save_obs:
            with open("obs_save.npy", "wb") as f:
                np.save(f, np.array(self.obs_save))
        if nsample > 0:
            observables /= nsample
            return obs_decode(observer.obs_info(self), observables)
        return None


def swap_configs(MCreplicas, rep, accept_count):
    """Exchange configuration and energy of replicas ``rep`` and ``rep + 1`` (mc.py:220-229; the helper of
    the reference's disabled single-process ``TemperatureReplicaExchange``).  Host-side: it swaps the
    objects, whatever engine they belong to."""
    a, b = MCreplicas[rep], MCreplicas[rep + 1]
    a.config, b.config = b.config, a.config
    a.energy, b.energy = b.energy, a.energy
    return MCreplicas, accept_count + 1


_default_observer = observer_base()  # one shared instance, as the reference's default argument (mc.py:188)
