"""Device-side observers.

``observer_base`` subclasses run on the host: every sample costs a stream synchronisation and one
Python call per replica, which is what ``TemperatureRX_MPI.run`` of the reference does per rank
(py_mc/mc_mpi.py:170-177).  ``device_observer`` keeps the same place in the API -- pass it as
``observer=`` -- but the accumulation ``obs[rank_to_T[rank]] += (E, |M|)`` happens in a kernel
(``stats_accumulate_kernel``), per temperature, together with E^2, M^2 and M^4, with no host round
trip per sample.  After ``run`` the observer holds

    .counts[T]           samples per temperature index
    .mean_E, .mean_absM  <E>, <|M|>            (what run() also returns, per temperature)
    .specific_heat       (<E^2> - <E>^2) / kT^2
    .binder              1 - <M^4> / (3 <M^2>^2)

all as totals for the lattice (divide by the number of sites for per-site values).  Time series
files (``obs_save.npy`` ...) are not written on this path.

``device_observer(series=True)`` also keeps the time series of (E, |M|) *by temperature index* on the
device (``series_record_kernel``) -- the regrouping ``examples/analyze_ising2D_mpi.py:28-34`` does
afterwards from the per-replica files -- and ``analyze()`` runs the reference's post-processing on it
there: mean after discarding the first 10 % and the blocking errors of ``binning`` (py_mc/mc.py:47-61)
for every temperature and observable in one kernel (``series_binning_kernel``).
"""
import numpy as np

from .mc import observer_base


def thermo(sums, counts, kTs):
    """Means and fluctuation observables from the per-temperature sums of pymc_stats_get."""
    n = np.maximum(np.asarray(counts, dtype=np.float64), 1.0)[:, None]
    m = np.asarray(sums, dtype=np.float64) / n
    kTs = np.asarray(kTs, dtype=np.float64)
    E, E2, absM, M2, M4, M = (m[:, k] for k in range(6))
    with np.errstate(divide="ignore", invalid="ignore"):
        binder = 1.0 - M4 / (3.0 * M2 * M2)
    return dict(mean_E=E, mean_E2=E2, mean_absM=absM, mean_M2=M2, mean_M4=M4, mean_M=M,
                specific_heat=(E2 - E * E) / (kTs * kTs), binder=binder)


class device_observer(observer_base):
    _pmc_device = True

    def __init__(self, series=False):
        super(device_observer, self).__init__()
        self.counts = None
        self.keep_series = bool(series)
        self._lat = None

    def analyze(self, discard_fraction=0.1, bin_level=None):
        """(means [T, 2], errors [T, 2], all blocking errors [T, 2, level]) of the recorded series, as
        examples/analyze_ising2D_mpi.py computes them: mean over the samples after the first 10 %, error = the
        largest blocking error over levels 0 .. int(log2(nsamples // 30))."""
        if self._lat is None or not self.keep_series:
            raise RuntimeError("no series was recorded: use device_observer(series=True) and run first")
        n = self._lat.series_count()
        if bin_level is None:
            bin_level = int(np.log2(n // 30))
        mean, err = self._lat.series_binning(int(n * discard_fraction), bin_level)
        return mean, err.max(axis=2), err

    def series(self):
        """[sample, T, (E, |M|)]"""
        return self._lat.series_get()

    # host fallback with the same two logged quantities (used for the probe call and obs_info)
    def logfunc(self, calc_state):
        second = calc_state.model._pmc_second(calc_state.config)
        return calc_state.energy, abs(second)

    def _finish(self, sums, counts, kTs):
        self.sums, self.counts = sums, counts
        for k, v in thermo(sums, counts, kTs).items():
            setattr(self, k, v)
