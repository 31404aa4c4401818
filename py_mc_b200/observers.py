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

    def __init__(self):
        super(device_observer, self).__init__()
        self.counts = None

    # host fallback with the same two logged quantities (used for the probe call and obs_info)
    def logfunc(self, calc_state):
        second = calc_state.model._pmc_second(calc_state.config)
        return calc_state.energy, abs(second)

    def _finish(self, sums, counts, kTs):
        self.sums, self.counts = sums, counts
        for k, v in thermo(sums, counts, kTs).items():
            setattr(self, k, v)
