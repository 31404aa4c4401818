"""Per-temperature regrouping and error analysis of a finished replica-exchange run: the job of
``examples/analyze_ising2D_mpi.py`` (lines 1-45), without MPI.

The reference script runs one MPI rank per replica; rank r loads ``r/obs_save.npy`` and
``r/Trank_hist.npy``, masks its time series to the samples taken at temperature index i
(``np.where(Trank_hist == i, obs[:, j], 0.0)``), and an ``MPI.Reduce(SUM)`` over ranks assembles the
series of temperature i.  Rank 0 writes the mean after discarding the first 10 % (``obs_mean.dat``)
and the blocking errors of the full series (``binning.dat``).  ``analyze`` does the same in one
process and writes byte-identical files; ``regroup`` returns the assembled series.
"""
import os

import numpy as np

from .mc import binning


def regroup(obs, Trank_hist):
    """obs [replica, sample, nobs], Trank_hist [replica, sample] -> series [T, sample, nobs]:
    the observable of whichever replica sat at temperature index T at each sample."""
    obs = np.asarray(obs, dtype=np.float64)
    th = np.asarray(Trank_hist)
    R, ns, nobs = obs.shape
    out = np.zeros((R, ns, nobs))
    for i in range(R):
        for j in range(nobs):
            total = np.zeros(ns)
            for r in range(R):  # rank order, like the Reduce
                total = total + np.where(th[r] == i, obs[r, :, j], 0.0)
            out[i, :, j] = total
    return out


def analyze(nreplicas, directory=".", write=True):
    """Returns (means [T, nobs], errors [T, nobs]); writes obs_mean.dat and binning.dat like the script."""
    obs = np.array([np.load(os.path.join(directory, str(r), "obs_save.npy")) for r in range(nreplicas)])
    th = np.array([np.load(os.path.join(directory, str(r), "Trank_hist.npy")) for r in range(nreplicas)])
    nsamples, nobs = obs.shape[1], obs.shape[2]
    throwout = int(nsamples * 0.1)
    bin_level = int(np.log2(nsamples // 30))
    series = regroup(obs, th)
    means = np.zeros((nreplicas, nobs))
    errors = np.zeros((nreplicas, nobs))
    mean_txt, bin_txt = [], []
    for i in range(nreplicas):
        for j in range(nobs):
            s = series[i, :, j]
            means[i, j] = s[throwout:].mean()
            bin_data = binning(s, bin_level)
            errors[i, j] = np.max(bin_data)
            mean_txt.append(str(means[i, j]) + "\t" + str(errors[i, j]) + "\t")
            bin_txt.append("# " + str(i) + "," + str(j) + "\n" + "\n".join(str(x) for x in bin_data) + "\n\n\n")
        mean_txt.append("\n")
    if write:
        with open(os.path.join(directory, "obs_mean.dat"), "w") as f:
            f.write("".join(mean_txt))
        with open(os.path.join(directory, "binning.dat"), "w") as f:
            f.write("".join(bin_txt))
    return means, errors
