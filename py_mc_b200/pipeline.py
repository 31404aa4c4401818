"""Independent lattices streamed through one GPU: configuration in, sweeps, configuration and exact sums out.

The reference runs one `CanonicalMonteCarlo` per process on arrays that never leave host memory
(py_mc/mc.py:163-214).  When the configurations live on the host and every job is "these sweeps on this
lattice" -- a parameter scan, an ensemble of independent starts -- the PCIe copies of a 256 MiB image cost
about as much as the sweeps themselves.  `SweepPipeline` keeps `depth` handles, each with its own CUDA stream
(include/pymc_b200.h: pymc_upload_config_async / pymc_download_config_async / pymc_get_tracked_async): while one
lattice is being swept, the next one is on its way up and the previous one on its way down, so a job costs
max(copy, sweeps) instead of their sum.

A job's result is a pure function of (input configuration, seed, first_sweep, nsweeps): all handles share the
seed and the replica index, so it does not depend on which handle ran it and equals what the synchronous calls
(`Lattices.upload`, `sweep`, `download`, `tracked`) give.
"""
import numpy as np

from . import _cabi


class SweepPipeline:
    def __init__(self, model_code, lenX, lenY, thresholds, nsweeps, depth=3, seed=0, layout=_cabi.LAYOUT_INT8, device=0,
                 kernel=_cabi.KERNEL_AUTO):
        """thresholds: one row of the acceptance table (`model._pmc_thresholds(kT)`); nsweeps: sweeps per job."""
        if depth < 1:
            raise ValueError("depth must be at least 1")
        self.shape = (int(lenX), int(lenY))
        self.nsweeps = int(nsweeps)
        self._slots = []
        thr = np.ascontiguousarray(thresholds, dtype=np.uint32)[None, :]
        for _ in range(depth):
            lat = _cabi.Lattices(model_code, lenX, lenY, nreplicas=1, ntemps=1, seed=seed, layout=layout, device=device,
                                 kernel=kernel)
            lat.set_accept_tables(thr)
            self._slots.append({"lat": lat, "sums": _cabi.pinned_empty((1, 2), np.int64), "job": None})
        self._next = 0

    def close(self):
        for s in self._slots:
            s["lat"].close()
        self._slots = []

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _finish(self, slot):
        job = slot["job"]
        if job is None:
            return None
        slot["lat"].synchronize()  # copies done; raises if a barrier inside the sweep kernel gave up
        slot["job"] = None
        tag, host_out = job
        return tag, host_out, int(slot["sums"][0, 0]), int(slot["sums"][0, 1])

    def submit(self, host_in, host_out, first_sweep=0, tag=None):
        """Enqueue one job: upload `host_in`, `nsweeps` sweeps starting at stream position `first_sweep`, download
        into `host_out` (both page-locked int8 [lenX][lenY] from `_cabi.pinned_empty`; they may be the same array).
        Returns the finished job whose handle is reused -- (tag, host_out, bond_sum, second) -- or None."""
        slot = self._slots[self._next]
        self._next = (self._next + 1) % len(self._slots)
        done = self._finish(slot)
        lat = slot["lat"]
        lat.upload_async(0, host_in)
        lat.sweep_counter = int(first_sweep)
        lat.sweep(self.nsweeps)
        lat.download_async(0, host_out)
        lat.tracked_async(slot["sums"])
        slot["job"] = (tag, host_out)
        return done

    def drain(self):
        """Wait for every job in flight; returns their results in submission order."""
        out = []
        n = len(self._slots)
        for k in range(n):
            r = self._finish(self._slots[(self._next + k) % n])
            if r is not None:
                out.append(r)
        return out
