"""Drop-in for ``py_mc/mc_mpi.py``: parallel replicas and temperature replica exchange, with the
replicas batched on GPUs instead of one MPI rank per replica.

Reference items mirrored (``py_mc/mc_mpi.py`` line ranges):

* ``RX_MPI_init``         9-34    command-line parsing and process-group set-up
* ``ParallelMC``          37-72   independent replicas at fixed temperatures
* ``TemperatureRX_MPI``   75-209  constructor, ``reload``, ``Xtrial``, ``run``

How the mapping works.  The reference runs one process per replica and ``rank`` means "replica
index".  Here one process drives one GPU and owns a contiguous block of replicas; the *global
replica index* plays the role the MPI rank plays in the reference (``rank_to_T[replica]``,
per-replica directories ``str(replica)/``, ``Trank_hist`` ...).  Per exchange the reference sends
one double and one int per pair and all-gathers the rank->temperature map; here the energies of
all replicas are all-gathered once (NCCL through ``torch.distributed`` when several GPUs take part,
nothing at all on one GPU) and every GPU runs the same decision kernel on them, so the map never
has to be communicated and the host is not involved in an exchange.

Step accounting is in sweeps (see ``py_mc_b200.mc``): ``nsteps``, ``RXtrial_frequency``,
``sample_frequency`` and ``print_frequency`` count full-lattice sweeps.
"""
import os
import pickle
import sys

import numpy as np

from . import _cabi
from ._device import _layout_code
from .mc import *  # noqa: F401,F403  (the reference does the same star import, mc_mpi.py:7)
from .mc import _default_observer, _is_observer, obs_decode, verylargeint
from .seeding import get_seed


# ------------------------------------------------------------------------------------------------
# process group: torch.distributed when present, a trivial one-process group otherwise
# ------------------------------------------------------------------------------------------------
class ReplicaComm:
    """Block partition of replicas over processes plus the two collectives the path needs.

    ``Get_rank`` / ``Get_size`` keep the mpi4py spelling because user scripts call them on the
    object returned by ``RX_MPI_init`` (examples/ising2D_mpi.py:38)."""

    def __init__(self, dist=None):
        self.dist = dist if (dist is not None and dist.is_available() and dist.is_initialized()) else None
        self.rank = self.dist.get_rank() if self.dist else 0
        self.size = self.dist.get_world_size() if self.dist else 1
        self.backend = self.dist.get_backend() if self.dist else "single"

    def Get_rank(self):
        return self.rank

    def Get_size(self):
        return self.size

    def partition(self, nreplicas):
        """(first, count) of this process's block; replicas must divide evenly (equal-size all-gather)."""
        if nreplicas % self.size:
            raise ValueError("%d replicas cannot be split evenly over %d GPUs" % (nreplicas, self.size))
        per = nreplicas // self.size
        return self.rank * per, per

    def allgather_f64_host(self, local):
        """Gather equal-length float64 blocks from all processes, in rank order (host arrays)."""
        local = np.ascontiguousarray(local, dtype=np.float64)
        if not self.dist:
            return local.copy()
        import torch
        dev = "cuda" if self.backend == "nccl" else "cpu"
        src = torch.from_numpy(local).to(dev)
        out = torch.empty(local.size * self.size, dtype=torch.float64, device=dev)
        self.dist.all_gather_into_tensor(out, src)
        return out.cpu().numpy()

    def allreduce_sum_host(self, array):
        """Sum a float64 array over processes (mirror of the closing Allreduce, mc_mpi.py:204)."""
        array = np.ascontiguousarray(array, dtype=np.float64)
        if not self.dist:
            return array.copy()
        import torch
        dev = "cuda" if self.backend == "nccl" else "cpu"
        t = torch.from_numpy(array.copy()).to(dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return t.cpu().numpy()

    def barrier(self):
        if self.dist:
            self.dist.barrier()


class _DeviceArray:
    """Zero-copy handle on library-owned device memory for ``torch.as_tensor``."""

    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 3,
                                         "strides": None}


def RX_MPI_init():
    """Parse ``argv[1] = nreplicas`` and ``argv[2] = nprocs_per_replica`` like the reference
    (mc_mpi.py:10-12) and join the process group if the script was started by ``torchrun``
    (one process per GPU).  Returns ``(comm, nreplicas, nprocs_per_replica)``."""
    nreplicas = int(sys.argv[1])
    nprocs_per_replica = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    dist = None
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch
        import torch.distributed as dist
        if not dist.is_initialized():
            local = int(os.environ.get("LOCAL_RANK", "0"))
            torch.cuda.set_device(local)
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return ReplicaComm(dist), nreplicas, nprocs_per_replica


class _ReplicaState:
    """What an observer sees for one replica: the ``calc_state`` of the reference
    (``.energy .kT .model .config``, py_mc/mc.py:124-160)."""

    def __init__(self, model, config, grid):
        self.model, self.config, self.grid = model, config, grid
        self.kT = None
        self.energy = None
        self.obs_save = []


class ParallelMC(object):
    """Independent replicas at fixed temperatures (mc_mpi.py:37-72), batched on the GPU(s)."""

    def __init__(self, comm, MCalgo, model, configs, kTs, grid=None, subdirs=True):
        self.comm = comm if isinstance(comm, ReplicaComm) else ReplicaComm(_maybe_dist())
        self.rank = self.comm.Get_rank()
        self.procs = self.comm.Get_size()
        self.kTs = kTs
        self.model = model
        self.grid = grid
        self.subdirs = subdirs
        self.configs = configs
        self.nreplicas = len(configs)
        self.MCalgo = MCalgo
        if self.nreplicas != len(kTs):
            # the reference also demands one temperature per replica (mc_mpi.py:47-52)
            if self.rank == 0:
                print("ERROR: You have to set the number of replicas equal to the number of temperatures")
            sys.exit(1)
        code = getattr(model, "_pmc_model", None)
        if code is None or not all(hasattr(c, "_bind") for c in configs):
            raise TypeError("py_mc_b200 runs device-backed models/configs only; there is no CPU fallback")
        self.first, self.nlocal = self.comm.partition(self.nreplicas)
        c0 = configs[0]
        device = int(os.environ.get("LOCAL_RANK", "0")) if self.procs > 1 else getattr(c0, "device", 0)
        self._lat = _cabi.Lattices(code, c0.lenX, c0.lenY, nreplicas=self.nlocal, ntemps=self.nreplicas,
                                   seed=get_seed(), layout=_layout_code(c0.layout, c0.lenY), device=device,
                                   replica_offset=self.first, nreplicas_global=self.nreplicas)
        if self.procs > 1 and self.comm.backend == "nccl":
            import torch
            self._lat.set_stream(torch.cuda.current_stream().cuda_stream)
        self.states = []
        for loc in range(self.nlocal):
            cfg = configs[self.first + loc]
            cfg._bind(self._lat, loc, code)
            self.states.append(_ReplicaState(model, cfg, grid))
        self.mycalc = self.states[0]  # the reference's per-rank engine object (first local replica here)
        self.rank_to_T = np.arange(0, self.nreplicas, 1, dtype=int)
        self._set_tables()
        self._refresh_states()

    # -- helpers -------------------------------------------------------------------------------
    def _set_tables(self):
        thr = np.array([self.model._pmc_thresholds(kT) for kT in self.kTs])
        self._lat.set_accept_tables(thr)
        self._lat.set_replica_to_T(np.asarray(self.rank_to_T, dtype=np.int32))

    def _refresh_states(self, with_energy=False):
        if with_energy:
            b, s = self._lat.tracked()
        for loc, st in enumerate(self.states):
            st.kT = self.kTs[self.rank_to_T[self.first + loc]]
            if with_energy and hasattr(self.model, "_pmc_energy_for"):
                st.energy = self.model._pmc_energy_for(st.config, int(b[loc]), int(s[loc]))
            elif with_energy:
                st.energy = self.model._pmc_energy(int(b[loc]), int(s[loc]))

    def _configs_changed(self):
        for st in self.states:
            st.config._device_changed()

    def _sync_host_edits(self):
        """Re-upload replicas whose host array was handed out (and possibly edited) since the last run."""
        for st in self.states:
            if not st.config._dev_valid:
                st.config._attach(self.model._pmc_model)

    def _replica_dir(self, loc):
        return str(self.first + loc)

    def run(self, nsteps, sample_frequency, observer=None):
        if observer is None:
            observer = _default_observer
        self._sync_host_edits()
        self._lat.reduce_observables()
        self._refresh_states(with_energy=True)
        outs = [open(self._path(loc, "obs.dat"), "a") for loc in range(self.nlocal)]
        observables = [0.0] * self.nlocal
        nsample = 0
        with open(os.devnull, "w") as devnull:
            probe = observer.observe(self.states[0], devnull)
        observe = hasattr(probe, "__add__") and bool(sample_frequency)
        done = 0
        while done < nsteps:
            chunk = nsteps - done
            if observe:
                chunk = min(chunk, sample_frequency - done % sample_frequency)
            self._lat.sweep(chunk)
            self._configs_changed()
            done += chunk
            if observe and done % sample_frequency == 0:
                self._refresh_states(with_energy=True)
                for loc in range(self.nlocal):
                    observables[loc] = observables[loc] + observer.observe(self.states[loc], outs[loc], False)
                nsample += 1
        for f in outs:
            f.close()
        for loc in range(self.nlocal):
            with open(self._path(loc, "config.pickle"), "wb") as f:
                pickle.dump(self.states[loc].config, f)
        if sample_frequency and nsample:
            local = np.array([o / nsample for o in observables], dtype=np.float64)
            nobs = local.shape[1]
            return self.comm.allgather_f64_host(local.ravel()).reshape(self.nreplicas, nobs)

    def _path(self, loc, name):
        if not self.subdirs:
            return name
        d = self._replica_dir(loc)
        os.makedirs(d, exist_ok=True)
        return os.path.join(d, name)


def _maybe_dist():
    try:
        import torch.distributed as dist
        return dist if dist.is_available() and dist.is_initialized() else None
    except ImportError:
        return None


class TemperatureRX_MPI(ParallelMC):
    """Temperature replica exchange (mc_mpi.py:75-209) with on-device exchange decisions."""

    def __init__(self, comm, MCalgo, model, configs, kTs, grid=None, subdirs=True):
        super(TemperatureRX_MPI, self).__init__(comm, MCalgo, model, configs, kTs, grid, subdirs)
        self.betas = 1.0 / np.array(kTs)
        self.obs_save = [[] for _ in range(self.nlocal)]
        self.Trank_hist = [[] for _ in range(self.nlocal)]
        self.kT_hist = [[] for _ in range(self.nlocal)]
        c_bond, c_second = model._pmc_energy_coeffs()
        self._lat.rx_setup(self.betas, c_bond, c_second)
        self._energies_t = None
        self.accept_count = 0
        # the exchange inside the sweep kernel: always on one GPU; on several after connecting them through peer memory
        self._fused = self._lat.sweep_kernel_info()[1] == "pbatch_kernel" and self.nreplicas <= 4096 * 16
        if self._fused and self.procs > 1:
            if self.comm.backend == "nccl" and os.environ.get("PYMC_RX_COMM", "peer") == "peer":
                self._lat.rx_connect_group(self.comm.dist)
            else:
                self._fused = False

    # -- restart (mc_mpi.py:86-92) ---------------------------------------------------------------
    def reload(self):
        with open("rank_to_T.pickle", "rb") as f:
            self.rank_to_T = np.asarray(pickle.load(f), dtype=int)
        self._lat.set_replica_to_T(self.rank_to_T.astype(np.int32))
        self.obs_save0, self.Trank_hist0, self.kT_hist0 = [], [], []
        code = self.model._pmc_model
        for loc in range(self.nlocal):
            d = self._replica_dir(loc)
            with open(os.path.join(d, "calc.pickle"), "rb") as f:
                cfg = pickle.load(f)
            cfg._bind(self._lat, loc, code)
            self.states[loc].config = cfg
            self.configs[self.first + loc] = cfg
            self.obs_save0.append(np.load(os.path.join(d, "obs_save.npy")))
            self.Trank_hist0.append(np.load(os.path.join(d, "Trank_hist.npy")))
            self.kT_hist0.append(np.load(os.path.join(d, "kT_hist.npy")))
        state = "rx_state.pickle"
        if os.path.exists(state):  # extension: random-stream position for bit-reproducible restarts
            with open(state, "rb") as f:
                st = pickle.load(f)
            self._lat.sweep_counter = st["sweep_counter"]
            self._lat.rx_round = st["rx_round"]
        self._refresh_states()

    def find_procrank_from_Trank(self, Trank):
        """Global replica index currently at temperature index ``Trank`` (mc_mpi.py:95-100)."""
        i = np.argwhere(np.asarray(self.rank_to_T) == Trank)
        if i.size == 0:
            sys.exit("Internal error in TemperatureRX_MPI.find_procrank_from_Trank")
        return int(i[0, 0])

    # -- one exchange round (mc_mpi.py:102-140) --------------------------------------------------
    def Xtrial(self, XCscheme=-1):
        self._exchange(XCscheme % 2)
        self.rank_to_T = self._lat.get_replica_to_T().astype(int)
        self._refresh_states()

    def _exchange(self, scheme):
        lat = self._lat
        lat.rx_energies()
        if self.procs > 1:
            self._allgather_energies()
        lat.rx_exchange(scheme)

    def _advance(self, n, done, every, scheme):
        """n sweeps from sweep number ``done`` with an exchange after every ``every``-th sweep (mc_mpi.py:165-169);
        returns the scheme of the next exchange.  Where the batch-resident kernel applies (one GPU, or several
        connected through peer memory) whole runs of exchange cycles are one launch: sweeps, exact energies and
        the pair decisions all happen inside it."""
        lat = self._lat
        while n > 0:
            to_x = every - done % every  # sweeps up to and including the next exchange point
            if n < to_x:
                lat.sweep(n)
                return scheme
            if self._fused:
                if to_x == every:
                    ncyc = n // every
                    lat.rx_cycle(ncyc, every, scheme)
                    scheme = (scheme + ncyc) % 2
                    k = ncyc * every
                else:
                    lat.rx_cycle(1, to_x, scheme)
                    scheme = (scheme + 1) % 2
                    k = to_x
            else:
                lat.sweep(to_x)
                self._exchange(scheme)
                scheme = (scheme + 1) % 2
                k = to_x
            n -= k
            done += k
        return scheme

    def _allgather_energies(self):
        if self.comm.backend == "nccl":
            import torch
            if self._energies_t is None:
                ptr, _ = self._lat.rx_buffers()
                self._energies_t = torch.as_tensor(_DeviceArray(ptr, self.nreplicas, "<f8"), device="cuda")
            full = self._energies_t
            mine = full[self.first:self.first + self.nlocal]
            self.comm.dist.all_gather_into_tensor(full, mine)  # in place, on the handle's stream
        else:  # host path (gloo): used by the CPU tests of the plumbing only
            raise RuntimeError("multi-process replica exchange needs the NCCL backend (one process per GPU)")

    def _run_device_stats(self, nsteps, RXtrial_frequency, sample_frequency, observer):
        """The run loop with sampling on the device: nothing in it synchronises with the host."""
        lat = self._lat
        lat.stats_reset(*self.model._pmc_energy_coeffs())
        keep = getattr(observer, "keep_series", False)
        if keep:
            lat.series_reset(max(1, nsteps // sample_frequency), *self.model._pmc_energy_coeffs())
        XCscheme, done, nsample = 0, 0, 0
        while done < nsteps:
            chunk = min(nsteps - done, sample_frequency - done % sample_frequency)
            XCscheme = self._advance(chunk, done, RXtrial_frequency, XCscheme)
            done += chunk
            if done % sample_frequency == 0:
                lat.stats_accumulate()
                if keep:
                    lat.series_record()
                nsample += 1
        sums, counts = lat.stats_get()
        if keep:
            if self.procs > 1:  # every GPU recorded the entries of its own replicas: add them up, everywhere
                full = self.comm.allreduce_sum_host(lat.series_get().ravel())
                lat.series_put(full)
            observer._lat = lat
        if self.procs > 1:
            sums = self.comm.allreduce_sum_host(sums)
            counts = np.rint(self.comm.allreduce_sum_host(counts.astype(np.float64))).astype(np.uint64)
        self.rank_to_T = lat.get_replica_to_T().astype(int)
        self._configs_changed()
        self._refresh_states(with_energy=True)
        _, accepted = lat.rx_stats()
        self.accept_count = int(accepted.sum())
        observer._finish(sums, counts, self.kTs)
        if nsample:
            return [[observer.mean_E[i], observer.mean_absM[i]] for i in range(len(self.kTs))]

    # -- the run loop (mc_mpi.py:143-209) --------------------------------------------------------
    def run(self, nsteps, RXtrial_frequency, sample_frequency=verylargeint, print_frequency=verylargeint,
            observer=None, subdirs=True, save_obs=True):
        if _is_observer(sample_frequency):
            observer, sample_frequency = sample_frequency, verylargeint
        if observer is None:
            observer = _default_observer
        self.subdirs = subdirs
        self.accept_count = 0
        self._sync_host_edits()
        self._lat.set_replica_to_T(np.asarray(self.rank_to_T, dtype=np.int32))
        self._lat.reduce_observables()  # mycalc.energy = model.energy(config), mc_mpi.py:153
        self._refresh_states(with_energy=True)
        with open(os.devnull, "w") as devnull:
            test_observe = observer.observe(self.states[0], devnull, lprint=False)
        observe = hasattr(test_observe, "__add__")
        obs = None  # one accumulator per replica, as every reference rank has its own (mc_mpi.py:157,172)
        if hasattr(test_observe, "__iter__"):
            obs = np.zeros([self.nlocal, len(self.kTs), len(test_observe)])
        nsample = 0
        XCscheme = 0
        if getattr(observer, "_pmc_device", False) and observe:
            return self._run_device_stats(nsteps, RXtrial_frequency, sample_frequency, observer)
        outs = [open(self._path(loc, "obs.dat"), "a") for loc in range(self.nlocal)]
        done = 0
        try:
            while done < nsteps:
                chunk = nsteps - done
                if observe:
                    chunk = min(chunk, sample_frequency - done % sample_frequency)
                XCscheme = self._advance(chunk, done, RXtrial_frequency, XCscheme)
                done += chunk
                if observe and done % sample_frequency == 0:
                    self.rank_to_T = self._lat.get_replica_to_T().astype(int)
                    self._configs_changed()
                    self._refresh_states(with_energy=True)
                    for loc in range(self.nlocal):
                        st = self.states[loc]
                        obs_step = observer.observe(st, outs[loc], done % print_frequency == 0)
                        T = self.rank_to_T[self.first + loc]
                        obs[loc, T] += obs_step
                        if save_obs:
                            self.obs_save[loc].append(obs_step)
                            self.Trank_hist[loc].append(T)
                            self.kT_hist[loc].append(st.kT)
                    nsample += 1
        finally:
            for f in outs:
                f.close()
        self.rank_to_T = self._lat.get_replica_to_T().astype(int)
        self._configs_changed()
        self._refresh_states(with_energy=True)
        tried, accepted = self._lat.rx_stats()
        self.accept_count = int(accepted.sum())

        for loc in range(self.nlocal):
            with open(self._path(loc, "calc.pickle"), "wb") as f:
                pickle.dump(self.states[loc].config, f)
            if save_obs:
                o = np.array(self.obs_save[loc])
                t = np.array(self.Trank_hist[loc])
                k = np.array(self.kT_hist[loc])
                if hasattr(self, "obs_save0"):
                    o = np.concatenate((self.obs_save0[loc], o)) if o.size else self.obs_save0[loc]
                    t = np.concatenate((self.Trank_hist0[loc], t)) if t.size else self.Trank_hist0[loc]
                    k = np.concatenate((self.kT_hist0[loc], k)) if k.size else self.kT_hist0[loc]
                for name, arr in (("obs_save.npy", o), ("Trank_hist.npy", t), ("kT_hist.npy", k)):
                    with open(self._path(loc, name), "wb") as f:
                        np.save(f, arr, False)
        if self.rank == 0:
            with open("rank_to_T.pickle", "wb") as f:
                pickle.dump(self.rank_to_T, f)
            with open("kTs.npy", "wb") as f:
                np.save(f, self.kTs, False)
            with open("rx_state.pickle", "wb") as f:  # extension, see reload()
                pickle.dump({"sweep_counter": self._lat.sweep_counter, "rx_round": self._lat.rx_round}, f)

        if nsample != 0:
            obs = np.array(obs)
            obs /= nsample
            partial = np.zeros(obs.shape[1:])
            for loc in range(self.nlocal):  # replica order = the rank order of the reference's Allreduce
                partial = partial + obs[loc]
            if self.procs > 1:
                parts = self.comm.allgather_f64_host(partial.ravel()).reshape((self.procs,) + partial.shape)
                obs_buffer = np.zeros(partial.shape)
                for g in range(self.procs):
                    obs_buffer = obs_buffer + parts[g]
            else:
                obs_buffer = partial
            args_info = observer.obs_info(self.states[0])
            return [obs_decode(args_info, obs_buffer[i]) for i in range(len(self.kTs))]
