"""ctypes binding of include/pymc_b200.h -- the only way Python reaches the kernels.

There is no CPU fallback: if the shared library has not been built, importing this module
raises, and so does every product class that needs it.
"""
import ctypes as C
import os

import numpy as np

from ._build import LIBPATH

ABI_VERSION = 1
MODEL_ISING, MODEL_LATGAS = 0, 1
LAYOUT_INT8, LAYOUT_BITS = 0, 1
KERNEL_AUTO, KERNEL_GENERIC, KERNEL_TWO_PASS, KERNEL_RESIDENT, KERNEL_PERSISTENT, KERNEL_PBATCH, KERNEL_ROWSTREAM = 0, 1, 2, 3, 4, 5, 6
NCLASS = {MODEL_ISING: 5, MODEL_LATGAS: 10}

# every symbol include/pymc_b200.h declares (tests check the library exports all of them)
SYMBOLS = [
    "pymc_abi_version", "pymc_last_error", "pymc_device_count", "pymc_create", "pymc_destroy",
    "pymc_set_stream", "pymc_reset_stream", "pymc_synchronize", "pymc_upload_config", "pymc_download_config",
    "pymc_init_random", "pymc_set_accept_tables", "pymc_set_replica_to_T", "pymc_get_replica_to_T",
    "pymc_sweep", "pymc_get_sweep_counter", "pymc_set_sweep_counter", "pymc_reduce_observables",
    "pymc_get_tracked", "pymc_propose", "pymc_timer_start", "pymc_timer_stop",
    "pymc_rx_setup", "pymc_rx_energies", "pymc_rx_buffers", "pymc_rx_exchange", "pymc_rx_cycle",
    "pymc_rx_get_energies", "pymc_rx_get_stats", "pymc_rx_get_round", "pymc_rx_set_round",
    "pymc_host_alloc", "pymc_host_free", "pymc_reduce_observables_static", "pymc_sweep_kernel_info",
    "pymc_set_pair_tables", "pymc_pair_sweep", "pymc_mixed_sweep", "pymc_get_pair_counter", "pymc_set_pair_counter",
    "pymc_discs_create", "pymc_discs_destroy", "pymc_discs_upload", "pymc_discs_download", "pymc_discs_run",
    "pymc_discs_propose", "pymc_discs_observe", "pymc_discs_counters", "pymc_discs_set_trial_counter",
    "pymc_create_slab", "pymc_slab_ipc_export", "pymc_slab_connect", "pymc_slab_connect_local",
    "pymc_slab_push_halos", "pymc_slab_status",
    "pymc_discs_timer", "pymc_stats_reset", "pymc_stats_accumulate", "pymc_stats_get",
    "pymc_debug_raise_async_error", "pymc_rx_ipc_export", "pymc_rx_connect",
    "pymc_upload_config_async", "pymc_download_config_async", "pymc_get_tracked_async", "pymc_sweep_kernel_for",
    "pymc_series_reset", "pymc_series_record", "pymc_series_count", "pymc_series_get", "pymc_series_put", "pymc_series_binning",
]


class Params(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32), ("device", C.c_int32), ("model", C.c_int32), ("layout", C.c_int32),
        ("lenX", C.c_int32), ("lenY", C.c_int32), ("nreplicas", C.c_int32), ("replica_offset", C.c_int32),
        ("nreplicas_global", C.c_int32), ("ntemps", C.c_int32), ("seed", C.c_uint64),
        ("kernel", C.c_int32), ("reserved", C.c_int32),
    ]


class PymcError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("pymc_b200 error %d: %s" % (code, message))
        self.code = code


def _load():
    global LIBPATH
    override = os.environ.get("PYMC_B200_LIB")  # alternative build of the same ABI (kernel experiments)
    if override:
        LIBPATH = override
    else:
        if not os.path.exists(LIBPATH):  # build an absent library when a compiler is at hand; never fall back
            try:
                from . import _build
                _build.build()
            except Exception as exc:
                raise ImportError("py_mc_b200: %s is missing and cannot be built: %s" % (LIBPATH, exc))
    if not os.path.exists(LIBPATH):
        raise ImportError(
            "py_mc_b200: %s is missing. Build it with `python -m py_mc_b200._build` "
            "(needs nvcc; there is no CPU fallback)." % LIBPATH)
    lib = C.CDLL(LIBPATH)
    P, I, I64 = C.c_void_p, C.c_int, C.c_int64
    lib.pymc_abi_version.restype = I
    lib.pymc_last_error.restype = C.c_char_p
    sigs = {
        "pymc_device_count": [C.POINTER(I)],
        "pymc_create": [C.POINTER(Params), C.POINTER(P)],
        "pymc_destroy": [P],
        "pymc_set_stream": [P, P],
        "pymc_reset_stream": [P],
        "pymc_synchronize": [P],
        "pymc_debug_raise_async_error": [P, I],
        "pymc_upload_config": [P, I, P, C.c_size_t],
        "pymc_download_config": [P, I, P, C.c_size_t],
        "pymc_upload_config_async": [P, I, P, C.c_size_t],
        "pymc_download_config_async": [P, I, P, C.c_size_t],
        "pymc_get_tracked_async": [P, P],
        "pymc_init_random": [P],
        "pymc_set_accept_tables": [P, P, I, I],
        "pymc_set_replica_to_T": [P, P, I],
        "pymc_get_replica_to_T": [P, P, I],
        "pymc_sweep": [P, I64],
        "pymc_sweep_kernel_info": [P, C.POINTER(I), C.POINTER(C.c_char_p)],
        "pymc_sweep_kernel_for": [P, I64, C.POINTER(I), C.POINTER(C.c_char_p)],
        "pymc_get_sweep_counter": [P, C.POINTER(C.c_uint64)],
        "pymc_set_sweep_counter": [P, C.c_uint64],
        "pymc_reduce_observables": [P, P, P],
        "pymc_get_tracked": [P, P, P],
        "pymc_reduce_observables_static": [P, P, P],
        "pymc_propose": [P, I, P, P, I, P, P],
        "pymc_set_pair_tables": [P, P, I, I],
        "pymc_pair_sweep": [P, I64, C.c_uint32],
        "pymc_mixed_sweep": [P, I64, C.c_uint32, C.c_uint32],
        "pymc_get_pair_counter": [P, C.POINTER(C.c_uint64)],
        "pymc_set_pair_counter": [P, C.c_uint64],
        "pymc_discs_create": [I, I, I, I, C.c_double, C.c_double, C.c_double, C.c_uint64, I, C.POINTER(P)],
        "pymc_discs_destroy": [P],
        "pymc_discs_upload": [P, I, P],
        "pymc_discs_download": [P, I, P],
        "pymc_discs_run": [P, I64],
        "pymc_discs_propose": [P, I, C.POINTER(I), P, C.POINTER(I)],
        "pymc_discs_observe": [P, C.c_double, I, P, P],
        "pymc_discs_counters": [P, C.POINTER(C.c_uint64), P],
        "pymc_discs_set_trial_counter": [P, C.c_uint64],
        "pymc_discs_timer": [P, I64, C.POINTER(C.c_float)],
        "pymc_create_slab": [C.POINTER(Params), I, I, C.POINTER(P)],
        "pymc_slab_ipc_export": [P, P, C.c_size_t],
        "pymc_slab_connect": [P, P, P, C.c_size_t],
        "pymc_slab_connect_local": [P, P, P],
        "pymc_slab_push_halos": [P],
        "pymc_slab_status": [P, C.POINTER(I)],
        "pymc_stats_reset": [P, C.c_double, C.c_double],
        "pymc_stats_accumulate": [P],
        "pymc_stats_get": [P, P, P, I, I],
        "pymc_series_reset": [P, I, C.c_double, C.c_double],
        "pymc_series_record": [P],
        "pymc_series_count": [P, C.POINTER(I)],
        "pymc_series_get": [P, P, C.c_size_t],
        "pymc_series_put": [P, P, I],
        "pymc_series_binning": [P, I, I, P, P],
        "pymc_rx_setup": [P, P, I, C.c_double, C.c_double],
        "pymc_rx_energies": [P],
        "pymc_rx_buffers": [P, C.POINTER(P), C.POINTER(P)],
        "pymc_rx_exchange": [P, I],
        "pymc_rx_ipc_export": [P, P, C.c_size_t],
        "pymc_rx_connect": [P, P, C.c_size_t, I, I],
        "pymc_rx_cycle": [P, I64, I64, I],
        "pymc_rx_get_energies": [P, P, I],
        "pymc_rx_get_stats": [P, P, P, I],
        "pymc_rx_get_round": [P, C.POINTER(C.c_uint64)],
        "pymc_rx_set_round": [P, C.c_uint64],
        "pymc_host_alloc": [C.c_size_t, C.POINTER(P)],
        "pymc_host_free": [P],
        "pymc_timer_start": [P],
        "pymc_timer_stop": [P, C.POINTER(C.c_float)],
    }
    for name, argtypes in sigs.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = I
    if lib.pymc_abi_version() != ABI_VERSION:
        raise ImportError("py_mc_b200: library ABI %d, binding expects %d; rebuild with "
                          "`python -m py_mc_b200._build --force`" % (lib.pymc_abi_version(), ABI_VERSION))
    return lib


lib = _load()


def check(rc):
    if rc != 0:
        raise PymcError(rc, lib.pymc_last_error().decode("utf-8", "replace"))


def device_count():
    n = C.c_int(0)
    check(lib.pymc_device_count(C.byref(n)))
    return n.value


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class _PinnedOwner:
    def __init__(self, ptr):
        self.ptr = ptr

    def __del__(self):
        try:
            lib.pymc_host_free(C.c_void_p(self.ptr))
        except Exception:
            pass


def pinned_empty(shape, dtype=np.int8):
    """numpy array in page-locked host memory (freed when the array and its views are gone)."""
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    p = C.c_void_p()
    check(lib.pymc_host_alloc(n, C.byref(p)))
    owner = _PinnedOwner(p.value)
    buf = (C.c_char * max(n, 1)).from_address(p.value)
    buf._owner = owner  # keeps the allocation alive as long as the ctypes buffer (hence the array) lives
    return np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)


class Lattices:
    """A batch of replicas of one lattice model on one GPU (thin owner of a pymc_handle)."""

    def __init__(self, model, lenX, lenY, nreplicas=1, ntemps=1, seed=0, layout=LAYOUT_INT8, device=0,
                 replica_offset=0, nreplicas_global=None, kernel=KERNEL_AUTO, slab=None):
        self._h = C.c_void_p()
        self.model, self.lenX, self.lenY = int(model), int(lenX), int(lenY)
        self.nreplicas, self.ntemps, self.layout = int(nreplicas), int(ntemps), int(layout)
        self.replica_offset = int(replica_offset)
        self.nreplicas_global = int(nreplicas if nreplicas_global is None else nreplicas_global)
        self.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
        self.device = int(device)
        p = Params(ABI_VERSION, self.device, self.model, self.layout, self.lenX, self.lenY, self.nreplicas,
                   self.replica_offset, self.nreplicas_global, self.ntemps, self.seed, int(kernel), 0)
        self.slab = slab  # (index, count): this handle owns one row slab of the lenX x lenY lattice
        if slab is None:
            check(lib.pymc_create(C.byref(p), C.byref(self._h)))
        else:
            check(lib.pymc_create_slab(C.byref(p), int(slab[0]), int(slab[1]), C.byref(self._h)))
            self.rows = self.lenX // int(slab[1])
        self._tracked_cache = None  # (bond, second) valid until the configurations change

    # -- lifetime
    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib.pymc_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def nsites(self):
        return self.lenX * self.lenY

    def set_stream(self, cuda_stream_ptr):
        """Run on the given cudaStream_t (an int; 0 = the legacy default stream)."""
        check(lib.pymc_set_stream(self._h, C.c_void_p(int(cuda_stream_ptr))))

    def reset_stream(self):
        check(lib.pymc_reset_stream(self._h))

    def synchronize(self):
        check(lib.pymc_synchronize(self._h))

    def debug_raise_async_error(self, which):
        """Test hook: raise the flag a timed-out in-kernel barrier raises (0 persistent kernel, 1 slab barrier)."""
        check(lib.pymc_debug_raise_async_error(self._h, int(which)))

    # -- configurations
    def _local_shape(self):
        return (self.rows if self.slab else self.lenX, self.lenY)

    def upload(self, replica, config_int8):
        a = np.ascontiguousarray(config_int8, dtype=np.int8)
        if a.shape != self._local_shape():
            raise ValueError("config shape %s, lattice (slab) is %dx%d" % ((a.shape,) + self._local_shape()))
        self._tracked_cache = None
        check(lib.pymc_upload_config(self._h, replica, _ptr(a), a.size))

    def download(self, replica, out=None):
        if out is None:
            out = np.empty(self._local_shape(), dtype=np.int8)
        assert out.dtype == np.int8 and out.flags.c_contiguous and out.size == int(np.prod(self._local_shape()))
        check(lib.pymc_download_config(self._h, replica, _ptr(out), out.size))
        return out

    def upload_async(self, replica, pinned_int8):
        """Enqueue the upload and return; ``pinned_int8`` (from ``pinned_empty``) must stay untouched until ``synchronize``."""
        a = pinned_int8
        if a.dtype != np.int8 or not a.flags.c_contiguous or a.shape != self._local_shape():
            raise ValueError("upload_async needs a C-contiguous int8 array of shape %s" % (self._local_shape(),))
        self._tracked_cache = None
        check(lib.pymc_upload_config_async(self._h, replica, _ptr(a), a.size))

    def download_async(self, replica, pinned_out):
        """Enqueue the download into ``pinned_out`` (from ``pinned_empty``); its contents are valid after ``synchronize``."""
        a = pinned_out
        if a.dtype != np.int8 or not a.flags.c_contiguous or a.shape != self._local_shape():
            raise ValueError("download_async needs a C-contiguous int8 array of shape %s" % (self._local_shape(),))
        check(lib.pymc_download_config_async(self._h, replica, _ptr(a), a.size))

    def tracked_async(self, pinned_pairs):
        """Enqueue the copy of the tracked (bond, second) pairs into ``pinned_pairs`` (int64 [nreplicas][2])."""
        a = pinned_pairs
        if a.dtype != np.int64 or not a.flags.c_contiguous or a.size != 2 * self.nreplicas:
            raise ValueError("tracked_async needs a C-contiguous int64 array of %d entries" % (2 * self.nreplicas))
        check(lib.pymc_get_tracked_async(self._h, _ptr(a)))

    def init_random(self):
        self._tracked_cache = None
        check(lib.pymc_init_random(self._h))

    # -- tables
    def set_accept_tables(self, thr):
        t = np.ascontiguousarray(thr, dtype=np.uint32).reshape(self.ntemps, NCLASS[self.model])
        check(lib.pymc_set_accept_tables(self._h, _ptr(t), t.shape[0], t.shape[1]))

    def set_replica_to_T(self, mapping):
        m = np.ascontiguousarray(mapping, dtype=np.int32)
        check(lib.pymc_set_replica_to_T(self._h, _ptr(m), m.size))

    def get_replica_to_T(self):
        m = np.empty(self.nreplicas_global, dtype=np.int32)
        check(lib.pymc_get_replica_to_T(self._h, _ptr(m), m.size))
        return m

    # -- hot path
    def sweep(self, nsweeps=1):
        self._tracked_cache = None
        check(lib.pymc_sweep(self._h, int(nsweeps)))

    def sweep_kernel_info(self):
        """(launches per sweep, kernel name) of the sweep path this handle uses."""
        n, name = C.c_int(0), C.c_char_p()
        check(lib.pymc_sweep_kernel_info(self._h, C.byref(n), C.byref(name)))
        return n.value, name.value.decode()

    def sweep_kernel_for(self, nsweeps):
        """(launches per sweep, kernel name) of a call of exactly ``nsweeps`` sweeps."""
        n, name = C.c_int(0), C.c_char_p()
        check(lib.pymc_sweep_kernel_for(self._h, int(nsweeps), C.byref(n), C.byref(name)))
        return n.value, name.value.decode()

    @property
    def sweep_counter(self):
        c = C.c_uint64(0)
        check(lib.pymc_get_sweep_counter(self._h, C.byref(c)))
        return c.value

    @sweep_counter.setter
    def sweep_counter(self, v):
        check(lib.pymc_set_sweep_counter(self._h, int(v)))

    # -- observables
    def reduce_observables(self):
        b = np.empty(self.nreplicas, dtype=np.int64)
        s = np.empty(self.nreplicas, dtype=np.int64)
        check(lib.pymc_reduce_observables(self._h, _ptr(b), _ptr(s)))
        self._tracked_cache = (b, s)
        return b, s

    def reduce_observables_static(self):
        """Same as ``reduce_observables`` through the entry point that allocates nothing (for hosts that
        may not call cudaMalloc while running); at most 4096 replicas, shares one static device array
        per GPU, so do not overlap two such calls on one GPU."""
        b = np.empty(self.nreplicas, dtype=np.int64)
        s = np.empty(self.nreplicas, dtype=np.int64)
        check(lib.pymc_reduce_observables_static(self._h, _ptr(b), _ptr(s)))
        self._tracked_cache = (b, s)
        return b, s

    def tracked(self):
        if self._tracked_cache is None:
            b = np.empty(self.nreplicas, dtype=np.int64)
            s = np.empty(self.nreplicas, dtype=np.int64)
            check(lib.pymc_get_tracked(self._h, _ptr(b), _ptr(s)))
            self._tracked_cache = (b, s)
        return self._tracked_cache

    def propose(self, replica, xs, ys):
        xs = np.ascontiguousarray(xs, dtype=np.int32)
        ys = np.ascontiguousarray(ys, dtype=np.int32)
        site = np.empty(xs.size, dtype=np.int32)
        nn = np.empty(xs.size, dtype=np.int32)
        check(lib.pymc_propose(self._h, replica, _ptr(xs), _ptr(ys), xs.size, _ptr(site), _ptr(nn)))
        return site, nn

    # -- particle-hole exchange (lattice gas)
    def set_pair_tables(self, thr):
        t = np.ascontiguousarray(thr, dtype=np.uint32).reshape(self.ntemps, 7)
        check(lib.pymc_set_pair_tables(self._h, _ptr(t), t.shape[0], t.shape[1]))

    def pair_sweep(self, nsweeps=1, class_mask=0xFFFF):
        self._tracked_cache = None
        check(lib.pymc_pair_sweep(self._h, int(nsweeps), int(class_mask)))

    def mixed_sweep(self, nsteps=1, mask_even=0x00FF, mask_odd=0xFF00):
        """nsteps x (flip sweep + pair pass over the classes the parity of the pair counter selects)."""
        self._tracked_cache = None
        check(lib.pymc_mixed_sweep(self._h, int(nsteps), int(mask_even), int(mask_odd)))

    @property
    def pair_counter(self):
        c = C.c_uint64(0)
        check(lib.pymc_get_pair_counter(self._h, C.byref(c)))
        return c.value

    @pair_counter.setter
    def pair_counter(self, v):
        check(lib.pymc_set_pair_counter(self._h, int(v)))

    # -- row slabs (one lattice over several GPUs)
    def slab_ipc_export(self):
        buf = (C.c_char * 256)()
        check(lib.pymc_slab_ipc_export(self._h, buf, 256))
        return bytes(buf)

    def slab_connect(self, up_handles, down_handles):
        check(lib.pymc_slab_connect(self._h, C.c_char_p(up_handles), C.c_char_p(down_handles), 256))

    def slab_connect_local(self, up, down):
        check(lib.pymc_slab_connect_local(self._h, up._h, down._h))

    def slab_push_halos(self):
        self._tracked_cache = None
        check(lib.pymc_slab_push_halos(self._h))

    def slab_status(self):
        st = C.c_int(0)
        check(lib.pymc_slab_status(self._h, C.byref(st)))
        return st.value

    # -- device-side statistics
    NSTAT = 6

    def stats_reset(self, c_bond, c_second=0.0):
        check(lib.pymc_stats_reset(self._h, float(c_bond), float(c_second)))

    def stats_accumulate(self):
        check(lib.pymc_stats_accumulate(self._h))

    def stats_get(self):
        """(sums [ntemps, 6] = sum of E, E^2, |M|, M^2, M^4, M;  counts [ntemps])."""
        sums = np.empty((self.ntemps, self.NSTAT), dtype=np.float64)
        counts = np.empty(self.ntemps, dtype=np.uint64)
        check(lib.pymc_stats_get(self._h, _ptr(sums), _ptr(counts), self.ntemps, self.NSTAT))
        return sums, counts

    # -- time series by temperature + blocking analysis on the device
    def series_reset(self, capacity, c_bond, c_second=0.0):
        check(lib.pymc_series_reset(self._h, int(capacity), float(c_bond), float(c_second)))

    def series_record(self):
        check(lib.pymc_series_record(self._h))

    def series_count(self):
        n = C.c_int(0)
        check(lib.pymc_series_count(self._h, C.byref(n)))
        return n.value

    def series_get(self):
        """[sample, T, (E, |M|)] as recorded on this GPU."""
        out = np.empty((self.series_count(), self.ntemps, 2), dtype=np.float64)
        check(lib.pymc_series_get(self._h, _ptr(out), out.size))
        return out

    def series_put(self, series):
        a = np.ascontiguousarray(series, dtype=np.float64).reshape(-1, self.ntemps, 2)
        check(lib.pymc_series_put(self._h, _ptr(a), a.shape[0]))

    def series_binning(self, discard, nlevels):
        """(mean [T, 2] over samples [discard:], errors [T, 2, nlevels + 1] of binning() over all samples)."""
        mean = np.empty((self.ntemps, 2), dtype=np.float64)
        err = np.empty((self.ntemps, 2, nlevels + 1), dtype=np.float64)
        check(lib.pymc_series_binning(self._h, int(discard), int(nlevels), _ptr(mean), _ptr(err)))
        return mean, err

    # -- replica exchange
    def rx_setup(self, betas, c_bond, c_second=0.0):
        b = np.ascontiguousarray(betas, dtype=np.float64)
        check(lib.pymc_rx_setup(self._h, _ptr(b), b.size, float(c_bond), float(c_second)))

    def rx_energies(self):
        check(lib.pymc_rx_energies(self._h))

    def rx_buffers(self):
        """Device pointers (ints) of energies[R] (float64) and replica_to_T[R] (int32)."""
        e, m = C.c_void_p(), C.c_void_p()
        check(lib.pymc_rx_buffers(self._h, C.byref(e), C.byref(m)))
        return e.value, m.value

    def rx_exchange(self, xcscheme):
        check(lib.pymc_rx_exchange(self._h, int(xcscheme)))

    def rx_ipc_export(self):
        """128 bytes: CUDA IPC handles of the array the other GPUs store their replicas' sums into (in-kernel exchange)."""
        buf = (C.c_char * 128)()
        check(lib.pymc_rx_ipc_export(self._h, buf, 128))
        return bytes(buf)

    def rx_connect(self, all_handles, nranks, rank):
        """all_handles: the exports of all GPUs concatenated in rank order."""
        check(lib.pymc_rx_connect(self._h, C.c_char_p(all_handles), len(all_handles), int(nranks), int(rank)))

    def rx_connect_group(self, dist):
        """Exchange the IPC handles over a torch.distributed group (once) and connect: afterwards rx_cycle on this
        partitioned handle is one launch per GPU with the exchange done through peer memory inside the kernel."""
        import torch
        mine = torch.frombuffer(bytearray(self.rx_ipc_export()), dtype=torch.uint8).cuda()
        world = dist.get_world_size()
        out = torch.empty(128 * world, dtype=torch.uint8, device="cuda")
        dist.all_gather_into_tensor(out, mine)
        self.rx_connect(bytes(out.cpu().numpy().tobytes()), world, dist.get_rank())
        dist.barrier()

    def rx_cycle(self, ncycles, sweeps_per_exchange, first_scheme=0):
        self._tracked_cache = None
        check(lib.pymc_rx_cycle(self._h, int(ncycles), int(sweeps_per_exchange), int(first_scheme)))

    def rx_get_energies(self):
        e = np.empty(self.nreplicas_global, dtype=np.float64)
        check(lib.pymc_rx_get_energies(self._h, _ptr(e), e.size))
        return e

    def rx_stats(self):
        t = np.empty(self.ntemps, dtype=np.uint64)
        a = np.empty(self.ntemps, dtype=np.uint64)
        check(lib.pymc_rx_get_stats(self._h, _ptr(t), _ptr(a), t.size))
        return t, a

    @property
    def rx_round(self):
        c = C.c_uint64(0)
        check(lib.pymc_rx_get_round(self._h, C.byref(c)))
        return c.value

    @rx_round.setter
    def rx_round(self, v):
        check(lib.pymc_rx_set_round(self._h, int(v)))

    # -- timing
    def timer_start(self):
        check(lib.pymc_timer_start(self._h))

    def timer_stop(self):
        ms = C.c_float(0)
        check(lib.pymc_timer_stop(self._h, C.byref(ms)))
        return ms.value


class Discs:
    """A batch of hard-disc replicas on one GPU (owner of a pymc_discs handle)."""

    def __init__(self, nreplicas, ndisc, Lcell, rdisc, maxstep, seed=0, isotropic=False, device=0, replica_offset=0):
        self._h = C.c_void_p()
        self.nreplicas, self.ndisc = int(nreplicas), int(ndisc)
        check(lib.pymc_discs_create(int(device), self.nreplicas, int(replica_offset), self.ndisc, float(Lcell),
                                    float(rdisc), float(maxstep), int(seed) & 0xFFFFFFFFFFFFFFFF,
                                    1 if isotropic else 0, C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib.pymc_discs_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, replica, coors):
        a = np.ascontiguousarray(coors, dtype=np.float64)
        if a.shape != (self.ndisc, 2):
            raise ValueError("coors shape %s, expected (%d, 2)" % (a.shape, self.ndisc))
        check(lib.pymc_discs_upload(self._h, replica, _ptr(a)))

    def download(self, replica):
        a = np.empty((self.ndisc, 2), dtype=np.float64)
        check(lib.pymc_discs_download(self._h, replica, _ptr(a)))
        return a

    def run(self, ntrials):
        check(lib.pymc_discs_run(self._h, int(ntrials)))

    def timed_run(self, ntrials):
        ms = C.c_float(0)
        check(lib.pymc_discs_timer(self._h, int(ntrials), C.byref(ms)))
        return ms.value

    def propose(self, replica):
        i, ov = C.c_int(0), C.c_int(0)
        new = np.empty(2, dtype=np.float64)
        check(lib.pymc_discs_propose(self._h, replica, C.byref(i), _ptr(new), C.byref(ov)))
        return i.value, new, bool(ov.value)

    def observe(self, dr=0.0, nbins=0):
        ov = np.empty(self.nreplicas, dtype=np.int32)
        counts = np.zeros((self.nreplicas, max(nbins, 0)), dtype=np.int64)
        check(lib.pymc_discs_observe(self._h, float(dr), int(nbins), _ptr(ov), _ptr(counts) if nbins else None))
        return ov.astype(bool), counts

    def counters(self):
        t = C.c_uint64(0)
        acc = np.empty(self.nreplicas, dtype=np.uint64)
        check(lib.pymc_discs_counters(self._h, C.byref(t), _ptr(acc)))
        return t.value, acc

    def set_trial_counter(self, v):
        check(lib.pymc_discs_set_trial_counter(self._h, int(v)))
