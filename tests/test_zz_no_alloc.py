"""GPU: the observables entry point that allocates nothing, pymc_reduce_observables_static.

Some hosts may not call cudaMalloc once they are running -- all device memory belongs to another
allocator or to a fixed budget, or the loop cannot take the device-wide synchronisation an allocation
implies.  For them the library keeps the reduction's accumulators in a static device array (64 KiB per
GPU, 4096 replicas at most, shared by every handle on that GPU).  This file checks, through the C ABI,
that the entry point returns exactly what pymc_reduce_observables returns (the sums are integers:
tolerance zero) on consecutive calls, that what one call leaves in the shared array never shows up in
the next result, and that a batch beyond the maximum is refused with a message instead of overrunning
the array; that the free device memory reported by the driver is the same before and after the calls
(the property the name promises); and that a row-slab handle, whose reduction reads a halo row, gives the
allocating entry point's numbers too.  It is named to be collected last, and it must pass both on its own in a fresh process and
after the rest of the suite.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _lattice(nrep, shape, layout, seed):
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D
    lat = _cabi.Lattices(_cabi.MODEL_ISING, shape[0], shape[1], nreplicas=nrep, ntemps=1, seed=seed, layout=layout)
    lat.init_random()
    lat.set_accept_tables(ising2D(-1.0)._pmc_thresholds(2.3)[None, :])
    return lat


def test_matches_allocating_entry_point_on_consecutive_calls():
    from py_mc_b200 import _cabi
    big = _lattice(9, (12, 20), _cabi.LAYOUT_INT8, seed=11)
    for call in range(2):  # two calls in a row on the same handle, the lattice changing in between
        big.sweep(3)
        want_b, want_m = big.reduce_observables()
        got_b, got_m = big.reduce_observables_static()
        assert np.array_equal(got_b, want_b) and np.array_equal(got_m, want_m), call
        # the re-based tracked values agree as well
        big._tracked_cache = None
        tb, tm = big.tracked()
        assert np.array_equal(tb, want_b) and np.array_equal(tm, want_m)
    # a smaller batch on another handle reuses the front of the same static array: nothing of the
    # nine-replica result above may leak into it
    small = _lattice(2, (16, 128), _cabi.LAYOUT_BITS, seed=12)
    small.sweep(2)
    want_b, want_m = small.reduce_observables()
    for call in range(2):
        got_b, got_m = small.reduce_observables_static()
        assert np.array_equal(got_b, want_b) and np.array_equal(got_m, want_m), call
    # and the first handle still gets its own numbers afterwards
    want_b, want_m = big.reduce_observables()
    got_b, got_m = big.reduce_observables_static()
    assert np.array_equal(got_b, want_b) and np.array_equal(got_m, want_m)
    big.close()
    small.close()


def test_batch_beyond_static_maximum_is_refused():
    from py_mc_b200 import _cabi
    lat = _cabi.Lattices(_cabi.MODEL_ISING, 2, 2, nreplicas=4097, ntemps=1, seed=1)
    lat.init_random()
    ok_b, ok_m = lat.reduce_observables()  # the allocating entry point has no such limit
    assert ok_b.shape == (4097,)
    with pytest.raises(_cabi.PymcError, match="4096"):
        lat.reduce_observables_static()
    lat.close()
    # exactly at the maximum it works
    lat = _cabi.Lattices(_cabi.MODEL_ISING, 2, 2, nreplicas=4096, ntemps=1, seed=1)
    lat.init_random()
    want = lat.reduce_observables()
    got = lat.reduce_observables_static()
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    lat.close()


def _free_device_memory():
    """cudaMemGetInfo through the CUDA runtime the library itself is linked against."""
    import ctypes as C
    rt = C.CDLL("libcudart.so.12")
    free, total = C.c_size_t(0), C.c_size_t(0)
    assert rt.cudaMemGetInfo(C.byref(free), C.byref(total)) == 0
    return free.value


def test_device_memory_is_untouched_by_the_call():
    """A host that may not allocate needs more than equal numbers: no device allocation may happen inside the call.
    The first call on the device has already been made above or is made here before measuring (it only looks up
    the static array's address)."""
    from py_mc_b200 import _cabi
    lat = _lattice(5, (64, 256), _cabi.LAYOUT_INT8, seed=3)
    lat.sweep(2)
    lat.reduce_observables_static()          # warm: address look-up, module fully loaded
    lat.sweep(1)
    lat.synchronize()
    before = _free_device_memory()
    for _ in range(3):
        got = lat.reduce_observables_static()
    after = _free_device_memory()
    assert after == before, (before, after)
    want = lat.reduce_observables()
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    lat.close()


def test_row_slab_handles_agree_with_the_allocating_entry_point():
    """Two row slabs of one 128 x 1024 lattice driven by this process (pymc_slab_connect_local): the reduction of a
    slab reads the halo row above its first owned row."""
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D
    thr = ising2D(-1.0)._pmc_thresholds(2.3)[None, :]
    slabs = [_cabi.Lattices(_cabi.MODEL_ISING, 128, 1024, seed=9, slab=(k, 2)) for k in range(2)]
    for k, s in enumerate(slabs):
        s.set_stream(0)  # one stream for both slabs: the sweeps below are ordered by issue order
        s.slab_connect_local(slabs[(k - 1) % 2], slabs[(k + 1) % 2])
    for s in slabs:
        s.init_random()
        s.set_accept_tables(thr)
    for _ in range(3):
        for s in slabs:
            s.sweep(1)
    for s in slabs:
        s.synchronize()
    for s in slabs:
        want = s.reduce_observables()
        got = s.reduce_observables_static()
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    for s in slabs:
        s.close()
