"""GPU (one device): row-slab decomposition of one lattice (row f4).  Several slabs driven from one process on
one stream must follow exactly the trajectory of the whole-lattice run: same random stream (indexed by
whole-lattice rows), halo rows written by the neighbours' sweep kernels.  The multi-process / multi-GPU form
(CUDA IPC peer mappings + flag barrier) is checked by tests/mgpu_slab_check.py under torchrun."""
import numpy as np
import pytest

from oracle import ising as o_ising

pytestmark = pytest.mark.gpu


def _whole(shape, seed, thr, nsweeps):
    from py_mc_b200 import _cabi
    lat = _cabi.Lattices(_cabi.MODEL_ISING, shape[0], shape[1], seed=seed, layout=_cabi.LAYOUT_INT8)
    lat.init_random()
    lat.set_accept_tables(thr)
    lat.sweep(nsweeps)
    return lat.download(0), lat.tracked()


@pytest.mark.parametrize("nslabs", [1, 2, 4])
def test_slabs_follow_the_whole_lattice_trajectory(nslabs):
    from py_mc_b200 import _cabi
    shape, seed, nsweeps = (128, 1024), 321, 7
    thr = o_ising.threshold_table(-1.0, 2.3)[1][None, :]
    want, (wb, wm) = _whole(shape, seed, thr, nsweeps)
    slabs = [_cabi.Lattices(_cabi.MODEL_ISING, shape[0], shape[1], seed=seed, layout=_cabi.LAYOUT_INT8,
                            slab=(g, nslabs)) for g in range(nslabs)]
    if nslabs > 1:
        for g, s in enumerate(slabs):
            s.set_stream(0)  # one stream for all slabs: the sweeps below are ordered by issue order
            s.slab_connect_local(slabs[(g - 1) % nslabs], slabs[(g + 1) % nslabs])
    for s in slabs:
        s.init_random()
        s.set_accept_tables(thr)
        assert s.sweep_kernel_for(1) == (1, "sweep_fused_i8_kernel")
    # the hot start is a function of whole-lattice rows: the slabs together are the single-GPU start
    start = np.concatenate([s.download(0) for s in slabs])
    assert np.array_equal(start, o_ising.init_config(seed, 0, *shape))
    for _ in range(nsweeps):
        for s in slabs:
            s.sweep(1)
    got = np.concatenate([s.download(0) for s in slabs])
    assert np.array_equal(got, want)
    b = sum(int(s.tracked()[0][0]) for s in slabs)
    m = sum(int(s.tracked()[1][0]) for s in slabs)
    assert (b, m) == (int(wb[0]), int(wm[0]))
    rb = sum(int(s.reduce_observables()[0][0]) for s in slabs)
    assert rb == o_ising.bond_sum(want)


@pytest.mark.parametrize("shape", [(128, 1024), (2048, 2048)])
def test_single_slab_calls_of_several_sweeps_stay_on_chip(shape):
    """One slab over the whole lattice (its neighbours are itself): calls of two or more sweeps run on the batch-resident
    kernel with the rows above row 0 / below the last row travelling through the exchange buffer like between GPUs; the
    halo rows of the tiled kernel are refreshed afterwards, so single-sweep calls can follow.  Same trajectory and sums as
    the plain lattice.  (Real neighbours over CUDA IPC: tests/mgpu_slab_check.py.)"""
    from py_mc_b200 import _cabi
    seed, calls = 77, (3, 1, 4, 1, 1, 2)
    thr = o_ising.threshold_table(-1.0, 2.3)[1][None, :]
    want, (wb, wm) = _whole(shape, seed, thr, sum(calls))
    s = _cabi.Lattices(_cabi.MODEL_ISING, shape[0], shape[1], seed=seed, layout=_cabi.LAYOUT_INT8, slab=(0, 1))
    s.init_random()
    s.set_accept_tables(thr)
    assert s.sweep_kernel_for(2) == (1, "pbatch_kernel") and s.sweep_kernel_for(1) == (1, "sweep_fused_i8_kernel")
    for n in calls:
        s.sweep(n)
    assert s.slab_status() == 0
    assert np.array_equal(s.download(0), want)
    b, m = s.tracked()
    assert (int(b[0]), int(m[0])) == (int(wb[0]), int(wm[0]))
    rb, rm = s.reduce_observables()
    assert (int(rb[0]), int(rm[0])) == (int(wb[0]), int(wm[0]))
    s.close()


def test_upload_then_push_halos():
    from py_mc_b200 import _cabi
    shape, seed, nslabs = (64, 1024), 5, 2
    thr = o_ising.threshold_table(-1.0, 2.0)[1][None, :]
    rng = np.random.default_rng(0)
    cfg = rng.choice(np.array([-1, 1], dtype=np.int8), size=shape)
    whole = _cabi.Lattices(_cabi.MODEL_ISING, shape[0], shape[1], seed=seed, layout=_cabi.LAYOUT_INT8)
    whole.upload(0, cfg)
    whole.set_accept_tables(thr)
    whole.sweep(3)
    slabs = [_cabi.Lattices(_cabi.MODEL_ISING, shape[0], shape[1], seed=seed, layout=_cabi.LAYOUT_INT8,
                            slab=(g, nslabs)) for g in range(nslabs)]
    rows = shape[0] // nslabs
    for g, s in enumerate(slabs):
        s.set_stream(0)
        s.slab_connect_local(slabs[(g - 1) % nslabs], slabs[(g + 1) % nslabs])
        s.upload(0, cfg[g * rows:(g + 1) * rows])
        s.set_accept_tables(thr)
    for s in slabs:
        s.slab_push_halos()
    for _ in range(3):
        for s in slabs:
            s.sweep(1)
    assert np.array_equal(np.concatenate([s.download(0) for s in slabs]), whole.download(0))


def test_slab_shape_rules():
    from py_mc_b200 import _cabi
    with pytest.raises(_cabi.PymcError, match="lenX/nslabs"):
        _cabi.Lattices(_cabi.MODEL_ISING, 96, 1024, slab=(0, 2))     # 48 rows per slab: not whole tiles
    with pytest.raises(_cabi.PymcError, match="int8"):
        _cabi.Lattices(_cabi.MODEL_ISING, 64, 1024, layout=_cabi.LAYOUT_BITS, slab=(0, 2))
