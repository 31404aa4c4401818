"""GPU: the persistent kernel (K11, pmc_persist.cuh) -- one lattice spread over the shared memory of all SMs for all
sweeps of a call -- against the per-sweep kernels: same trajectory, same tracked sums, both layouts and models,
uneven slabs, fewer rows than SMs, and the 4096 x 4096 case against the fused kernel."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def run_pair(model, shape, layout, thr, seed=5, calls=(2, 3), nreplicas=1):
    """Same seed on a persistent and a two-pass handle; returns [(configs, tracked, reduced, kernel name)] x 2."""
    from py_mc_b200 import _cabi
    out = []
    for kernel in (_cabi.KERNEL_PERSISTENT, _cabi.KERNEL_TWO_PASS):
        lat = _cabi.Lattices(model, shape[0], shape[1], nreplicas=nreplicas, seed=seed, layout=layout, kernel=kernel)
        lat.init_random()
        lat.set_accept_tables(thr[None, :])
        for n in calls:
            lat.sweep(n)
        lat.synchronize()
        out.append(([lat.download(r) for r in range(nreplicas)], lat.tracked(), lat.reduce_observables(),
                    lat.sweep_kernel_info()[1], lat.sweep_counter))
        lat.close()
    return out


def check_pair(out):
    a, b = out
    assert a[3] == "persistent_kernel" and b[3] != "persistent_kernel"
    for x, y in zip(a[0], b[0]):
        assert np.array_equal(x, y)
    for k in (0, 1):
        assert np.array_equal(a[1][k], b[1][k]) and np.array_equal(a[1][k], a[2][k])
    assert a[4] == b[4]


@pytest.mark.parametrize("shape", [(300, 256), (64, 128)])
@pytest.mark.parametrize("layout", [0, 1])
def test_ising_matches_two_pass_kernel(shape, layout):
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D
    check_pair(run_pair(_cabi.MODEL_ISING, shape, layout, ising2D(-1.0)._pmc_thresholds(2.269185314213022)))


def test_headline_lattice_matches_the_fused_kernel():
    """16384 x 16384 int8: the persistent kernel for calls of two sweeps or more, the fused kernel for one-sweep
    calls on the same buffers; trajectory and tracked sums equal those of the two-pass kernel.  (PYMC_KERNEL_AUTO
    takes the batch-resident kernel K12 here: tests/test_zx_gpu_pbatch.py.)"""
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D
    thr = ising2D(-1.0)._pmc_thresholds(2.269185314213022)
    L = 16384
    a = _cabi.Lattices(_cabi.MODEL_ISING, L, L, seed=20260611, kernel=_cabi.KERNEL_PERSISTENT)
    b = _cabi.Lattices(_cabi.MODEL_ISING, L, L, seed=20260611, kernel=_cabi.KERNEL_TWO_PASS)
    assert a.sweep_kernel_info()[1] == "persistent_kernel" and b.sweep_kernel_info()[1] == "sweep_tiled_i8_kernel"
    for lat in (a, b):
        lat.init_random()
        lat.set_accept_tables(thr[None, :])
    a.sweep(5)          # persistent
    a.sweep(1)          # a single sweep: fused kernel on the same buffers
    a.sweep(2)          # persistent again
    b.sweep(8)
    a.synchronize()
    assert np.array_equal(a.download(0), b.download(0))
    ta, tb = a.tracked(), b.tracked()
    assert int(ta[0][0]) == int(tb[0][0]) and int(ta[1][0]) == int(tb[1][0])
    ra = a.reduce_observables()
    assert int(ra[0][0]) == int(ta[0][0]) and int(ra[1][0]) == int(ta[1][0])
    a.close()
    b.close()


@pytest.mark.parametrize("shape", [(148, 384), (1024, 1024)])
@pytest.mark.parametrize("layout", [0, 1])
def test_ising_more_shapes(shape, layout):
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D
    check_pair(run_pair(_cabi.MODEL_ISING, shape, layout, ising2D(-1.0)._pmc_thresholds(2.269185314213022)))


@pytest.mark.parametrize("layout", [0, 1])
def test_lattice_gas_and_several_replicas(layout):
    from py_mc_b200 import _cabi
    from py_mc_b200.latgas import latticegas
    thr = latticegas(Eads=-0.3, J=-1.0, mu=0.2)._pmc_thresholds(0.9)
    check_pair(run_pair(_cabi.MODEL_LATGAS, (200, 256), layout, thr, calls=(4, 2), nreplicas=3))


def test_4096_against_the_fused_kernel():
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D
    thr = ising2D(-1.0)._pmc_thresholds(2.269185314213022)
    got = []
    for kernel in (_cabi.KERNEL_PERSISTENT, _cabi.KERNEL_TWO_PASS):
        lat = _cabi.Lattices(_cabi.MODEL_ISING, 4096, 4096, seed=21, kernel=kernel)
        lat.init_random()
        lat.set_accept_tables(thr[None, :])
        lat.sweep(6)
        lat.synchronize()
        got.append((lat.download(0), lat.tracked(), lat.sweep_kernel_info()[1]))
        lat.close()
    assert got[0][2] == "persistent_kernel"
    assert np.array_equal(got[0][0], got[1][0])
    assert int(got[0][1][0][0]) == int(got[1][1][0][0]) and int(got[0][1][1][0]) == int(got[1][1][1][0])


def test_headline_lattice_bit_packed_matches_the_row_kernel():
    """16384 x 16384 on the 1-bit layout: the persistent kernel (forced) against sweep_bits_rows_kernel (what
    PYMC_KERNEL_TWO_PASS selects for this layout), which is pinned to the scalar C oracle at 4096 x 4096 in
    test_gpu_ising.py -- the pin of this case to the reference is through that chain."""
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D
    thr = ising2D(-1.0)._pmc_thresholds(2.269185314213022)
    L = 16384
    got = []
    for kernel in (_cabi.KERNEL_PERSISTENT, _cabi.KERNEL_TWO_PASS):
        lat = _cabi.Lattices(_cabi.MODEL_ISING, L, L, seed=77, layout=_cabi.LAYOUT_BITS, kernel=kernel)
        lat.init_random()
        lat.set_accept_tables(thr[None, :])
        lat.sweep(4)
        lat.sweep(3)
        lat.synchronize()
        got.append((lat.download(0), lat.tracked(), lat.reduce_observables(), lat.sweep_kernel_info()[1]))
        lat.close()
    assert got[0][3] == "persistent_kernel" and got[1][3] == "sweep_bits_rows_kernel"
    assert np.array_equal(got[0][0], got[1][0])
    for k in (0, 1):
        assert int(got[0][1][k][0]) == int(got[1][1][k][0]) == int(got[0][2][k][0])


@pytest.mark.parametrize("layout", [0, 1])
@pytest.mark.parametrize("which", ["persistent_kernel", "pbatch_kernel"])
def test_lattice_gas_on_a_large_lattice(layout, which):
    """MODEL_LATGAS on a single lattice of 2^27 sites on both shared-memory-resident kernels (PYMC_KERNEL_AUTO takes
    the batch-resident one for calls of two sweeps or more); same trajectory and tracked sums as the per-colour
    kernels."""
    from py_mc_b200 import _cabi
    from py_mc_b200.latgas import latticegas
    thr = latticegas(Eads=-0.3, J=-1.0, mu=0.2)._pmc_thresholds(0.9)
    shape = (8192, 16384)
    got = []
    for kernel in (_cabi.KERNEL_PERSISTENT if which == "persistent_kernel" else _cabi.KERNEL_AUTO, _cabi.KERNEL_TWO_PASS):
        lat = _cabi.Lattices(_cabi.MODEL_LATGAS, shape[0], shape[1], seed=31, layout=layout, kernel=kernel)
        lat.init_random()
        lat.set_accept_tables(thr[None, :])
        lat.sweep(3)
        lat.sweep(2)
        lat.synchronize()
        got.append((lat.download(0), lat.tracked(), lat.reduce_observables(), lat.sweep_kernel_info()[1]))
        lat.close()
    assert got[0][3] == which and got[1][3] not in ("persistent_kernel", "pbatch_kernel")
    assert np.array_equal(got[0][0], got[1][0])
    for k in (0, 1):
        assert int(got[0][1][k][0]) == int(got[1][1][k][0]) == int(got[0][2][k][0])


def test_refused_where_it_cannot_run():
    from py_mc_b200 import _cabi
    with pytest.raises(_cabi.PymcError, match="persistent"):
        _cabi.Lattices(_cabi.MODEL_ISING, 8, 12, kernel=_cabi.KERNEL_PERSISTENT)          # not whole words
    with pytest.raises(_cabi.PymcError, match="persistent"):
        _cabi.Lattices(_cabi.MODEL_ISING, 32768, 32768, kernel=_cabi.KERNEL_PERSISTENT)   # 128 MiB of bits
