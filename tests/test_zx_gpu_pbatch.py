"""GPU: the batch-resident kernel (K12, pmc_pbatch.cuh) -- the rows of all replicas of a handle cut into one piece
per SM, resident in shared memory for all sweeps and (on one GPU) all replica exchanges of a call -- against the
per-sweep kernels: same trajectories, same exact sums, same exchange decisions, maps, energies and counters.

The per-sweep kernels are pinned to the reference by tests/test_gpu_ising.py, test_gpu_latgas.py and test_gpu_rx.py
(fixtures recorded from the unmodified reference); this file extends that pin to K12 by comparing the two on the
same seeds, at the sizes of BASELINE configs 2-4 and on shapes that exercise every case of the block geometry."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TC = 2.269185314213022


def _ising_thr(kTs):
    from py_mc_b200.ising2D import ising2D
    m = ising2D(-1.0)
    return np.array([m._pmc_thresholds(kT) for kT in kTs])


def _latgas_thr(kTs):
    from py_mc_b200.latgas import latticegas
    m = latticegas(Eads=-0.3, J=-1.0, mu=0.2)
    return np.array([m._pmc_thresholds(kT) for kT in kTs])


def _run(model, shape, layout, thr, kernel, nrep, calls, seed=5):
    """Every replica at its own temperature (row r of thr), no exchange."""
    from py_mc_b200 import _cabi
    lat = _cabi.Lattices(model, shape[0], shape[1], nreplicas=nrep, ntemps=nrep, seed=seed, layout=layout, kernel=kernel)
    lat.init_random()
    lat.set_accept_tables(thr)
    for n in calls:
        lat.sweep(n)
    lat.synchronize()
    out = ([lat.download(r) for r in range(nrep)], lat.tracked(), lat.reduce_observables(), lat.sweep_kernel_info()[1],
           lat.sweep_counter)
    lat.close()
    return out


def _same(a, b):
    assert a[3] == "pbatch_kernel" and b[3] != "pbatch_kernel", (a[3], b[3])
    for r, (x, y) in enumerate(zip(a[0], b[0])):
        assert np.array_equal(x, y), r
    for k in (0, 1):
        assert np.array_equal(a[1][k], b[1][k]) and np.array_equal(a[1][k], a[2][k]) and np.array_equal(b[1][k], b[2][k])
    assert a[4] == b[4]


@pytest.mark.parametrize("nrep,shape", [(1, (300, 256)), (1, (64, 128)), (3, (200, 256)), (5, (100, 384)), (148, (4, 128)),
                                        (7, (1024, 1024)), (300, (8, 256))])
@pytest.mark.parametrize("layout", [0, 1])
def test_ising_matches_two_pass_kernel(nrep, shape, layout):
    from py_mc_b200 import _cabi
    thr = _ising_thr(np.linspace(1.8, 3.0, nrep))
    a = _run(_cabi.MODEL_ISING, shape, layout, thr, _cabi.KERNEL_PBATCH, nrep, (2, 3))
    b = _run(_cabi.MODEL_ISING, shape, layout, thr, _cabi.KERNEL_TWO_PASS, nrep, (5,))
    _same(a, b)


@pytest.mark.parametrize("layout", [0, 1])
def test_lattice_gas_batch_of_config_4(layout):
    """512 lattice-gas replicas of 256 x 256 (BASELINE config 4): 3.46 replicas per block."""
    from py_mc_b200 import _cabi
    nrep, shape = 512, (256, 256)
    thr = _latgas_thr(np.linspace(0.7, 1.4, nrep))
    a = _run(_cabi.MODEL_LATGAS, shape, layout, thr, _cabi.KERNEL_PBATCH, nrep, (3, 2))
    b = _run(_cabi.MODEL_LATGAS, shape, layout, thr, _cabi.KERNEL_TWO_PASS, nrep, (5,))
    _same(a, b)


@pytest.mark.parametrize("nrep", [8, 32, 256])
def test_replica_exchange_batches_of_config_3(nrep):
    """R x 1024^2 int8 (BASELINE config 3 on 8 / 32 / 256 replicas per GPU): the automatic choice, and the same
    trajectory as the per-colour kernel."""
    from py_mc_b200 import _cabi
    thr = _ising_thr(np.linspace(2.0, 2.6, nrep))
    a = _run(_cabi.MODEL_ISING, (1024, 1024), 0, thr, _cabi.KERNEL_AUTO, nrep, (4, 2))
    b = _run(_cabi.MODEL_ISING, (1024, 1024), 0, thr, _cabi.KERNEL_TWO_PASS, nrep, (6,))
    _same(a, b)


def _rx(shape, layout, kernel, R, ncycles, every, first_scheme, model_code, thr, coeffs, kTs, seed=11):
    from py_mc_b200 import _cabi
    lat = _cabi.Lattices(model_code, shape[0], shape[1], nreplicas=R, ntemps=R, seed=seed, layout=layout, kernel=kernel)
    lat.init_random()
    lat.set_accept_tables(thr)
    lat.rx_setup(1.0 / np.asarray(kTs), *coeffs)
    scheme = first_scheme
    for n in ncycles:  # several calls: the maps, the round counter and the scheme carry over
        lat.rx_cycle(n, every, scheme)
        scheme ^= n & 1
    lat.synchronize()
    out = dict(configs=[lat.download(r) for r in range(R)], map=lat.get_replica_to_T(), energies=lat.rx_get_energies(),
               stats=lat.rx_stats(), round=lat.rx_round, tracked=lat.tracked(), reduced=lat.reduce_observables(),
               counter=lat.sweep_counter, kernel=lat.sweep_kernel_info()[1])
    lat.close()
    return out


@pytest.mark.parametrize("R,shape,every,ncycles", [(6, (64, 128), 1, (5, 4)), (16, (128, 256), 3, (3, 2)), (40, (32, 128), 2, (7,)),
                                                   (5, (200, 384), 2, (4, 3))])
@pytest.mark.parametrize("layout", [0, 1])
def test_exchange_inside_the_kernel_matches_the_decision_kernel(R, shape, every, ncycles, layout):
    """pymc_rx_cycle on a K12 handle is ONE launch (sweeps, exact energies, grid barrier, pair decisions); on a
    two-pass handle it is the loop [sweeps, rx_energy_kernel, rx_decide_kernel].  Everything a caller can observe
    must agree: configurations, replica -> temperature map, energies of the last exchange, tried / accepted per
    pair, round and sweep counters, exact sums."""
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D
    kTs = np.linspace(2.25, 2.25 + 0.004 * R, R)  # close enough for exchanges to be accepted on every lattice size here
    model = ising2D(-1.0)
    thr = _ising_thr(kTs)
    a = _rx(shape, layout, _cabi.KERNEL_PBATCH, R, ncycles, every, 1, _cabi.MODEL_ISING, thr, model._pmc_energy_coeffs(), kTs)
    b = _rx(shape, layout, _cabi.KERNEL_TWO_PASS, R, ncycles, every, 1, _cabi.MODEL_ISING, thr, model._pmc_energy_coeffs(), kTs)
    assert a["kernel"] == "pbatch_kernel" and b["kernel"] != "pbatch_kernel"
    assert np.array_equal(a["map"], b["map"]) and sorted(a["map"]) == list(range(R))
    assert int(a["stats"][1].sum()) > 0, "no exchange was accepted: the case tests nothing"
    for k in (0, 1):
        assert np.array_equal(a["stats"][k], b["stats"][k])
        assert np.array_equal(a["tracked"][k], b["tracked"][k]) and np.array_equal(a["tracked"][k], a["reduced"][k])
    assert np.array_equal(a["energies"], b["energies"])
    assert (a["round"], a["counter"]) == (b["round"], b["counter"])
    for r in range(R):
        assert np.array_equal(a["configs"][r], b["configs"][r]), r


def test_exchange_of_the_hard_core_lattice_gas():
    """Config 5 read as a lattice model: nearest-neighbour exclusion, energy has a particle-number term."""
    from py_mc_b200 import _cabi
    from py_mc_b200.latgas import hardcore_latticegas
    R, shape = 8, (128, 256)
    kTs = np.linspace(0.5, 2.0, R)
    model = hardcore_latticegas(mu=1.0)
    thr = np.array([model._pmc_thresholds(kT) for kT in kTs])
    a = _rx(shape, 0, _cabi.KERNEL_PBATCH, R, (6,), 2, 0, _cabi.MODEL_LATGAS, thr, model._pmc_energy_coeffs(), kTs)
    b = _rx(shape, 0, _cabi.KERNEL_TWO_PASS, R, (6,), 2, 0, _cabi.MODEL_LATGAS, thr, model._pmc_energy_coeffs(), kTs)
    assert np.array_equal(a["map"], b["map"]) and np.array_equal(a["energies"], b["energies"])
    for k in (0, 1):
        assert np.array_equal(a["stats"][k], b["stats"][k]) and np.array_equal(a["tracked"][k], b["tracked"][k])
    for r in range(R):
        assert np.array_equal(a["configs"][r], b["configs"][r])


def test_headline_lattice_against_the_fused_kernel():
    """16384 x 16384 int8: PYMC_KERNEL_AUTO takes K12 for calls of two sweeps or more and the fused kernel for
    single sweeps, on the same buffers; trajectory and sums equal those of the two-pass kernel."""
    from py_mc_b200 import _cabi
    thr = _ising_thr([TC])
    L = 16384
    a = _cabi.Lattices(_cabi.MODEL_ISING, L, L, seed=20260611)
    b = _cabi.Lattices(_cabi.MODEL_ISING, L, L, seed=20260611, kernel=_cabi.KERNEL_TWO_PASS)
    assert a.sweep_kernel_info()[1] == "pbatch_kernel"
    for lat in (a, b):
        lat.init_random()
        lat.set_accept_tables(thr)
    a.sweep(5)
    a.sweep(1)
    a.sweep(2)
    b.sweep(8)
    a.synchronize()
    assert np.array_equal(a.download(0), b.download(0))
    ta, tb, ra = a.tracked(), b.tracked(), a.reduce_observables()
    for k in (0, 1):
        assert int(ta[k][0]) == int(tb[k][0]) == int(ra[k][0])
    a.close()
    b.close()


def test_refused_where_it_cannot_run():
    from py_mc_b200 import _cabi
    with pytest.raises(_cabi.PymcError, match="batch-resident"):
        _cabi.Lattices(_cabi.MODEL_ISING, 8, 12, kernel=_cabi.KERNEL_PBATCH)                       # not whole words
    with pytest.raises(_cabi.PymcError, match="batch-resident"):
        _cabi.Lattices(_cabi.MODEL_ISING, 1024, 1024, nreplicas=512, kernel=_cabi.KERNEL_PBATCH)   # 64 MiB of bits
    with pytest.raises(_cabi.PymcError, match="batch-resident"):
        _cabi.Lattices(_cabi.MODEL_ISING, 2, 128, nreplicas=4096, kernel=_cabi.KERNEL_PBATCH)      # > 8 replicas per block
