"""GPU: the row-streaming int8 sweep kernel (K13, pmc_rowstream.cuh) -- rows of a piece stream through shared memory by
bulk asynchronous copies, both colours in one pass, one launch per sweep -- against the oracle and against the
two-pass kernels (which tests/test_gpu_ising.py and test_gpu_latgas.py pin to the reference's trajectories): same
configurations after every sweep, same tracked sums, for one piece and several, ragged pieces, batches of replicas at
different temperatures, both models."""
import os

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


def _run(model, shape, thr, kernel, nrep, nsweeps, seed=11, pieces=None):
    from py_mc_b200 import _cabi
    if pieces:
        os.environ["PYMC_ROWSTREAM_PIECES"] = str(pieces)
    try:
        lat = _cabi.Lattices(model, shape[0], shape[1], nreplicas=nrep, ntemps=nrep, seed=seed, kernel=kernel)
    finally:
        os.environ.pop("PYMC_ROWSTREAM_PIECES", None)
    lat.init_random()
    lat.set_accept_tables(thr)
    snaps = []
    for _ in range(nsweeps):
        lat.sweep(1)
        snaps.append([lat.download(r) for r in range(nrep)])
    out = (snaps, lat.tracked(), lat.reduce_observables(), lat.sweep_kernel_info()[1], lat.sweep_counter)
    lat.close()
    return out


@pytest.mark.parametrize("nrep,shape,pieces", [(1, (260, 1024), 1), (1, (526, 1024), 2), (3, (270, 2048), 2), (1, (60, 16384), 3),
                                               (2, (70, 4096), 1), (1, (74, 8192), 2), (1, (700, 2048), None)])
@pytest.mark.parametrize("model", [0, 1])
def test_matches_two_pass_kernel_after_every_sweep(nrep, shape, pieces, model):
    from py_mc_b200 import _cabi
    thr = (_ising_thr if model == 0 else _latgas_thr)(np.linspace(1.9, 2.9, nrep))
    a = _run(model, shape, thr, _cabi.KERNEL_ROWSTREAM, nrep, 4, pieces=pieces)
    b = _run(model, shape, thr, _cabi.KERNEL_TWO_PASS, nrep, 4)
    assert a[3] == "rowstream_kernel" and b[3] != "rowstream_kernel", (a[3], b[3])
    for sw, (x, y) in enumerate(zip(a[0], b[0])):
        for r in range(nrep):
            assert np.array_equal(x[r], y[r]), (sw, r)
    for k in (0, 1):
        assert np.array_equal(a[1][k], b[1][k]) and np.array_equal(a[1][k], a[2][k])
    assert a[4] == b[4] == 4


def test_against_the_oracle():
    from py_mc_b200 import _cabi
    from oracle import ising as o_ising
    seed, shape = 20260611, (260, 1024)
    thr = o_ising.threshold_table(-1.0, TC)
    lat = _cabi.Lattices(_cabi.MODEL_ISING, shape[0], shape[1], seed=seed, kernel=_cabi.KERNEL_ROWSTREAM)
    lat.init_random()
    lat.set_accept_tables(thr[1][None, :])
    ref = o_ising.init_config(seed, 0, *shape)
    B = o_ising.bond_sum(ref)
    for sw in range(3):
        lat.sweep(1)
        _, dB, _ = o_ising.sweep(ref, -1.0, TC, seed, 0, sw, thr)
        B += dB
        assert np.array_equal(lat.download(0), ref), sw
    b, m = lat.tracked()
    assert int(b[0]) == B and int(m[0]) == int(ref.astype(np.int64).sum())
    lat.close()


def test_auto_takes_it_for_single_sweeps_of_large_lattices_and_calls_of_many_sweeps_agree():
    """16384 x 8192: PYMC_KERNEL_AUTO runs single-sweep calls on K13 and calls of several sweeps on the batch-resident
    kernel; both must walk the same trajectory, and so must the tiled kernel (PYMC_NO_ROWSTREAM=1)."""
    from py_mc_b200 import _cabi
    thr = _ising_thr([TC])
    outs = []
    for calls, env in (((1, 1, 1, 1), {}), ((4,), {}), ((2, 1, 1), {}), ((1, 1, 1, 1), {"PYMC_NO_ROWSTREAM": "1"})):
        os.environ.update(env)
        try:
            lat = _cabi.Lattices(_cabi.MODEL_ISING, 16384, 8192, seed=3)
        finally:
            for k in env:
                os.environ.pop(k, None)
        lat.init_random()
        lat.set_accept_tables(thr)
        for n in calls:
            lat.sweep(n)
        outs.append((lat.download(0), lat.tracked()))
        lat.close()
    for o in outs[1:]:
        assert np.array_equal(o[0], outs[0][0])
        assert np.array_equal(o[1][0], outs[0][1][0]) and np.array_equal(o[1][1], outs[0][1][1])


@pytest.mark.parametrize("shape", [(64, 384), (8, 16384), (64, 256)])
def test_unsupported_shapes_are_refused(shape):
    from py_mc_b200 import _cabi
    with pytest.raises(_cabi.PymcError) as info:
        _cabi.Lattices(_cabi.MODEL_ISING, shape[0], shape[1], kernel=_cabi.KERNEL_ROWSTREAM)
    assert info.value.code == -3
