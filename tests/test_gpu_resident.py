"""GPU: the whole-lattice-resident kernel (K10, pmc_resident.cuh) against the CPU specification and against
the per-pass kernels: flip sweeps of both models, pair-exchange passes, the mixed step, several steps per
launch, both layouts, per-replica temperatures, and the automatic choice for batches of small lattices."""
import numpy as np
import pytest

from oracle import ising as o_ising
from oracle import latgas as o_latgas
from oracle import pairmove as o_pair

pytestmark = pytest.mark.gpu

SHAPES = [(4, 128), (6, 128), (16, 384), (32, 256)]


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("layout", [0, 1])
def test_ising_flip_sweeps_match_specification(shape, layout):
    from py_mc_b200 import _cabi
    seed, R, J = 31, 3, -1.0
    kTs = [1.5, 2.27, 4.0]
    lat = _cabi.Lattices(_cabi.MODEL_ISING, shape[0], shape[1], nreplicas=R, ntemps=R, seed=seed, layout=layout,
                         kernel=_cabi.KERNEL_RESIDENT)
    assert lat.sweep_kernel_info()[1] == "resident_kernel"
    lat.init_random()
    from py_mc_b200.ising2D import ising2D
    lat.set_accept_tables(np.array([ising2D(J)._pmc_thresholds(kT) for kT in kTs]))
    lat.set_replica_to_T(np.array([2, 0, 1], dtype=np.int32))
    refs = [o_ising.init_config(seed, r, *shape) for r in range(R)]
    done = 0
    for chunk in (1, 3, 2):                       # one launch per call, whatever the number of sweeps
        lat.sweep(chunk)
        for r, t in enumerate((2, 0, 1)):
            for s in range(done, done + chunk):
                o_ising.sweep(refs[r], J, kTs[t], seed, r, s)
            assert np.array_equal(lat.download(r), refs[r]), (shape, layout, r, done)
        done += chunk
        b, m = lat.tracked()
        assert [int(x) for x in b] == [o_ising.bond_sum(c) for c in refs]
        assert [int(x) for x in m] == [int(c.sum()) for c in refs]
    assert lat.sweep_counter == 6


@pytest.mark.parametrize("shape", [(4, 128), (8, 256), (16, 384)])
@pytest.mark.parametrize("layout", [0, 1])
def test_latgas_flip_pair_and_mixed_steps_match_specification(shape, layout):
    from py_mc_b200 import _cabi
    seed, J, Eads, mu, kT = 77, -1.0, -0.3, 0.2, 0.9
    thr = o_latgas.threshold_table(J, Eads, mu, kT)
    pthr = o_pair.threshold_table(J, kT)
    lat = _cabi.Lattices(_cabi.MODEL_LATGAS, shape[0], shape[1], nreplicas=2, seed=seed, layout=layout,
                         kernel=_cabi.KERNEL_RESIDENT)
    lat.init_random()
    lat.set_accept_tables(thr[None, :])
    lat.set_pair_tables(pthr[None, :])
    refs = [o_latgas.init_config(seed, r, *shape) for r in range(2)]
    sweep = pair = 0

    def check(tag):
        for r in range(2):
            assert np.array_equal(lat.download(r), refs[r]), (shape, layout, tag, r)
        b, n = lat.tracked()
        assert [int(x) for x in b] == [o_latgas.bond_sum(c) for c in refs], tag
        assert [int(x) for x in n] == [int(c.sum()) for c in refs], tag
        assert (lat.sweep_counter, lat.pair_counter) == (sweep, pair)

    lat.sweep(2)
    for r in range(2):
        for s in range(2):
            o_latgas.sweep(refs[r], thr, seed, r, sweep + s)
    sweep += 2
    check("flip")

    lat.pair_sweep(2)
    for r in range(2):
        for s in range(2):
            o_pair.sweep_classes(refs[r], pthr, seed, r, pair + s)
    pair += 2
    check("pair")

    lat.pair_sweep(1, 0x0F0F)
    for r in range(2):
        o_pair.sweep_classes(refs[r], pthr, seed, r, pair, classes=[0, 1, 2, 3, 8, 9, 10, 11])
    pair += 1
    check("pair subset")

    lat.mixed_sweep(3, 0x00FF, 0xFF00)            # pair counter is odd here: vertical, horizontal, vertical
    for r in range(2):
        for s in range(3):
            o_latgas.sweep(refs[r], thr, seed, r, sweep + s)
            classes = range(8) if (pair + s) % 2 == 0 else range(8, 16)
            o_pair.sweep_classes(refs[r], pthr, seed, r, pair + s, classes=classes)
    sweep += 3
    pair += 3
    check("mixed")


def test_mixed_sweep_on_per_pass_kernels_equals_separate_calls():
    from py_mc_b200 import _cabi
    seed, shape = 5, (8, 12)                      # not a whole word: the per-pass kernels
    thr = o_latgas.threshold_table(1.0, -0.5, 0.0, 0.8)
    pthr = o_pair.threshold_table(1.0, 0.8)
    lats = []
    for _ in range(2):
        lat = _cabi.Lattices(_cabi.MODEL_LATGAS, shape[0], shape[1], seed=seed)
        lat.init_random()
        lat.set_accept_tables(thr[None, :])
        lat.set_pair_tables(pthr[None, :])
        lats.append(lat)
    assert lats[0].sweep_kernel_info()[1] != "resident_kernel"
    lats[0].mixed_sweep(4, 0x00FF, 0xFF00)
    for s in range(4):
        lats[1].sweep(1)
        lats[1].pair_sweep(1, 0x00FF if s % 2 == 0 else 0xFF00)
    assert np.array_equal(lats[0].download(0), lats[1].download(0))
    assert lats[0].pair_counter == lats[1].pair_counter == 4


def test_auto_takes_the_resident_kernel_for_a_batch_of_small_lattices():
    from py_mc_b200 import _cabi
    seed, shape, R = 8, (32, 128), 192            # more replicas than SMs, 4 KiB of bits each
    from py_mc_b200.ising2D import ising2D
    thr = ising2D(-1.0)._pmc_thresholds(2.3)
    out = {}
    for kernel in (_cabi.KERNEL_AUTO, _cabi.KERNEL_GENERIC):
        lat = _cabi.Lattices(_cabi.MODEL_ISING, shape[0], shape[1], nreplicas=R, seed=seed, kernel=kernel)
        lat.init_random()
        lat.set_accept_tables(thr[None, :])
        lat.sweep(7)
        out[kernel] = (lat.sweep_kernel_info()[1], [lat.download(r) for r in (0, 95, 191)], lat.tracked())
    assert out[_cabi.KERNEL_AUTO][0] == "resident_kernel" and out[_cabi.KERNEL_GENERIC][0] == "sweep_word_kernel"
    for a, b in zip(out[_cabi.KERNEL_AUTO][1], out[_cabi.KERNEL_GENERIC][1]):
        assert np.array_equal(a, b)
    assert np.array_equal(out[_cabi.KERNEL_AUTO][2][0], out[_cabi.KERNEL_GENERIC][2][0])
    few = _cabi.Lattices(_cabi.MODEL_ISING, shape[0], shape[1], nreplicas=4, seed=seed)
    assert few.sweep_kernel_info()[1] != "resident_kernel"    # too few replicas to fill the GPU that way


@pytest.mark.parametrize("layout", [0, 1])
def test_large_image_needs_the_raised_shared_memory_limit(layout):
    """1024 x 1024 = 128 KiB of bits per replica: forced, against the fused / row kernels."""
    from py_mc_b200 import _cabi
    seed, L = 12, 1024
    from py_mc_b200.ising2D import ising2D
    thr = ising2D(-1.0)._pmc_thresholds(2.269)
    got = []
    for kernel in (_cabi.KERNEL_RESIDENT, _cabi.KERNEL_AUTO):
        lat = _cabi.Lattices(_cabi.MODEL_ISING, L, L, nreplicas=2, seed=seed, layout=layout, kernel=kernel)
        lat.init_random()
        lat.set_accept_tables(thr[None, :])
        lat.sweep(3)
        got.append((lat.download(1), lat.tracked()))
        lat.close()
    assert np.array_equal(got[0][0], got[1][0])
    assert np.array_equal(got[0][1][0], got[1][1][0]) and np.array_equal(got[0][1][1], got[1][1][1])


def test_resident_is_refused_where_it_cannot_run():
    from py_mc_b200 import _cabi
    with pytest.raises(_cabi.PymcError, match="resident"):
        _cabi.Lattices(_cabi.MODEL_ISING, 8, 12, kernel=_cabi.KERNEL_RESIDENT)        # not whole words
    with pytest.raises(_cabi.PymcError, match="resident"):
        _cabi.Lattices(_cabi.MODEL_ISING, 2048, 2048, kernel=_cabi.KERNEL_RESIDENT)   # 512 KiB of bits


@pytest.mark.parametrize("which", ["pbatch_kernel", "resident_kernel"])
def test_auto_on_large_int8_batches(which, monkeypatch):
    """256 replicas of 1024 x 1024 int8 (BASELINE config 3 on one GPU): calls of one sweep take the fused kernel,
    longer calls a shared-memory-resident one -- the batch-resident kernel K12, or, with PYMC_NO_PERSISTENT=1 (no
    cooperative launches wanted), one block per replica; the trajectory is that of the two-pass kernel either way."""
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D
    seed, L, R, probe = 3, 1024, 256, 200
    thr = ising2D(-1.0)._pmc_thresholds(2.269)
    if which == "resident_kernel":
        monkeypatch.setenv("PYMC_NO_PERSISTENT", "1")
    big = _cabi.Lattices(_cabi.MODEL_ISING, L, L, nreplicas=R, seed=seed)
    assert big.sweep_kernel_info()[1] == which
    one = _cabi.Lattices(_cabi.MODEL_ISING, L, L, nreplicas=1, seed=seed, replica_offset=probe, nreplicas_global=R,
                         kernel=_cabi.KERNEL_TWO_PASS)
    for lat in (big, one):
        lat.init_random()
        lat.set_accept_tables(thr[None, :])
    big.sweep(1)
    big.sweep(3)
    big.sweep(1)
    one.sweep(5)
    assert np.array_equal(big.download(probe), one.download(0))
    b, m = big.tracked()
    b1, m1 = one.tracked()
    assert (int(b[probe]), int(m[probe])) == (int(b1[0]), int(m1[0]))
    b2, m2 = big.reduce_observables()
    assert np.array_equal(b, b2) and np.array_equal(m, m2)
    big.close()
    one.close()
