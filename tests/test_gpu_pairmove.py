"""GPU: nearest-neighbour particle-hole exchange kernel (K6) against its bit-exact CPU specification
(oracle/pairmove.py), conservation laws, and exact enumeration of a small canonical ensemble.
The reference's own particle move (examples/latgas_move.py:56-83) is non-local and its dE is defective
(SURVEY.md 2.3), so there is no reference trajectory to compare with for this kernel."""
import itertools

import numpy as np
import pytest

from oracle import latgas as o_latgas
from oracle import pairmove as o_pair

pytestmark = pytest.mark.gpu


def _layouts(lenY):
    return [0, 1] if lenY % 128 == 0 else [0]


@pytest.mark.parametrize("shape,J,kT", [((8, 12), -1.0, 1.0), ((4, 4), 1.0, 0.7), ((12, 132), -0.5, 0.6),
                                        ((8, 128), -1.0, 0.8), ((16, 256), 2.0, 1.5), ((4, 260), -1.0, 1.2)])
def test_pair_sweeps_match_specification(shape, J, kT):
    from py_mc_b200 import _cabi
    seed = 909
    thr = o_pair.threshold_table(J, kT)
    ref = o_latgas.init_config(seed, 0, *shape)
    N0, B = int(ref.sum()), o_latgas.bond_sum(ref)
    lats = []
    for layout in _layouts(shape[1]):
        lat = _cabi.Lattices(_cabi.MODEL_LATGAS, shape[0], shape[1], seed=seed, layout=layout)
        lat.init_random()
        lat.set_pair_tables(thr[None, :])
        lats.append(lat)
    for sw in range(4):
        ntry, nacc, dB = o_pair.sweep_classes(ref, thr, seed, 0, sw)
        B += dB
        assert nacc > 0
        for lat in lats:
            lat.pair_sweep(1)
            assert np.array_equal(lat.download(0), ref), (shape, lat.layout, sw)
            b, n = lat.tracked()
            assert (int(b[0]), int(n[0])) == (B, N0)          # tracked bond count, conserved particle number
            b2, n2 = lat.reduce_observables()
            assert (int(b2[0]), int(n2[0])) == (B, N0)
    assert lats[0].pair_counter == 4


def test_class_mask_runs_a_subset_of_classes():
    from py_mc_b200 import _cabi
    shape, seed = (8, 16), 4
    thr = o_pair.threshold_table(-1.0, 1.0)
    ref = o_latgas.init_config(seed, 0, *shape)
    lat = _cabi.Lattices(_cabi.MODEL_LATGAS, shape[0], shape[1], seed=seed)
    lat.init_random()
    lat.set_pair_tables(thr[None, :])
    lat.pair_sweep(1, 0x00FF)  # horizontal bonds only
    o_pair.sweep_classes(ref, thr, seed, 0, 0, classes=range(8))
    assert np.array_equal(lat.download(0), ref)
    lat.pair_sweep(1, 0xFF00)  # vertical bonds, next sweep index
    o_pair.sweep_classes(ref, thr, seed, 0, 1, classes=range(8, 16))
    assert np.array_equal(lat.download(0), ref)


def test_shapes_and_tables_are_validated():
    from py_mc_b200 import _cabi
    lat = _cabi.Lattices(_cabi.MODEL_LATGAS, 6, 8)
    with pytest.raises(_cabi.PymcError, match="% 4"):
        lat.set_pair_tables(o_pair.threshold_table(-1.0, 1.0)[None, :])
    lat = _cabi.Lattices(_cabi.MODEL_LATGAS, 8, 8)
    with pytest.raises(_cabi.PymcError, match="pymc_set_pair_tables first"):
        lat.pair_sweep(1)
    bad = np.full((1, 7), 0xFFFFFFFF, dtype=np.uint32)
    bad[0, 1] = bad[0, 5] = 7
    with pytest.raises(_cabi.PymcError, match="both sides"):
        lat.set_pair_tables(bad)
    ising = _cabi.Lattices(_cabi.MODEL_ISING, 8, 8)
    with pytest.raises(_cabi.PymcError, match="lattice-gas"):
        ising.set_pair_tables(o_pair.threshold_table(-1.0, 1.0)[None, :])


def test_canonical_average_matches_exact_enumeration():
    """4x4 lattice, 8 particles, J = -1: the chain must sample exp(-E/kT) over the C(16,8) = 12870
    configurations of fixed particle number (detailed balance of the exchange move)."""
    from py_mc_b200 import _cabi
    J, kT, L, N = -1.0, 1.2, 4, 8
    sites = list(itertools.combinations(range(L * L), N))
    Bs = np.empty(len(sites))
    for n, occ in enumerate(sites):
        c = np.zeros(L * L, dtype=np.int64)
        c[list(occ)] = 1
        c = c.reshape(L, L)
        Bs[n] = (c * (np.roll(c, 1, 0) + np.roll(c, 1, 1))).sum()
    w = np.exp(-J * Bs / kT)
    exact = float((Bs * w).sum() / w.sum())
    R = 256  # independent replicas started from the same configuration
    lat = _cabi.Lattices(_cabi.MODEL_LATGAS, L, L, nreplicas=R, ntemps=1, seed=2024)
    start = np.zeros((L, L), dtype=np.int8)
    start.ravel()[:N] = 1
    for r in range(R):
        lat.upload(r, start)
    lat.set_pair_tables(o_pair.threshold_table(J, kT)[None, :])
    lat.pair_sweep(200)
    samples = []
    for _ in range(200):
        lat.pair_sweep(5)
        b, n = lat.tracked()
        assert np.all(n == N)
        samples.append(b.astype(np.float64))
    s = np.array(samples)          # [time, replica]
    per_rep = s.mean(axis=0)       # replicas are independent: their spread gives the error bar
    mean, err = per_rep.mean(), per_rep.std(ddof=1) / np.sqrt(R)
    assert abs(mean - exact) < 5 * err + 1e-3, (mean, exact, err)
