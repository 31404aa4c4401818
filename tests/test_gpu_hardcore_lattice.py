"""BASELINE config 5 read as a lattice model (note on C5): the nearest-neighbour-exclusion
("hard-core", hard-square) lattice gas is the reference's ``latticegas`` (examples/latgas.py:7-57) with a pair
energy whose Boltzmann factor underflows to 0.0.  The bit-exact trajectory check against the reference is the
``lg_hardcore`` case of tests/golden/latgas.npz (run by tests/test_gpu_latgas.py); here: physics with stated
error bars and temperature replica exchange over the activity ladder."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _violations(occ):
    occ = np.asarray(occ).astype(np.int64)
    return int((occ * (np.roll(occ, 1, 0) + np.roll(occ, 1, 1))).sum())


def exact_particle_number(n, m, z):
    """<N> of the nearest-neighbour-exclusion gas on the periodic n x m lattice at activity z, from the
    transfer matrix over columns of n sites: T[a,b] = [a & b == 0] z^|b| on ring-compatible states;
    symmetrised as D^1/2 A D^1/2, Z = sum_i lambda_i^m, <N> = z dlnZ/dz (central difference)."""
    full = (1 << n) - 1
    states = np.array([s for s in range(1 << n) if not (s & (((s << 1) | (s >> (n - 1))) & full))])
    A = ((states[:, None] & states[None, :]) == 0).astype(np.float64)
    pop = np.array([bin(s).count("1") for s in states], dtype=np.float64)

    def lnZ(zz):
        d = zz ** (0.5 * pop)
        lam = np.linalg.eigvalsh(A * d[:, None] * d[None, :])
        top = np.abs(lam).max()
        return m * np.log(top) + np.log(np.sum((lam / top) ** m))

    h = 1e-5
    return (lnZ(z * (1 + h)) - lnZ(z * (1 - h))) / (2 * h)


def test_transfer_matrix_helper_against_enumeration():
    """The checker itself: 4 x 4 by brute force."""
    z, n = 1.7, 4
    tot = num = 0.0
    for s in range(1 << 16):
        occ = np.array([(s >> k) & 1 for k in range(16)]).reshape(n, n)
        if (occ * (np.roll(occ, 1, 0) + np.roll(occ, 1, 1))).sum() == 0:
            w = z ** occ.sum()
            tot += w
            num += w * occ.sum()
    assert exact_particle_number(4, 4, z) == pytest.approx(num / tot, rel=1e-7)


@pytest.mark.parametrize("shape,layout,mu", [((8, 8), "int8", 0.5), ((16, 128), "int8", 0.5), ((16, 128), "bits", -0.7)])
def test_density_matches_exact_transfer_matrix(shape, layout, mu):
    """Independent streams against the exact finite-lattice average, error bar from the replica spread."""
    from py_mc_b200 import _cabi
    from py_mc_b200.latgas import hardcore_latticegas
    lenX, lenY = shape
    kT, R, nsamp = 1.0, 64, 1500
    model = hardcore_latticegas(mu=mu)
    lat = _cabi.Lattices(_cabi.MODEL_LATGAS, lenX, lenY, nreplicas=R, ntemps=1, seed=2024,
                         layout=_cabi.LAYOUT_BITS if layout == "bits" else _cabi.LAYOUT_INT8)
    for r in range(R):
        lat.upload(r, np.zeros((lenX, lenY), dtype=np.int8))   # start on the exclusion manifold
    lat.set_accept_tables(model._pmc_thresholds(kT)[None, :])
    lat.sweep(300)
    acc = np.zeros(R)
    for _ in range(nsamp):
        lat.sweep(1)
        b, n = lat.tracked()
        assert not b.any()
        acc += n
    lat.close()
    per_replica = acc / nsamp
    mean, err = per_replica.mean(), per_replica.std(ddof=1) / np.sqrt(R)
    exact = exact_particle_number(lenX, lenY, np.exp(mu / kT))
    assert abs(mean - exact) < 5 * err, (mean, exact, err)
    assert err < 0.01 * exact


def test_random_start_relaxes_onto_exclusion_manifold():
    from py_mc_b200 import _cabi
    from py_mc_b200.latgas import hardcore_latticegas
    model = hardcore_latticegas(mu=0.5)
    lat = _cabi.Lattices(_cabi.MODEL_LATGAS, 64, 256, nreplicas=3, ntemps=1, seed=5)
    lat.init_random()
    lat.set_accept_tables(model._pmc_thresholds(1.0)[None, :])
    b0, _ = lat.reduce_observables()
    assert np.all(b0 > 0)
    lat.sweep(4)
    b, n = lat.reduce_observables()
    assert np.all(b == 0) and np.all(n > 0)
    lat.sweep(50)
    b, n = lat.tracked()
    assert np.all(b == 0)
    for r in range(3):
        assert _violations(lat.download(r)) == 0
    lat.close()


def test_replica_exchange_over_activity_ladder(tmp_path):
    """TemperatureRX_MPI (mc_mpi.py:76-209) on the hard-core lattice gas: E = -mu*N on the manifold, so the kT
    ladder is an activity ladder; the density must fall with kT and agree with a single-temperature run."""
    import py_mc_b200
    from py_mc_b200.latgas import hardcore_latticegas, latgas_config, observer
    from py_mc_b200.mc import CanonicalMonteCarlo
    from py_mc_b200.mc_mpi import TemperatureRX_MPI
    old = os.getcwd()
    os.chdir(tmp_path)
    try:
        py_mc_b200.set_seed(123)
        R, L, mu = 8, 32, 1.0
        kTs = np.linspace(1.0, 2.4, R)            # z = e^(1/kT) from 2.72 down to 1.52, all below z_c = 3.796
        model = hardcore_latticegas(mu=mu)
        configs = []
        for _ in range(R):
            c = latgas_config(L, L)
            c.config[:, :] = 0
            configs.append(c)
        rx = TemperatureRX_MPI(None, CanonicalMonteCarlo, model, configs, kTs)
        rx.run(400, 2)
        out = rx.run(4000, 2, 2, observer=observer())
        assert rx.accept_count > 0
        e = np.array([o[0] for o in out])
        n = np.array([o[1] for o in out])
        assert np.allclose(e, -mu * n, rtol=1e-12)            # no occupied pair ever contributed HARD_J
        assert np.all(np.diff(n) < 0)                          # lower kT = higher activity = denser
        for c in configs:
            assert _violations(c.config) == 0
        # one temperature of the ladder against an independent canonical run
        i = 3
        py_mc_b200.set_seed(456)
        cfg = latgas_config(L, L)
        cfg.config[:, :] = 0
        calc = CanonicalMonteCarlo(model, kTs[i], cfg)
        calc.run(400)
        single = calc.run(4000, 2, observer())
        assert abs(single[1] - n[i]) / (L * L) < 4e-3, (single[1], n[i])
    finally:
        os.chdir(old)
