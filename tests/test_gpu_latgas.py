"""GPU parity for the lattice-gas flip path (examples/latgas.py) through the C ABI and the drop-in API."""
import os

import numpy as np
import pytest

from oracle import latgas as o_latgas

pytestmark = pytest.mark.gpu


def _layouts(lenY):
    return [0, 1] if lenY % 128 == 0 else [0]


def test_trajectories_and_observables_match_reference(golden_latgas):
    from py_mc_b200 import _cabi
    from py_mc_b200.latgas import latticegas
    z, meta = golden_latgas
    for m in meta:
        n = m["name"]
        model = latticegas(Eads=m["Eads"], J=m["J"], mu=m["mu"])
        thr = model._pmc_thresholds(m["kT"])
        assert np.array_equal(thr, o_latgas.threshold_table(m["J"], m["Eads"], m["mu"], m["kT"]))
        for layout in _layouts(m["lenY"]):
            lat = _cabi.Lattices(_cabi.MODEL_LATGAS, m["lenX"], m["lenY"], seed=m["seed"], layout=layout,
                                 replica_offset=m["replica"], nreplicas_global=m["replica"] + 1)
            lat.init_random()
            assert np.array_equal(lat.download(0), z[n + "_init"])
            xs, ys = np.meshgrid(np.arange(m["lenX"]), np.arange(m["lenY"]), indexing="ij")
            c, nn = lat.propose(0, xs.ravel(), ys.ravel())
            dE = np.array([model._pmc_dE(a, b) for a, b in zip(c, nn)]).reshape(m["lenX"], m["lenY"])
            assert np.array_equal(dE, z[n + "_dE"])  # bit-exact, every site
            b0, n0 = lat.reduce_observables()
            assert model._pmc_energy(int(b0[0]), int(n0[0])) == pytest.approx(float(z[n + "_e0"]), abs=1e-9)
            lat.set_accept_tables(thr[None, :])
            for sw in range(m["nsweeps"]):
                lat.sweep(1)
                assert np.array_equal(lat.download(0), z[n + "_configs"][sw]), (n, layout, sw)
                b, N = lat.tracked()
                assert float(N[0]) == z[n + "_density"][sw]
                assert model._pmc_energy(int(b[0]), int(N[0])) == pytest.approx(z[n + "_energy"][sw], abs=1e-9)
                b2, N2 = lat.reduce_observables()
                assert (b2[0], N2[0]) == (b[0], N[0])
            lat.close()


@pytest.mark.parametrize("shape", [(64, 256), (20, 36), (128, 1024)])
def test_medium_lattices_match_oracle(shape):
    from py_mc_b200 import _cabi
    seed = 77
    thr = o_latgas.threshold_table(-0.75, -0.5, 0.25, 0.9)
    ref = o_latgas.init_config(seed, 0, *shape)
    lats = []
    for layout in _layouts(shape[1]):
        lat = _cabi.Lattices(_cabi.MODEL_LATGAS, shape[0], shape[1], seed=seed, layout=layout)
        lat.init_random()
        lat.set_accept_tables(thr[None, :])
        lats.append(lat)
    B, N = o_latgas.bond_sum(ref), int(ref.sum())
    for sw in range(4):
        _, dB, dN = o_latgas.sweep(ref, thr, seed, 0, sw)
        B, N = B + dB, N + dN
        for lat in lats:
            lat.sweep(1)
            assert np.array_equal(lat.download(0), ref), (shape, lat.layout, sw)
            b, n = lat.tracked()
            assert (int(b[0]), int(n[0])) == (B, N)


def test_inconsistent_table_is_rejected():
    from py_mc_b200 import _cabi
    lat = _cabi.Lattices(_cabi.MODEL_LATGAS, 8, 8)
    bad = np.full((1, 10), 0xFFFFFFFF, dtype=np.uint32)
    bad[0, 2] = bad[0, 7] = 1000  # dE > 0 for both occupancies at nn = 2
    with pytest.raises(_cabi.PymcError, match="both occupancies"):
        lat.set_accept_tables(bad)


def test_langmuir_isotherm_J0(tmp_path):
    """J = 0: sites are independent and the coverage is 1/(1+exp((Eads-mu)/kT))."""
    import py_mc_b200
    from py_mc_b200.latgas import latgas_config, latticegas, observer
    from py_mc_b200.mc import CanonicalMonteCarlo
    old = os.getcwd()
    os.chdir(tmp_path)
    try:
        py_mc_b200.set_seed(5)
        for x in (-2.5608714821730496, 0.7, 2.0):  # (Eads - mu)/kT
            cfg = latgas_config(64, 64)
            cfg.prepare_random()
            calc = CanonicalMonteCarlo(latticegas(Eads=x, J=0, mu=0.0), 1.0, cfg)
            calc.run(50)
            obs = observer()
            out = calc.run(400, 1, obs)  # observer in the third slot, as the shipped script passes it
            theta = out[1] / 4096.0
            exact = 1.0 / (1.0 + np.exp(x))
            assert abs(theta - exact) < 6 * np.sqrt(exact * (1 - exact) / (4096 * 400)) + 1e-4, (x, theta, exact)
    finally:
        os.chdir(old)
