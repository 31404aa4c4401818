"""CPU: pin oracle/latgas.py to the fixture recorded from the reference's examples/latgas.py."""
import numpy as np
import pytest

from oracle import latgas as o_latgas


def test_dE_energy_and_trajectories_match_reference(golden_latgas):
    z, meta = golden_latgas
    for m in meta:
        n = m["name"]
        J, Eads, mu, kT = m["J"], m["Eads"], m["mu"], m["kT"]
        cfg = z[n + "_init"].copy()
        assert np.array_equal(cfg, o_latgas.init_config(m["seed"], m["replica"], m["lenX"], m["lenY"]))
        nn = o_latgas.neighbour_sum(cfg)
        dE = np.array([[o_latgas.delta_e(J, Eads, mu, cfg[x, y], nn[x, y]) for y in range(m["lenY"])]
                       for x in range(m["lenX"])])
        assert np.array_equal(dE, z[n + "_dE"])  # bit-exact float64
        assert o_latgas.energy(cfg, J, Eads, mu) == pytest.approx(float(z[n + "_e0"]), abs=1e-9)
        thr = o_latgas.threshold_table(J, Eads, mu, kT)
        # a consistent table never has dE > 0 for both occupancies at the same neighbour count
        assert all(not (thr[k] != o_latgas.ALWAYS and thr[5 + k] != o_latgas.ALWAYS) for k in range(5))
        for sw in range(m["nsweeps"]):
            o_latgas.sweep(cfg, thr, m["seed"], m["replica"], sw)
            assert np.array_equal(cfg, z[n + "_configs"][sw]), (n, sw)
            assert float(cfg.sum()) == z[n + "_density"][sw]
            assert o_latgas.energy(cfg, J, Eads, mu) == pytest.approx(z[n + "_energy"][sw], abs=1e-9)
