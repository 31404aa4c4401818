"""CPU: the C restatement (oracle/pymc_oracle.c) agrees with the numpy oracle and with the fixtures recorded
from the reference, so it can stand in for them on lattices the numpy oracle is too slow for."""
import numpy as np

from oracle import cbuild
from oracle import ising as o_ising


def test_c_oracle_matches_reference_fixtures(golden_ising):
    co = cbuild.load()
    z, meta = golden_ising
    for m in meta:
        n = m["name"]
        cfg = co.init(m["lenX"], m["lenY"], m["seed"], m["replica"])
        assert np.array_equal(cfg, z[n + "_init"])
        thr = o_ising.threshold_table(m["J"], m["kT"])[1]
        B, M = co.observables(cfg)
        assert (B, M) == (o_ising.bond_sum(cfg), o_ising.magnetization(cfg))
        for sw in range(m["nsweeps"]):
            nacc, dB, dM = co.sweep(cfg, thr, m["seed"], m["replica"], sw)
            B, M = B + dB, M + dM
            assert np.array_equal(cfg, z[n + "_configs"][sw]), (n, sw)
            assert (B, M) == co.observables(cfg)
            assert M == z[n + "_mag"][sw]


def test_c_oracle_matches_numpy_oracle_medium():
    co = cbuild.load()
    seed, shape, kT = 31, (96, 388), 2.2
    a = co.init(shape[0], shape[1], seed, 4)
    b = o_ising.init_config(seed, 4, *shape)
    assert np.array_equal(a, b)
    thr = o_ising.threshold_table(-1.0, kT)
    for sw in range(3):
        ra = co.sweep(a, thr[1], seed, 4, sw)
        rb = o_ising.sweep(b, -1.0, kT, seed, 4, sw, thr)
        assert ra == rb and np.array_equal(a, b)
