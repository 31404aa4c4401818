"""GPU: hard-disc kernels (K8) and the 2D_hardcore drop-in against the fixture recorded from the
reference: coordinates are float64 and compared bit-exactly."""
import os

import numpy as np
import pytest

from oracle import harddisc as o_hd

pytestmark = pytest.mark.gpu


def test_batched_trials_match_reference(golden_discs):
    from py_mc_b200 import _cabi
    z, meta = golden_discs
    for m in meta:
        n = m["name"]
        # three replicas; the fixture's global replica index sits at local position `me`
        me = 1 if m["replica"] > 0 else 0
        d = _cabi.Discs(3, m["ndisc"], m["L"], m["rdisc"], m["maxstep"], seed=m["seed"],
                        replica_offset=m["replica"] - me)
        for r in range(3):
            d.upload(r, z[n + "_start"])
        ov, counts = d.observe(m["dr"], len(z[n + "_gridx"]))
        assert not ov.any()
        g0 = counts[me].astype(np.float64) * (m["g_pref"] / (z[n + "_gridx"] * m["dr"]))
        assert np.array_equal(g0, z[n + "_g0"])
        for k in range(m["ntrials"] // 500):
            d.run(500)
            assert np.array_equal(d.download(me), z[n + "_snaps"][k]), (n, k)  # bit-exact
        t, acc = d.counters()
        assert t == m["ntrials"] and int(acc[me]) == m["nacc"]
        ov, counts = d.observe(m["dr"], len(z[n + "_gridx"]))
        assert not ov.any()
        g = counts[me].astype(np.float64) * (m["g_pref"] / (z[n + "_gridx"] * m["dr"]))
        assert np.array_equal(g, z[n + "_g_end"])
        assert not np.array_equal(d.download(me), d.download(2))  # replicas use different streams
        d.close()


def test_isotropic_mode_matches_specification():
    from py_mc_b200 import _cabi
    L, r, N, step, seed = 1.5, 0.1, 18, 0.08, 5
    c = o_hd.prepare_ordered(L, N, r)
    d = _cabi.Discs(1, N, L, r, step, seed=seed, isotropic=True)
    d.upload(0, c)
    d.run(800)
    for t in range(800):
        o_hd.trial(c, L, r, step, seed, 0, t, isotropic=True)
    assert np.array_equal(d.download(0), c)


def test_dropin_api(golden_discs, tmp_path):
    import py_mc_b200
    from py_mc_b200.hardcore import HC_2D, HC_2D_config, g_r, observables
    from py_mc_b200.mc import CanonicalMonteCarlo, grid_1D
    z, meta = golden_discs
    m = next(c for c in meta if c["replica"] == 0)
    n = m["name"]
    old = os.getcwd()
    os.chdir(tmp_path)
    try:
        py_mc_b200.set_seed(m["seed"])
        cfg = HC_2D_config(m["L"], m["ndisc"], m["rdisc"])
        cfg.prepare_ordered()
        assert np.array_equal(cfg.coors, z[n + "_start"])
        model = HC_2D(m["maxstep"])
        grid = grid_1D(m["dr"], m["dr"], m["L"] / 2.0)
        assert np.array_equal(g_r(cfg, grid), z[n + "_g0"])
        assert model.energy(cfg) == 0
        calc = CanonicalMonteCarlo(model, 1, cfg, grid=grid)
        dconfig, dE = model.trialstep(cfg, 0)  # the next trial, not applied
        assert dconfig[0] == z[n + "_ids"][0] and dE in (0, float("inf"))
        assert np.array_equal(cfg.coors, z[n + "_start"])
        calc.run(500)
        assert np.array_equal(cfg.coors, z[n + "_snaps"][0])
        for _ in range(10):
            calc.MCstep()  # one trial each, as in the reference
        calc.run(490)
        assert np.array_equal(cfg.coors, z[n + "_snaps"][1])
        assert calc.energy == 0 and calc.attempts_per_step == 1
        assert observables(calc, None).shape == (len(grid.x),)
        bad = HC_2D_config(m["L"], 2, m["rdisc"])
        bad.coors[:] = [[0.1, 0.1], [0.1 + m["rdisc"], 0.1]]
        assert model.energy(bad) == float("inf")
    finally:
        os.chdir(old)


def test_disc_batch_keeps_valid_states():
    from py_mc_b200.hardcore import DiscBatch
    from py_mc_b200.mc import grid_1D
    b = DiscBatch(64, 1.0, 20, 1.0 / 14.0, 1.0 / 28.0, isotropic=True, seed=3)
    b.run(5000)
    assert not b.any_overlap().any()
    acc = b.acceptance()
    assert 0.2 < acc.mean() < 1.0
    g = b.g_r(grid_1D(0.01, 0.01, 0.5))
    assert g.shape == (64, 49)
    assert np.all(g[:, :13] == 0)  # no pair closer than one diameter (0.1428)
