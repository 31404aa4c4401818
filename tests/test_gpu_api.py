"""GPU: the drop-in Python API (py_mc_b200.mc / py_mc_b200.ising2D) reproduces the reference's
results -- trajectories from the golden fixtures, return values of run(), exact 4x4 physics."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

# exact canonical averages of the 4x4 periodic Ising model, J = -1
EXACT_4x4 = {
    1.0: (-1.997158440256, 0.999274583701),
    2.269185314: (-1.565623787638, 0.843860444813),
    3.0: (-1.017069626955, 0.601291326039),
    5.0: (-0.456135369543, 0.342765627554),
}


@pytest.fixture()
def workdir(tmp_path):
    old = os.getcwd()
    os.chdir(tmp_path)
    yield tmp_path
    os.chdir(old)


def test_run_reproduces_reference_trajectory(golden_ising, workdir):
    import py_mc_b200
    from py_mc_b200.ising2D import ising2D, ising2D_config
    from py_mc_b200.mc import CanonicalMonteCarlo
    z, meta = golden_ising
    for m in [c for c in meta if c["replica"] == 0]:
        py_mc_b200.set_seed(m["seed"])
        cfg = ising2D_config(m["lenX"], m["lenY"])
        cfg.config[:, :] = z[m["name"] + "_init"]
        model = ising2D(m["J"])
        assert model.energy(cfg) == pytest.approx(float(z[m["name"] + "_e0"]), abs=1e-12)
        calc = CanonicalMonteCarlo(model, m["kT"], cfg)
        assert calc.attempts_per_step == m["lenX"] * m["lenY"]
        assert calc.run(m["nsweeps"]) is None  # nothing sampled -> None (mc.py:213-214)
        assert np.array_equal(calc.config.config.astype(np.int8), z[m["name"] + "_configs"][-1])
        if float(m["J"]).is_integer():
            assert calc.energy == z[m["name"] + "_energy"][-1]
        assert model.magnetization(calc.config) == z[m["name"] + "_mag"][-1]


def test_trialstep_newconfig_protocol(golden_ising_de, workdir):
    from py_mc_b200 import ising2D as mod
    z, meta = golden_ising_de
    m = meta[0]
    cfg = mod.ising2D_config(m["lenX"], m["lenY"])
    cfg.config[:, :] = z[m["name"] + "_init"]
    model = mod.ising2D(m["J"])

    class Fixed:
        def __init__(self, seq):
            self.seq = list(seq)

        def randrange(self, n):
            return self.seq.pop(0)

    old = mod.rand
    try:
        mod.rand = Fixed([3, 7, 0, 0, 9, 9])
        for (x, y) in [(3, 7), (0, 0), (9, 9)]:
            dconfig, dE = model.trialstep(cfg, 0.0)
            assert dconfig == [x, y]
            assert dE == z[m["name"] + "_dE"][x, y]
    finally:
        mod.rand = old
    before = cfg.config.copy()
    model.newconfig(cfg, [3, 7])
    assert cfg.config[3, 7] == -before[3, 7]
    # energy after the host-side edit is recomputed from the edited array
    assert model.energy(cfg) == float(z[m["name"] + "_energy"]) + z[m["name"] + "_dE"][3, 7]


def test_run_with_observer_returns_means(workdir):
    import py_mc_b200
    from py_mc_b200.ising2D import ising2D, ising2D_config, observer
    from py_mc_b200.mc import CanonicalMonteCarlo, binning
    py_mc_b200.set_seed(12345)
    for kT, (e_exact, m_exact) in EXACT_4x4.items():
        cfg = ising2D_config(4, 4)
        cfg.prepare_random()
        model = ising2D(-1.0)
        calc = CanonicalMonteCarlo(model, kT, cfg)
        calc.run(2000)
        obs = observer()
        nsamp = 2 ** 14
        out = calc.run(nsamp, 1, 10 ** 9, obs)
        assert len(out) == 2
        assert len(obs.energy_obs) == nsamp + 2  # + the pre-run probe (mc.py:193) and obs_info's call (mc.py:125,211)
        e = np.asarray(obs.energy_obs[1:-1]) / 16
        mag = np.asarray(obs.magnet_obs[1:-1]) / 16
        assert out[0] / 16 == pytest.approx(e.mean(), rel=1e-12)
        err_e = max(binning(e, 8)) + 1e-9
        err_m = max(binning(mag, 8)) + 1e-9
        assert abs(e.mean() - e_exact) < 5 * err_e, (kT, e.mean(), e_exact, err_e)
        assert abs(mag.mean() - m_exact) < 5 * err_m, (kT, mag.mean(), m_exact, err_m)
    assert os.path.exists("obs.dat")


def test_observer_in_third_positional_slot(workdir):
    from py_mc_b200.ising2D import ising2D, ising2D_config, observer
    from py_mc_b200.mc import CanonicalMonteCarlo
    cfg = ising2D_config(6, 6)
    cfg.prepare_random()
    calc = CanonicalMonteCarlo(ising2D(-1.0), 2.0, cfg)
    out = calc.run(10, 2, observer())  # the form ParallelMC.run uses (mc_mpi.py:66)
    assert out is not None and len(out) == 2


def test_non_device_model_is_rejected(workdir):
    from py_mc_b200.mc import CanonicalMonteCarlo, model
    with pytest.raises(TypeError, match="no CPU fallback"):
        CanonicalMonteCarlo(model(), 1.0, object()).run(1)


def test_swap_configs_helper(workdir):
    from py_mc_b200.ising2D import ising2D, ising2D_config
    from py_mc_b200.mc import CanonicalMonteCarlo, swap_configs
    model = ising2D(-1.0)
    reps = []
    for kT in (1.5, 3.0):
        c = ising2D_config(6, 6)
        c.prepare_random()
        calc = CanonicalMonteCarlo(model, kT, c)
        calc.run(3)
        reps.append(calc)
    c0, c1, e0, e1 = reps[0].config, reps[1].config, reps[0].energy, reps[1].energy
    reps, n = swap_configs(reps, 0, 0)
    assert n == 1 and reps[0].config is c1 and reps[1].config is c0
    assert (reps[0].energy, reps[1].energy) == (e1, e0)
    reps[0].run(2)  # the engines keep working on the swapped configurations
    assert reps[0].energy == model.energy(reps[0].config)
