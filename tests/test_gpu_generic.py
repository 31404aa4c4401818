"""GPU: table-driven generic lattice models (row f3) -- user-written plain-Python models run on the
sweep kernels after their (state, neighbours) -> dE table has been probed from their own energy()."""
import itertools
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


class plain_ising:
    """A user model in the reference's style: Ising with an external field, numpy only."""
    model_name = "plain_ising"

    def __init__(self, J, h=0.0):
        self.J, self.h = J, h

    def energy(self, cfg):
        c = cfg.config
        return float((self.J * c * (np.roll(c, 1, 0) + np.roll(c, 1, 1))).sum() - self.h * c.sum())


class next_nearest:
    """Outside the supported family: diagonal couplings."""

    def energy(self, cfg):
        c = cfg.config
        return float((c * (np.roll(c, 1, 0) + np.roll(np.roll(c, 1, 0), 1, 1))).sum())


@pytest.fixture()
def workdir(tmp_path):
    old = os.getcwd()
    os.chdir(tmp_path)
    yield tmp_path
    os.chdir(old)


def test_wrapped_ising_follows_the_native_trajectory(workdir):
    """Same thresholds, same random stream: the generic (10-class) path and the native Ising path agree bit for bit."""
    import py_mc_b200
    from py_mc_b200.generic import nn_lattice_model
    from py_mc_b200.ising2D import ising2D, ising2D_config
    from py_mc_b200.mc import CanonicalMonteCarlo
    from oracle import ising as o_ising
    seed, kT = 515, 2.1
    init = o_ising.init_config(seed, 0, 12, 20)
    py_mc_b200.set_seed(seed)
    gm = nn_lattice_model(plain_ising(-1.0))
    gcfg = gm.make_config(12, 20)
    gcfg.config[:, :] = init
    assert gm.energy(gcfg) == -1.0 * o_ising.bond_sum(init)
    gcalc = CanonicalMonteCarlo(gm, kT, gcfg)
    gcalc.run(9)
    py_mc_b200.set_seed(seed)
    ncfg = ising2D_config(12, 20)
    ncfg.config[:, :] = init
    ncalc = CanonicalMonteCarlo(ising2D(-1.0), kT, ncfg)
    ncalc.run(9)
    assert np.array_equal(gcfg.config, ncfg.config)
    assert gcalc.energy == pytest.approx(ncalc.energy, abs=1e-9)
    dconfig, dE = gm.trialstep(gcfg, 0.0)
    x, y = dconfig
    c = ncfg.peek()
    nn = c[x - 1, y] + c[(x + 1) % 12, y] + c[x, y - 1] + c[x, (y + 1) % 20]
    assert dE == pytest.approx(-2.0 * -1.0 * c[x, y] * nn, abs=1e-9)


def test_ising_in_a_field_matches_exact_enumeration(workdir):
    import py_mc_b200
    from py_mc_b200.generic import nn_lattice_model
    from py_mc_b200.mc import CanonicalMonteCarlo, observer_base
    J, h, kT, L = -1.0, 0.3, 2.5, 4
    user = plain_ising(J, h)
    # exact canonical averages over the 2^16 states
    states = np.array(list(itertools.product((-1.0, 1.0), repeat=L * L))).reshape(-1, L, L)
    E = (J * states * (np.roll(states, 1, 1) + np.roll(states, 1, 2))).sum(axis=(1, 2)) - h * states.sum(axis=(1, 2))
    w = np.exp(-(E - E.min()) / kT)
    e_exact = float((E * w).sum() / w.sum())
    m_exact = float((states.sum(axis=(1, 2)) * w).sum() / w.sum())

    class rec(observer_base):
        def logfunc(self, cs):
            up = cs.model.upper_count(cs.config)
            return cs.energy, 2.0 * up - L * L

    gm = nn_lattice_model(user)
    es, ms = [], []
    for rep in range(24):  # independent short chains give the error bar
        py_mc_b200.set_seed(77 + rep)  # a seed per chain: chains that share a random stream coalesce
        cfg = gm.make_config(L, L)
        cfg.prepare_random()
        calc = CanonicalMonteCarlo(gm, kT, cfg)
        calc.run(200)
        out = calc.run(1500, 1, 10 ** 9, rec())
        es.append(out[0])
        ms.append(out[1])
    es, ms = np.array(es), np.array(ms)
    assert abs(es.mean() - e_exact) < 5 * es.std(ddof=1) / np.sqrt(len(es)) + 1e-3, (es.mean(), e_exact)
    assert abs(ms.mean() - m_exact) < 5 * ms.std(ddof=1) / np.sqrt(len(ms)) + 1e-3, (ms.mean(), m_exact)


def test_models_outside_the_family_are_refused():
    from py_mc_b200.generic import nn_lattice_model
    with pytest.raises(TypeError, match="not a two-state nearest-neighbour"):
        nn_lattice_model(next_nearest())
