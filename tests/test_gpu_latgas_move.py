"""GPU: drop-in for examples/latgas_move.py -- trialstep (flip and move branches, the reference's
defective move dE reproduced in 'compat' mode), newconfig, and the engine's flip + exchange step."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


class Scripted:
    def __init__(self):
        self.script = []

    def random(self):
        kind, v = self.script.pop(0)
        assert kind == "random"
        return v

    def randrange(self, n):
        kind, v = self.script.pop(0)
        assert kind == "randrange"
        return int(v)

    def choice(self, seq):
        kind, v = self.script.pop(0)
        assert kind == "choice"
        return seq[int(v) % len(seq)]


def test_trialstep_and_newconfig_match_reference(golden_latgas_move):
    from py_mc_b200 import _cabi, latgas as flipmod, latgas_move as mod
    z, meta = golden_latgas_move
    inj = Scripted()
    old = (mod.rand, flipmod.rand)
    mod.rand = flipmod.rand = inj
    try:
        for m in meta:
            n = m["name"]
            cfg = mod.latgas_config(m["lenX"], m["lenY"])
            cfg.config[:, :] = z[n + "_init"]
            model = mod.latticegas(Eads=m["Eads"], J=m["J"], mu=m["mu"])
            if m["lenX"] % 2:  # the shipped 5x5 lattice: odd edges cannot be checkerboarded; refused loudly
                with pytest.raises(_cabi.PymcError, match="even"):
                    model.density(cfg)
                continue
            for script, want_dc, want_dE in zip(m["scripts"], z[n + "_dconfig"], z[n + "_dE"]):
                inj.script = [(k, v) for k, v in script]
                dconfig, dE = model.trialstep(cfg, 0.0)
                assert not inj.script
                assert [int(v) for v in dconfig] == [int(v) for v in want_dc if v >= 0]
                assert dE == want_dE  # bit-exact, including the defective move branch
            assert np.array_equal(cfg.peek().astype(np.int8), z[n + "_init"])
            model.newconfig(cfg, [int(v) for v in z[n + "_move"]])
            assert np.array_equal(cfg.config.astype(np.int8), z[n + "_after_move"])
    finally:
        mod.rand, flipmod.rand = old


def test_physical_move_dE_equals_energy_difference():
    from py_mc_b200 import latgas_move as mod
    import random
    random.seed(7)
    cfg = mod.latgas_config(8, 8)
    cfg.prepare_random()
    model = mod.latticegas(Eads=-0.5, J=-1.0, mu=0.25, move_dE="physical")
    for _ in range(30):
        e0 = model.energy(cfg)
        dconfig, dE = model.trialstep(cfg, e0)
        if len(dconfig) != 4:
            continue
        trial = mod.latgas_config(8, 8)
        trial.config[:, :] = cfg.peek()
        model.newconfig(trial, dconfig)
        assert model.energy(trial) - e0 == pytest.approx(dE, abs=1e-12)


def test_engine_step_is_flip_sweep_plus_half_exchange_sweep(tmp_path):
    import py_mc_b200
    from py_mc_b200 import latgas_move as mod
    from py_mc_b200.mc import CanonicalMonteCarlo
    from oracle import latgas as o_latgas, pairmove as o_pair
    old = os.getcwd()
    os.chdir(tmp_path)
    try:
        seed, J, Eads, mu, kT = 606, -1.0, -0.5, 0.25, 0.9
        py_mc_b200.set_seed(seed)
        ref = o_latgas.init_config(seed, 0, 8, 12)
        cfg = mod.latgas_config(8, 12)
        cfg.config[:, :] = ref
        model = mod.latticegas(Eads=Eads, J=J, mu=mu)
        calc = CanonicalMonteCarlo(model, kT, cfg)
        assert calc.attempts_per_step == 2 * 96
        calc.run(6)
        thr_f = o_latgas.threshold_table(J, Eads, mu, kT)
        thr_p = o_pair.threshold_table(J, kT)
        for s in range(6):
            o_latgas.sweep(ref, thr_f, seed, 0, s)
            o_pair.sweep_classes(ref, thr_p, seed, 0, s, classes=range(8) if s % 2 == 0 else range(8, 16))
        assert np.array_equal(cfg.config.astype(np.int8), ref)
        assert calc.energy == pytest.approx(o_latgas.energy(ref, J, Eads, mu), abs=1e-12)
    finally:
        os.chdir(old)
