#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ from the UNMODIFIED reference.

Run in the build container (where /root/reference exists):

    python tests/golden/make_golden.py

The reference (``py_mc/mc.py``, ``examples/ising2D.py``, ``examples/latgas*.py``,
``py_mc/mc_mpi.py``) is imported as-is and driven with *injected* random draws:
its module-level ``random`` / ``rand`` names are replaced by objects that return
the checkerboard site order and the per-site uniforms of the counter-based
stream (``oracle/philox.py``).  What it then computes -- dE, accept/reject,
tracked energies, final configurations, exchange histories -- is recorded.
Nothing of the reference is copied; only its outputs are stored.
The GPU box has no /root/reference: tests there read these files only.
"""
import json
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.environ.get("PYMC_REFERENCE", "/root/reference")
sys.path.insert(0, REF)
sys.path.insert(0, os.path.join(REF, "examples"))

from oracle import philox, ising as o_ising  # noqa: E402

import py_mc.mc as refmc  # noqa: E402
import ising2D as ref_ising  # noqa: E402


class SiteInjector:
    """Stands in for the ``random`` module inside a reference example module and for
    ``py_mc.mc.random``: returns scheduled sites from ``randrange`` and the uniform of the
    *current* site from ``random()`` (the reference draws it only when dE >= 0)."""

    def __init__(self):
        self.sites = []
        self.us = []
        self.pos = -1
        self.phase = 0
        self.ndraws = 0

    def load(self, sites, us):
        self.sites = sites
        self.us = us
        self.pos = -1
        self.phase = 0

    def randrange(self, n):
        if self.phase == 0:
            self.pos += 1
            self.phase = 1
            v = self.sites[self.pos][0]
        else:
            self.phase = 0
            v = self.sites[self.pos][1]
        assert 0 <= v < n
        return int(v)

    def random(self):
        self.ndraws += 1
        return float(self.us[self.pos])


def checkerboard_schedule(seed, replica, sweep, lenX, lenY):
    sites, us = [], []
    for colour in (0, 1):
        u = philox.x32_to_uniform(o_ising.site_x32(seed, replica, sweep, colour, lenX, lenY))
        for i in range(lenX):
            q = (i + colour) & 1
            for k in range(lenY // 2):
                sites.append((i, 2 * k + q))
                us.append(u[i, k])
    return sites, us


ISING_CASES = [
    # name, lenX, lenY, J, kT, seed, replica, nsweeps
    ("c1_10x10_T5", 10, 10, -1.0, 5.0, 20260611, 0, 12),
    ("c1_10x10_Tc", 10, 10, -1.0, 2.269185314213022, 20260611, 1, 12),
    ("c1_10x10_T1", 10, 10, -1.0, 1.0, 7, 2, 12),
    ("c1_10x10_T001", 10, 10, -1.0, 0.01, 7, 3, 8),
    ("afm_6x8_T2", 6, 8, 1.0, 2.0, 99, 0, 10),
    ("nondyadic_8x12", 8, 12, 0.3, 0.7, 1234567890123, 5, 10),
    ("word_16x128_Tc", 16, 128, -1.0, 2.269185314213022, 42, 0, 6),
    ("partial_4x260_T3", 4, 260, -1.0, 3.0, 43, 7, 6),
    ("multi_8x384_T2", 8, 384, -1.0, 2.0, 44, 300, 5),
    ("tall_64x4_T25", 64, 4, -1.0, 2.5, 45, 1, 8),
]


def gen_ising(out):
    inj = SiteInjector()
    ref_ising.rand = inj  # site selection (ising2D.py:32-33)
    refmc.random = inj.random  # Metropolis uniform (mc.py:182)
    meta = []
    cwd = os.getcwd()
    tmp = tempfile.mkdtemp()
    os.chdir(tmp)
    try:
        for name, lenX, lenY, J, kT, seed, replica, nsweeps in ISING_CASES:
            cfg = ref_ising.ising2D_config(lenX, lenY)
            init = o_ising.init_config(seed, replica, lenX, lenY)
            cfg.config[:, :] = init
            model = ref_ising.ising2D(J)
            calc = refmc.CanonicalMonteCarlo(model, kT, cfg)
            e0 = model.energy(cfg)
            energies, mags, configs = [], [], []
            for sw in range(nsweeps):
                sites, us = checkerboard_schedule(seed, replica, sw, lenX, lenY)
                inj.load(sites, us)
                calc.run(lenX * lenY)  # nsteps single-site trials = one sweep
                assert inj.pos == lenX * lenY - 1
                energies.append(calc.energy)
                mags.append(model.magnetization(cfg))
                configs.append(cfg.config.astype(np.int8).copy())
            out[name + "_init"] = init
            out[name + "_e0"] = np.float64(e0)
            out[name + "_energy"] = np.array(energies, dtype=np.float64)
            out[name + "_mag"] = np.array(mags, dtype=np.float64)
            out[name + "_configs"] = np.array(configs, dtype=np.int8)
            meta.append(dict(name=name, lenX=lenX, lenY=lenY, J=J, kT=kT, seed=seed,
                             replica=replica, nsweeps=nsweeps))
            print("ising", name, "E0", e0, "E_end", energies[-1], "M_end", mags[-1])
    finally:
        os.chdir(cwd)
    out["meta"] = np.array(json.dumps(meta))


def gen_ising_de(out):
    """dE / dconfig for every site of fixed configurations, straight from ising2D.trialstep,
    and energy()/magnetization() of the same configurations."""
    inj = SiteInjector()
    ref_ising.rand = inj
    meta = []
    for name, lenX, lenY, J, seed in [("de_10x10", 10, 10, -1.0, 5), ("de_6x130", 6, 130, -1.0, 6),
                                      ("de_8x8_J03", 8, 8, 0.3, 7), ("de_4x4_afm", 4, 4, 2.5, 8)]:
        cfg = ref_ising.ising2D_config(lenX, lenY)
        init = o_ising.init_config(seed, 0, lenX, lenY)
        cfg.config[:, :] = init
        model = ref_ising.ising2D(J)
        sites = [(x, y) for x in range(lenX) for y in range(lenY)]
        inj.load(sites, [0.0] * len(sites))
        dEs = np.zeros((lenX, lenY))
        for (x, y) in sites:
            dconfig, dE = model.trialstep(cfg, 0.0)
            assert dconfig == [x, y]
            dEs[x, y] = dE
        assert np.array_equal(cfg.config, init)  # trialstep leaves config unchanged
        out[name + "_init"] = init
        out[name + "_dE"] = dEs
        out[name + "_energy"] = np.float64(model.energy(cfg))
        out[name + "_mag"] = np.float64(model.magnetization(cfg))
        meta.append(dict(name=name, lenX=lenX, lenY=lenY, J=J, seed=seed))
        print("ising_de", name, out[name + "_energy"], out[name + "_mag"])
    out["meta"] = np.array(json.dumps(meta))


def main():
    which = sys.argv[1:] or ["ising", "ising_de"]
    gens = {"ising": gen_ising, "ising_de": gen_ising_de}
    try:
        from make_golden_more import MORE  # later families (rx, latgas, discs)
        gens.update(MORE)
        if not sys.argv[1:]:
            which += list(MORE)
    except ImportError:
        pass
    for w in which:
        out = {}
        gens[w](out)
        path = os.path.join(HERE, w + ".npz")
        np.savez_compressed(path, **out)
        print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
