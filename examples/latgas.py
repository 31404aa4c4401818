#!/usr/bin/env python
"""Adsorption isotherm of the lattice gas -- the experiment of the reference's ``examples/latgas.py``
``__main__`` (lines 99-129): coverage against pressure at T = 300 K for J = 0 (Langmuir), on the GPU engine.
The reference uses a 5 x 5 lattice, which cannot be checkerboarded; the default here is 6 x 6.

    python examples/latgas.py [--size 6] [--sweeps 1000]
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from py_mc_b200.latgas import latgas_config, latticegas, observer
from py_mc_b200.mc import CanonicalMonteCarlo, binning

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=6)
    ap.add_argument("--sweeps", type=int, default=1000)
    a = ap.parse_args()
    kb, T, h = 1.38064852e-23, 300, 6.62607004e-34
    m = 1.0 / 6.023 * 1e-26
    nsite = a.size * a.size
    cfg = latgas_config(a.size, a.size)
    cfg.prepare_random()
    mu0 = -kb * T * np.log((2.0 * np.pi * m * kb * T / h ** 2) ** (3 / 2) * kb * T)
    Eads = -0.4 * 1.602e-19
    for p in np.linspace(1e4, 0.0001, 20):
        mu = mu0 + kb * T * np.log(p)
        model = latticegas(Eads=Eads, J=0, mu=mu)
        calc = CanonicalMonteCarlo(model, kb * T, cfg)
        calc.run(a.sweeps)
        rec = observer()
        obs = calc.run(a.sweeps, 1, rec)
        theta_exact = 1.0 / (1.0 + np.exp((Eads - mu) / (kb * T)))
        err = binning(rec.density_obs, 6)
        print(p, "\t", obs[1] / nsite, "\t", theta_exact, "\t", max(err) / nsite, flush=True)
