#!/usr/bin/env python
"""2D Ising temperature scan on the GPU engine -- the experiment of the reference's
``examples/ising2D.py`` ``__main__`` block (lines 97-129): one configuration annealed from kT = 5|J| down
to 0.01|J| in ten steps, equilibration then sampling at every temperature, blocking errors to
``binningE.dat`` / ``binningM.dat``, one result line per temperature.

Steps are sweeps here (one sweep = size*size single-spin trials), so the defaults do the same number
of *trials* as the reference's 40960 + 163840 single-spin steps on its 10 x 10 lattice.

    python examples/ising2D.py [--size 10] [--eq-sweeps 410] [--mc-sweeps 1638] [--seed 1]
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import py_mc_b200
from py_mc_b200.ising2D import ising2D, ising2D_config, observer
from py_mc_b200.mc import CanonicalMonteCarlo, binning


def scan(size, eq_sweeps, mc_sweeps, J=-1.0, temps=None, out=sys.stdout):
    nspin = size * size
    temps = np.linspace(5.0, 0.01, 10) if temps is None else temps
    binlevel = max(int(np.log2(max(mc_sweeps // 30, 1))), 0)
    cfg = ising2D_config(size, size)
    cfg.prepare_random()
    model = ising2D(J)
    rows = []
    with open("binningE.dat", "w") as fe, open("binningM.dat", "w") as fm:
        for t in temps:
            kT = abs(J) * t
            calc = CanonicalMonteCarlo(model, kT, cfg)
            calc.run(eq_sweeps)
            rec = observer()
            means = calc.run(mc_sweeps, 1, 10 ** 9, rec)
            errE = binning(np.asarray(rec.energy_obs) / nspin, binlevel)
            errM = binning(np.asarray(rec.magnet_obs) / nspin, binlevel)
            fe.write("\n".join(str(x) for x in errE) + "\n\n\n")
            fm.write("\n".join(str(x) for x in errM) + "\n\n\n")
            row = (kT, means[0] / nspin, means[1] / nspin, float(np.max(errE)), float(np.max(errM)))
            rows.append(row)
            print("\t".join(str(v) for v in row), file=out, flush=True)
    return rows


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=10)
    ap.add_argument("--eq-sweeps", type=int, default=410)
    ap.add_argument("--mc-sweeps", type=int, default=1638)
    ap.add_argument("--seed", type=int, default=None)
    a = ap.parse_args()
    if a.seed is not None:
        py_mc_b200.set_seed(a.seed)
        import random
        random.seed(a.seed)
    scan(a.size, a.eq_sweeps, a.mc_sweeps)
