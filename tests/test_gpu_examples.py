"""GPU: the example drivers (the reference's L3 scripts on the GPU engine) run and give sane physics."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_ising2d_scan_matches_exact_10x10_limits(tmp_path):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "ising2D.py"), "--size", "10",
                          "--eq-sweeps", "300", "--mc-sweeps", "1200", "--seed", "7"],
                         cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    rows = np.array([[float(v) for v in ln.split("\t")] for ln in out.stdout.strip().splitlines()])
    assert rows.shape == (10, 5)
    assert rows[0, 0] == 5.0 and abs(rows[-1, 0] - 0.01) < 1e-12
    assert -0.6 < rows[0, 1] < -0.3            # hot: E/N near the 10x10 value at kT = 5
    assert rows[-1, 1] == -2.0 and rows[-1, 2] == 1.0   # ground state reached at kT = 0.01
    assert np.all(np.diff(rows[:, 1]) < 0.05)   # energy falls as the system is cooled
    assert os.path.exists(tmp_path / "binningE.dat") and os.path.exists(tmp_path / "obs.dat")


def test_ising2d_mpi_script_single_gpu(tmp_path):
    env = dict(os.environ, PYMC_EQ_SWEEPS="200", PYMC_MC_SWEEPS="400")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "ising2D_mpi.py"), "8", "1"],
                         cwd=tmp_path, capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = out.stdout.strip().splitlines()
    assert len(lines) == 8
    e = [float(ln.split("\t")[1]) for ln in lines]
    assert e[0] == -2.0 and e[-1] > -0.7
    for name in ("rank_to_T.pickle", "kTs.npy", "0/obs_save.npy", "7/Trank_hist.npy"):
        assert os.path.exists(tmp_path / name)
