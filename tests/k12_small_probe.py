"""GPU probe (not a pytest file): us per sweep of R replicas of 1024 x 1024 on the batch-resident kernel, 400 sweeps per call."""
import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np
from py_mc_b200 import _cabi
from py_mc_b200.ising2D import ising2D
R = int(sys.argv[1]) if len(sys.argv) > 1 else 32
m = ising2D(-1.0)
kTs = np.linspace(2.0, 2.6, R)
lat = _cabi.Lattices(_cabi.MODEL_ISING, 1024, 1024, nreplicas=R, ntemps=R, seed=2)
lat.init_random()
lat.set_accept_tables(np.array([m._pmc_thresholds(k) for k in kTs]))
lat.sweep(200); lat.synchronize()
lat.timer_start(); lat.sweep(400); ms = lat.timer_stop()
print("R", R, "us per sweep", 1e3 * ms / 400, lat.sweep_kernel_info())
