"""GPU probe (not a pytest file): ms per sweep of a single L x L lattice on the batch-resident kernel, 200 sweeps per call."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from py_mc_b200 import _cabi  # noqa: E402
from py_mc_b200.ising2D import ising2D  # noqa: E402

L = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
lat = _cabi.Lattices(_cabi.MODEL_ISING, L, L, seed=1)
lat.init_random()
lat.set_accept_tables(ising2D(-1.0)._pmc_thresholds(2.269185314213022)[None, :])
lat.sweep(200)
lat.synchronize()
best = 1e9
for _ in range(3):
    lat.timer_start()
    lat.sweep(200)
    best = min(best, lat.timer_stop() / 200)
print("L", L, "us per sweep %.3f" % (1e3 * best), lat.sweep_kernel_info()[1])
