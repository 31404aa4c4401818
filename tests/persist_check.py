"""One-shot GPU check + timing of the persistent kernel (parity on small shapes, 16384^2 against the fused kernel,
then the time per sweep).  Run on the GPU box: python tests/persist_check.py"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from py_mc_b200 import _cabi  # noqa: E402
from py_mc_b200.ising2D import ising2D  # noqa: E402
import test_zy_gpu_persist as T  # noqa: E402

res = {}
t0 = time.time()
try:
    thr = ising2D(-1.0)._pmc_thresholds(2.269185314213022)
    for shape in ((300, 256), (64, 128)):
        for layout in (0, 1):
            T.check_pair(T.run_pair(_cabi.MODEL_ISING, shape, layout, thr))
    res["small_parity"] = "ok"
    L = 16384
    lats = {}
    for name, kernel in (("persistent", _cabi.KERNEL_PERSISTENT), ("fused", _cabi.KERNEL_AUTO)):
        lat = _cabi.Lattices(_cabi.MODEL_ISING, L, L, seed=20260611, kernel=kernel)
        lat.init_random()
        lat.set_accept_tables(thr[None, :])
        lat.sweep(8)
        lat.synchronize()
        lats[name] = lat
    res["kernels"] = [lats[k].sweep_kernel_info()[1] for k in ("persistent", "fused")]
    a, b = lats["persistent"].download(0), lats["fused"].download(0)
    res["big_parity"] = bool(np.array_equal(a, b))
    res["tracked_equal"] = bool(all(np.array_equal(x, y) for x, y in zip(lats["persistent"].tracked(), lats["fused"].tracked())))
    del a, b
    for name in ("persistent", "fused"):
        lat = lats[name]
        lat.sweep(50)
        lat.synchronize()
        lat.timer_start()
        lat.sweep(400)
        ms = lat.timer_stop()
        lat.synchronize()
        res[name + "_ms_per_sweep"] = ms / 400
        res[name + "_attempts_per_s"] = L * L * 400 / (ms * 1e-3)
    res["still_equal"] = bool(np.array_equal(lats["persistent"].download(0), lats["fused"].download(0)))
except Exception as exc:  # report whatever happened
    res["error"] = repr(exc)
res["seconds"] = time.time() - t0
print(json.dumps(res))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "persist_check.json"), "w"))
