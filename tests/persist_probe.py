"""GPU probe (not a pytest file): per-sweep time of the on-chip kernels against the per-sweep kernels over lattice
sizes, sweeps per call and temperatures.  Used to place the automatic thresholds and for profiles/.

    python tests/persist_probe.py [--out gpurun_out/persist_probe.json]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from py_mc_b200 import _cabi  # noqa: E402
from py_mc_b200.ising2D import ising2D  # noqa: E402

TC = 2.269185314213022


def timed(lat, nsweeps, calls, warm=2):
    for _ in range(warm):
        lat.sweep(nsweeps)
    lat.synchronize()
    lat.timer_start()
    for _ in range(calls):
        lat.sweep(nsweeps)
    return lat.timer_stop() / (calls * nsweeps)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "persist_probe.json"))
    ap.add_argument("--sizes", default="1024,2048,4096,8192,16384")
    args = ap.parse_args()
    res = {"sizes": {}, "kT": {}, "sweeps_per_call": {}}
    t0 = time.time()
    model = ising2D(-1.0)
    for L in [int(x) for x in args.sizes.split(",")]:
        row = {}
        for layout, lname in ((_cabi.LAYOUT_INT8, "int8"), (_cabi.LAYOUT_BITS, "bits")):
            for kname, kernel in (("persistent", _cabi.KERNEL_PERSISTENT), ("per_sweep", _cabi.KERNEL_AUTO)):
                try:
                    lat = _cabi.Lattices(_cabi.MODEL_ISING, L, L, seed=1, layout=layout, kernel=kernel)
                    lat.init_random()
                    lat.set_accept_tables(model._pmc_thresholds(TC)[None, :])
                    if kname == "per_sweep":
                        ms = timed(lat, 1, 200)  # single-sweep calls never take the on-chip kernels
                    else:
                        ms = timed(lat, 200, 2)
                    row["%s_%s" % (lname, kname)] = {"ms_per_sweep": ms, "attempts_per_s": L * L / (ms * 1e-3),
                                                    "kernel": lat.sweep_kernel_info()[1]}
                    lat.close()
                except Exception as exc:
                    row["%s_%s" % (lname, kname)] = {"error": repr(exc)}
        res["sizes"][str(L)] = row
        print(L, json.dumps(row), flush=True)
    # sweeps per call at the headline size: fixed cost of one launch (load + pack, unpack + store)
    L = 16384
    lat = _cabi.Lattices(_cabi.MODEL_ISING, L, L, seed=1, kernel=_cabi.KERNEL_PERSISTENT)
    lat.init_random()
    lat.set_accept_tables(model._pmc_thresholds(TC)[None, :])
    for n in (2, 4, 8, 16, 64, 256, 1000):
        ms = timed(lat, n, max(2, 400 // n))
        res["sweeps_per_call"][str(n)] = {"ms_per_sweep": ms, "ms_per_call": ms * n}
        print("sweeps/call", n, ms, flush=True)
    # SURVEY 8d input set: kT in {2.0, Tc, 5.0}, >= 1000 sweeps after 100 warm-up sweeps
    for kT in (2.0, TC, 5.0):
        lat.set_accept_tables(model._pmc_thresholds(kT)[None, :])
        lat.sweep(100)
        lat.synchronize()
        lat.timer_start()
        lat.sweep(1000)
        ms = lat.timer_stop() / 1000
        b, m = lat.tracked()
        res["kT"]["%.6f" % kT] = {"ms_per_sweep": ms, "attempts_per_s": L * L / (ms * 1e-3),
                                  "energy_per_site": -float(b[0]) / (L * L), "abs_m_per_site": abs(float(m[0])) / (L * L)}
        print("kT", kT, res["kT"]["%.6f" % kT], flush=True)
    lat.close()
    res["seconds"] = time.time() - t0
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(res, open(args.out, "w"), indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
