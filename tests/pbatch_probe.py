"""GPU probe (not a pytest file): time per sweep of the batch-resident kernel K12 against the other kernels on single
lattices, replica-exchange batches and the lattice-gas batch.  For profiles/ and for placing the automatic choice.

    python tests/pbatch_probe.py [--out gpurun_out/pbatch_probe.json]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from py_mc_b200 import _cabi  # noqa: E402
from py_mc_b200.ising2D import ising2D  # noqa: E402

TC = 2.269185314213022
K = {"pbatch": _cabi.KERNEL_PBATCH, "persistent": _cabi.KERNEL_PERSISTENT, "resident": _cabi.KERNEL_RESIDENT,
     "per_sweep": _cabi.KERNEL_AUTO}


def timed(fn, n_per_call, calls, lat):
    fn()
    lat.synchronize()
    lat.timer_start()
    for _ in range(calls):
        fn()
    return lat.timer_stop() / (calls * n_per_call)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "pbatch_probe.json"))
    args = ap.parse_args()
    model = ising2D(-1.0)
    res = {"single": {}, "rx": {}, "latgas": {}}
    t0 = time.time()
    for L in (128, 256, 512, 1024, 2048, 4096, 8192, 16384):
        row = {}
        for layout, lname in ((0, "int8"), (1, "bits")):
            for kname in ("pbatch", "persistent", "per_sweep"):
                try:
                    lat = _cabi.Lattices(_cabi.MODEL_ISING, L, L, seed=1, layout=layout, kernel=K[kname])
                    lat.init_random()
                    lat.set_accept_tables(model._pmc_thresholds(TC)[None, :])
                    if kname == "per_sweep":
                        os.environ["PYMC_NO_PERSISTENT"] = "1"
                        lat.close()
                        lat = _cabi.Lattices(_cabi.MODEL_ISING, L, L, seed=1, layout=layout, kernel=K[kname])
                        del os.environ["PYMC_NO_PERSISTENT"]
                        lat.init_random()
                        lat.set_accept_tables(model._pmc_thresholds(TC)[None, :])
                        ms = timed(lambda: lat.sweep(1), 1, 200, lat)
                    else:
                        ms = timed(lambda: lat.sweep(200), 200, 2, lat)
                    row["%s_%s" % (lname, kname)] = {"ms_per_sweep": ms, "attempts_per_s": L * L / (ms * 1e-3),
                                                    "kernel": lat.sweep_kernel_info()[1]}
                    lat.close()
                except Exception as exc:
                    row["%s_%s" % (lname, kname)] = {"error": repr(exc)}
        res["single"][str(L)] = row
        print("single", L, json.dumps(row), flush=True)
    # replica exchange batches R x 1024^2 int8, exchange every `every` sweeps, all on one GPU
    for R in (8, 32, 64, 256):
        kTs = np.linspace(2.0, 2.6, R)
        thr = np.array([model._pmc_thresholds(kT) for kT in kTs])
        row = {}
        for kname in ("pbatch", "resident", "per_sweep"):
            for every in (1, 10):
                try:
                    if kname == "per_sweep":
                        os.environ["PYMC_NO_PERSISTENT"] = "1"
                    lat = _cabi.Lattices(_cabi.MODEL_ISING, 1024, 1024, nreplicas=R, ntemps=R, seed=2, kernel=K[kname])
                    os.environ.pop("PYMC_NO_PERSISTENT", None)
                    lat.init_random()
                    lat.set_accept_tables(thr)
                    lat.rx_setup(1.0 / kTs, *model._pmc_energy_coeffs())
                    ncyc = 200 // every
                    ms = timed(lambda: lat.rx_cycle(ncyc, every, 0), ncyc * every, 3, lat)
                    tried, acc = lat.rx_stats()
                    row["%s_every%d" % (kname, every)] = {"ms_per_sweep": ms, "attempts_per_s": R * 1024 * 1024 / (ms * 1e-3),
                                                         "kernel": lat.sweep_kernel_info()[1],
                                                         "acceptance": float(acc.sum()) / max(1.0, float(tried.sum()))}
                    lat.close()
                except Exception as exc:
                    os.environ.pop("PYMC_NO_PERSISTENT", None)
                    row["%s_every%d" % (kname, every)] = {"error": repr(exc)}
            # sweeps only (no exchange): what an exchange costs on top
            try:
                if kname == "per_sweep":
                    os.environ["PYMC_NO_PERSISTENT"] = "1"
                lat = _cabi.Lattices(_cabi.MODEL_ISING, 1024, 1024, nreplicas=R, ntemps=R, seed=2, kernel=K[kname])
                os.environ.pop("PYMC_NO_PERSISTENT", None)
                lat.init_random()
                lat.set_accept_tables(thr)
                ms = timed(lambda: lat.sweep(200), 200, 3, lat) if kname != "per_sweep" else timed(lambda: lat.sweep(1), 1, 200, lat)
                row["%s_sweeps_only" % kname] = {"ms_per_sweep": ms, "attempts_per_s": R * 1024 * 1024 / (ms * 1e-3)}
                lat.close()
            except Exception as exc:
                os.environ.pop("PYMC_NO_PERSISTENT", None)
                row["%s_sweeps_only" % kname] = {"error": repr(exc)}
        res["rx"][str(R)] = row
        print("rx", R, json.dumps(row), flush=True)
    # lattice gas 512 x 256^2, flip sweeps
    from py_mc_b200.latgas_move import latticegas
    lg = latticegas(Eads=0.0, J=-1.0, mu=0.0)
    for kname in ("pbatch", "resident"):
        try:
            lat = _cabi.Lattices(_cabi.MODEL_LATGAS, 256, 256, nreplicas=512, ntemps=1, seed=3, kernel=K[kname])
            lat.init_random()
            lat.set_accept_tables(lg._pmc_thresholds(1.0)[None, :])
            ms = timed(lambda: lat.sweep(200), 200, 3, lat)
            res["latgas"][kname] = {"ms_per_sweep": ms, "attempts_per_s": 512 * 256 * 256 / (ms * 1e-3)}
            lat.close()
        except Exception as exc:
            res["latgas"][kname] = {"error": repr(exc)}
    print("latgas", json.dumps(res["latgas"]), flush=True)
    res["seconds"] = time.time() - t0
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(res, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
