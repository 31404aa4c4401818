"""GPU probe (not a pytest file): the row-streaming kernel K13 against the tiled per-sweep kernel K1f on single-sweep
calls -- ms per sweep over lattice sizes, pieces per replica and batches; parity of the two at 16384^2."""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
TC = 2.269185314213022


def make(L0, L1, nrep, kernel, env=None, model=0):
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D
    env = env or {}
    for k, v in env.items():
        os.environ[k] = str(v)
    try:
        lat = _cabi.Lattices(model, L0, L1, nreplicas=nrep, ntemps=1, seed=5, kernel=kernel)
    finally:
        for k in env:
            os.environ.pop(k, None)
    lat.init_random()
    lat.set_accept_tables(ising2D(-1.0)._pmc_thresholds(TC)[None, :])
    return lat


def time_single(lat, n):
    for _ in range(10):
        lat.sweep(1)
    lat.synchronize()
    lat.timer_start()
    for _ in range(n):
        lat.sweep(1)
    return lat.timer_stop() / n


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/rowstream_probe.json")
    ap.add_argument("--n", type=int, default=200)
    ap.add_argument("--profile", action="store_true", help="a few single-sweep calls at 16384^2 only (for ncu)")
    args = ap.parse_args()
    from py_mc_b200 import _cabi
    if args.profile:
        a = make(16384, 16384, 1, _cabi.KERNEL_ROWSTREAM)
        print("ms per sweep", time_single(a, 20))
        a.close()
        return
    out = {"single": {}, "pieces": {}, "batch": {}}
    for L in (16384, 8192, 4096, 2048):
        row = {}
        a = make(L, L, 1, _cabi.KERNEL_ROWSTREAM)
        row["rowstream_ms"] = time_single(a, args.n)
        row["kernel"] = a.sweep_kernel_info()[1]
        b = make(L, L, 1, _cabi.KERNEL_AUTO, {"PYMC_NO_ROWSTREAM": 1})
        row["tiled_ms"] = time_single(b, args.n)
        row["tiled_kernel"] = "sweep_fused_i8_kernel"
        if L == 16384:  # same seed, same number of single-sweep calls: the two must hold the same lattice
            row["parity"] = bool(np.array_equal(a.download(0), b.download(0)))
            ta, tb = a.tracked(), b.tracked()
            row["tracked_equal"] = bool(ta[0][0] == tb[0][0] and ta[1][0] == tb[1][0])
        row["attempts_per_s"] = L * L / (row["rowstream_ms"] * 1e-3)
        row["roofline_frac"] = 2.0 * L * L / (row["rowstream_ms"] * 1e-3) / 1e9 / 6548.5
        a.close()
        b.close()
        out["single"][str(L)] = row
        print(L, row, flush=True)
    for P in (74, 148, 222, 296, 444, 592, 888):
        a = make(16384, 16384, 1, _cabi.KERNEL_ROWSTREAM, {"PYMC_ROWSTREAM_PIECES": P})
        out["pieces"][str(P)] = time_single(a, args.n)
        a.close()
        print("pieces", P, out["pieces"][str(P)], flush=True)
    for R, L in ((256, 1024), (32, 1024), (64, 2048), (16, 4096)):
        a = make(L, L, R, _cabi.KERNEL_ROWSTREAM)
        b = make(L, L, R, _cabi.KERNEL_AUTO, {"PYMC_NO_ROWSTREAM": 1})
        out["batch"]["%dx%d" % (R, L)] = {"rowstream_ms": time_single(a, args.n), "tiled_ms": time_single(b, args.n),
                                           "tiled_kernel": b.sweep_kernel_info()[1]}
        a.close()
        b.close()
        print(R, L, out["batch"]["%dx%d" % (R, L)], flush=True)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(out, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
