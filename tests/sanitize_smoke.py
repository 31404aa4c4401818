#!/usr/bin/env python
"""Tiny run of every kernel family, meant to be executed under compute-sanitizer (one tool per run):

    compute-sanitizer --tool memcheck  python tests/sanitize_smoke.py
    compute-sanitizer --tool racecheck python tests/sanitize_smoke.py

Not collected by pytest.  Results are also checked against the oracle so that a sanitizer-clean but wrong
run cannot pass.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ising as o_ising, latgas as o_latgas, pairmove as o_pair  # noqa: E402
from py_mc_b200 import _cabi  # noqa: E402


def main():
    seed, kT = 3, 2.3
    thr = o_ising.threshold_table(-1.0, kT)
    # fused int8 (one tile), two-pass tiled, bit-packed rows, generic word kernel
    for shape, layout, kernel in (((32, 1024), 0, _cabi.KERNEL_AUTO), ((64, 256), 0, _cabi.KERNEL_AUTO),
                                  ((16, 256), 1, _cabi.KERNEL_AUTO), ((10, 12), 0, _cabi.KERNEL_AUTO)):
        lat = _cabi.Lattices(_cabi.MODEL_ISING, shape[0], shape[1], nreplicas=2, ntemps=1, seed=seed, layout=layout,
                             kernel=kernel)
        lat.init_random()
        lat.set_accept_tables(thr[1][None, :])
        lat.sweep(2)
        ref = o_ising.init_config(seed, 1, *shape)
        for sw in range(2):
            o_ising.sweep(ref, -1.0, kT, seed, 1, sw, thr)
        assert np.array_equal(lat.download(1), ref), (shape, layout)
        b, m = lat.reduce_observables()
        assert int(b[1]) == o_ising.bond_sum(ref)
        print("ok", lat.sweep_kernel_info()[1], shape, flush=True)
        lat.close()
    # slabs in one process (peer stores into the neighbour slab's halo rows)
    slabs = [_cabi.Lattices(_cabi.MODEL_ISING, 64, 1024, seed=seed, layout=0, slab=(g, 2)) for g in range(2)]
    for g, s in enumerate(slabs):
        s.set_stream(0)
        s.slab_connect_local(slabs[1 - g], slabs[1 - g])
        s.init_random()
        s.set_accept_tables(thr[1][None, :])
    for _ in range(2):
        for s in slabs:
            s.sweep(1)
    ref = o_ising.init_config(seed, 0, 64, 1024)
    for sw in range(2):
        o_ising.sweep(ref, -1.0, kT, seed, 0, sw, thr)
    assert np.array_equal(np.concatenate([s.download(0) for s in slabs]), ref)
    print("ok slabs", flush=True)
    # lattice gas flip + pair exchange, both layouts
    for shape, layout in (((8, 12), 0), ((8, 128), 1)):
        lat = _cabi.Lattices(_cabi.MODEL_LATGAS, shape[0], shape[1], seed=seed, layout=layout)
        lat.init_random()
        tf = o_latgas.threshold_table(-1.0, -0.5, 0.25, 0.9)
        tp = o_pair.threshold_table(-1.0, 0.9)
        lat.set_accept_tables(tf[None, :])
        lat.set_pair_tables(tp[None, :])
        ref = o_latgas.init_config(seed, 0, *shape)
        lat.sweep(1)
        o_latgas.sweep(ref, tf, seed, 0, 0)
        lat.pair_sweep(1)
        o_pair.sweep_classes(ref, tp, seed, 0, 0)
        assert np.array_equal(lat.download(0), ref), (shape, layout)
        print("ok latgas + pair", shape, flush=True)
        lat.close()
    # replica exchange and statistics
    R = 4
    kTs = np.linspace(1.5, 3.0, R)
    lat = _cabi.Lattices(_cabi.MODEL_ISING, 8, 8, nreplicas=R, ntemps=R, seed=seed)
    lat.init_random()
    lat.set_accept_tables(np.array([o_ising.threshold_table(-1.0, t)[1] for t in kTs]))
    lat.rx_setup(1.0 / kTs, -1.0)
    lat.rx_cycle(4, 1, 0)
    lat.stats_reset(-1.0)
    lat.stats_accumulate()
    lat.stats_get()
    lat.reduce_observables_static()
    print("ok rx + stats", lat.get_replica_to_T(), flush=True)
    # hard discs
    d = _cabi.Discs(3, 20, 1.0, 1.0 / 14.0, 1.0 / 28.0, seed=seed)
    from oracle import harddisc as o_hd
    c = o_hd.prepare_ordered(1.0, 20, 1.0 / 14.0)
    for r in range(3):
        d.upload(r, c)
    d.run(50)
    d.observe(0.01, 49)
    for t in range(50):
        o_hd.trial(c, 1.0, 1.0 / 14.0, 1.0 / 28.0, seed, 0, t)
    assert np.array_equal(d.download(0), c)
    print("ok discs", flush=True)
    print("SANITIZE_SMOKE PASS", flush=True)


if __name__ == "__main__":
    main()
