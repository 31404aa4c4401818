#!/usr/bin/env python
"""Multi-GPU check of the row-slab decomposition, run under torchrun on N GPUs (not collected by pytest):

    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29561 tests/mgpu_slab_check.py

One lattice is split over the ranks; halo rows travel as peer stores from inside the sweep kernel.  Rank 0 also
runs the whole lattice on its own GPU and the gathered slabs must equal it bit for bit.  Prints
MGPU_SLAB_CHECK PASS/FAIL and, when asked with --bench, the strong-scaling rate.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D
    from py_mc_b200.slab import SlabLattice
    ok = True
    thr = ising2D(-1.0)._pmc_thresholds(2.269185314213022)[None, :]
    # calls: single sweeps take the tiled kernel (halo rows stored into the neighbours' arrays), calls of two or more sweeps
    # the batch-resident kernel (first / last rows to the neighbouring GPUs as packets); both, and mixtures, must give the
    # whole-lattice trajectory
    for (lenX, lenY, calls, seed) in [(256, 1024, (9,), 11), (1024, 2048, (5,), 12), (512, 1024, (1, 1, 3, 1, 2, 1), 13),
                                       (2048, 4096, (2, 4), 14)]:
        if lenX % (32 * world):
            continue
        nsweeps = sum(calls)
        sl = SlabLattice(_cabi.MODEL_ISING, lenX, lenY, seed, dist)
        sl.init_random()
        sl.set_accept_tables(thr)
        for n in calls:
            sl.sweep(n)
        sl.check()
        got = sl.gather()
        b, m = sl.observables()
        good = True
        if rank == 0:
            whole = _cabi.Lattices(_cabi.MODEL_ISING, lenX, lenY, seed=seed, layout=_cabi.LAYOUT_INT8, device=local)
            whole.init_random()
            whole.set_accept_tables(thr)
            whole.sweep(nsweeps)
            want = whole.download(0)
            wb, wm = whole.tracked()
            good = np.array_equal(got, want) and (b, m) == (int(wb[0]), int(wm[0]))
            # upload path: continue from the downloaded state on the slabs
            print("slab %dx%d over %d GPUs, calls %s (%s / %s): %s" % (lenX, lenY, world, calls, sl.lat.sweep_kernel_for(1)[1],
                                                                     sl.lat.sweep_kernel_for(2)[1], "ok" if good else "MISMATCH"),
                  flush=True)
        flag = torch.tensor([1 if good else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok &= bool(flag.item())
        # upload + halo push + more sweeps must continue the same trajectory
        rows = got[rank * sl.rows:(rank + 1) * sl.rows].copy()
        sl2 = SlabLattice(_cabi.MODEL_ISING, lenX, lenY, seed, dist)
        sl2.set_accept_tables(thr)
        sl2.upload_rows(rows)
        sl2.lat.sweep_counter = nsweeps
        sl.sweep(3)
        sl2.sweep(3)
        sl.check()
        sl2.check()
        same = np.array_equal(sl.gather(), sl2.gather())
        flag = torch.tensor([1 if same else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if rank == 0:
            print("   upload + halo push continues the trajectory: %s" % ("ok" if flag.item() else "MISMATCH"), flush=True)
        ok &= bool(flag.item())
        del sl, sl2
    if "--bench" in sys.argv:
        L = 16384
        sl = SlabLattice(_cabi.MODEL_ISING, L, L, 20260611, dist)
        sl.init_random()
        sl.set_accept_tables(thr)
        sl.sweep(50)
        sl.lat.synchronize()
        dist.barrier()
        sl.lat.timer_start()
        sl.sweep(500)
        ms = sl.lat.timer_stop()
        sl.check()
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print("SLAB_BENCH gpus=%d attempts_per_s=%.4g ms_per_sweep=%.4f" % (world, L * L * 500 / (t.item() * 1e-3),
                                                                                 t.item() / 500), flush=True)
    if rank == 0:
        print("MGPU_SLAB_CHECK", "PASS" if ok else "FAIL", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
