#!/usr/bin/env python
"""Multi-GPU parity check, run under torchrun on N GPUs (not collected by pytest):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tests/mgpu_rx_check.py

Runs the replica-exchange fixture cases with the replicas split over the ranks (NCCL all-gather of
the energies, decision kernel replicated) and compares every rank's share with the fixture recorded
from the unmodified reference.  Prints MGPU_RX_CHECK PASS/FAIL on rank 0; exit code != 0 on failure.
"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import py_mc_b200
    from conftest import load_golden
    from py_mc_b200.ising2D import ising2D, ising2D_config, observer
    from py_mc_b200.mc import CanonicalMonteCarlo
    from py_mc_b200.mc_mpi import ReplicaComm, TemperatureRX_MPI
    from oracle import ising as o_ising
    z, meta = load_golden("rx")
    ok = True
    work = [tempfile.mkdtemp() if rank == 0 else None]
    dist.broadcast_object_list(work, src=0)
    os.chdir(work[0])
    for case in meta:
        n, R = case["name"], case["R"]
        if R % world:
            continue
        os.makedirs(n, exist_ok=True)
        dist.barrier()
        os.chdir(os.path.join(work[0], n))
        py_mc_b200.set_seed(case["seed"])
        kTs = np.linspace(case["kT_lo"], case["kT_hi"], R)
        configs = []
        for r in range(R):
            cfg = ising2D_config(case["lenX"], case["lenY"])
            cfg.config[:, :] = o_ising.init_config(case["seed"], r, case["lenX"], case["lenY"])
            configs.append(cfg)
        rx = TemperatureRX_MPI(ReplicaComm(dist), CanonicalMonteCarlo, ising2D(case["J"]), configs, kTs)
        out = rx.run(case["nsweeps"], case["rx_every"], 1, 10 ** 12, observer(), True, True)
        good = np.array_equal(rx.rank_to_T, z[n + "_rank_to_T"])
        good &= np.allclose(np.array(out, dtype=np.float64), z[n + "_obs_list"], rtol=1e-13, atol=0)
        for loc in range(rx.nlocal):
            r = rx.first + loc
            good &= np.array_equal(configs[r].config.astype(np.int8), z[n + "_configs"][r])
            good &= np.array_equal(np.load(os.path.join(str(r), "obs_save.npy")), z[n + "_obs_save"][r])
            good &= np.array_equal(np.load(os.path.join(str(r), "Trank_hist.npy")), z[n + "_Trank_hist"][r])
        flag = torch.tensor([1 if good else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if rank == 0:
            print("case %s on %d GPUs: %s" % (n, world, "ok" if flag.item() else "MISMATCH"), flush=True)
        ok &= bool(flag.item())
        os.chdir(work[0])
    # device-side observer and ParallelMC on the partitioned replicas: results must not depend on the
    # number of GPUs (compared with a single-process run of the same seed on rank 0)
    from py_mc_b200.mc_mpi import ParallelMC
    from py_mc_b200.observers import device_observer

    def rx_device(comm_obj, R=8):
        py_mc_b200.set_seed(1234)
        kTs = np.linspace(1.8, 3.0, R)
        cfgs = []
        for r in range(R):
            c = ising2D_config(8, 8)
            c.config[:, :] = o_ising.init_config(1234, r, 8, 8)
            cfgs.append(c)
        rx = TemperatureRX_MPI(comm_obj, CanonicalMonteCarlo, ising2D(-1.0), cfgs, kTs)
        obs = device_observer()
        out = rx.run(300, 2, 3, observer=obs)
        return np.array(out), obs.counts.copy(), rx.rank_to_T.copy()

    if world in (2, 4, 8):
        os.makedirs("devobs", exist_ok=True)
        dist.barrier()
        os.chdir(os.path.join(work[0], "devobs"))
        multi = rx_device(ReplicaComm(dist))
        good = True
        if rank == 0:
            single = rx_device(ReplicaComm(None))
            good = (np.allclose(multi[0], single[0], rtol=1e-12) and np.array_equal(multi[1], single[1])
                    and np.array_equal(multi[2], single[2]))
            print("device observer on %d GPUs vs 1 GPU: %s" % (world, "ok" if good else "MISMATCH"), flush=True)
        flag = torch.tensor([1 if good else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok &= bool(flag.item())
        os.chdir(work[0])
    if rank == 0:
        print("MGPU_RX_CHECK", "PASS" if ok else "FAIL", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
