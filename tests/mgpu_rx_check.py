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
    # exchange inside the sweep kernel through peer memory (pymc_rx_ipc_export / pymc_rx_connect): one launch per GPU
    # for all sweeps and exchanges; everything a caller can observe must equal the one-GPU run of the same ladder
    from py_mc_b200 import _cabi
    for (R, shape, every, ncyc) in ((16, (128, 256), 2, (5, 4)), (8 * world, (64, 128), 1, (7,)), (2 * world, (1024, 1024), 3, (3,))):
        if R % world:
            continue
        kTs = np.linspace(2.25, 2.25 + 0.004 * R, R)
        model = ising2D(-1.0)
        thr = np.array([model._pmc_thresholds(kT) for kT in kTs])

        def ladder(nloc, offset, group):
            lat = _cabi.Lattices(_cabi.MODEL_ISING, shape[0], shape[1], nreplicas=nloc, ntemps=R, seed=99, device=local,
                                 replica_offset=offset, nreplicas_global=R, kernel=_cabi.KERNEL_PBATCH)
            lat.init_random()
            lat.set_accept_tables(thr)
            lat.rx_setup(1.0 / kTs, *model._pmc_energy_coeffs())
            if group is not None:
                lat.rx_connect_group(group)
            scheme = 1
            for n in ncyc:
                lat.rx_cycle(n, every, scheme)
                scheme ^= n & 1
            lat.synchronize()
            res = dict(map=lat.get_replica_to_T(), energies=lat.rx_get_energies(), stats=lat.rx_stats(), tracked=lat.tracked(),
                       reduced=lat.reduce_observables(), configs=[lat.download(r) for r in range(nloc)], round=lat.rx_round)
            lat.close()
            return res

        nloc = R // world
        multi = ladder(nloc, rank * nloc, dist)
        good = all(np.array_equal(multi["tracked"][k], multi["reduced"][k]) for k in (0, 1))
        # every rank must hold the same map, energies and counters
        t = torch.tensor(multi["map"], device="cuda", dtype=torch.int32)
        lo, hi = t.clone(), t.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        good &= bool((lo == hi).all().item())
        if rank == 0:
            single = ladder(R, 0, None)
            good &= np.array_equal(multi["map"], single["map"]) and np.array_equal(multi["energies"], single["energies"])
            good &= all(np.array_equal(multi["stats"][k], single["stats"][k]) for k in (0, 1)) and multi["round"] == single["round"]
            good &= int(single["stats"][1].sum()) > 0
            for r in range(nloc):
                good &= np.array_equal(multi["configs"][r], single["configs"][r])
                good &= all(int(multi["tracked"][k][r]) == int(single["tracked"][k][r]) for k in (0, 1))
            print("exchange through peer memory, %d replicas of %dx%d on %d GPUs vs 1 GPU: %s"
                  % (R, shape[0], shape[1], world, "ok" if good else "MISMATCH"), flush=True)
        flag = torch.tensor([1 if good else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok &= bool(flag.item())
    if rank == 0:
        print("MGPU_RX_CHECK", "PASS" if ok else "FAIL", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
