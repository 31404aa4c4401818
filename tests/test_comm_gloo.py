"""CPU, world_size 2 over gloo: the host-side plumbing of the multi-GPU replica-exchange path --
block partition of replicas, the energy all-gather (rank order) and the closing all-reduce -- and the
property the design rests on: every process that sees the same gathered energies derives the same
new replica->temperature map (checked here with the oracle's decision rule; the device kernel is
checked against that rule in tests/test_gpu_rx.py)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from py_mc_b200.mc_mpi import ReplicaComm
    from oracle import rx as o_rx
    comm = ReplicaComm(dist)
    R = 8
    first, n = comm.partition(R)
    # each process "owns" the energies of its block only
    rng = np.random.default_rng(5)
    energies_all = -rng.integers(0, 200, size=R).astype(np.float64)
    local = energies_all[first:first + n]
    gathered = comm.allgather_f64_host(local)
    kTs = np.linspace(1.0, 3.0, R)
    maps = []
    m = np.arange(R)
    for rnd in range(6):
        o_rx.xtrial(m, gathered, 1.0 / kTs, rnd & 1, 99, rnd)
        maps.append(m.copy())
    # per-temperature accumulators: each process fills only the rows of its own replicas
    acc = np.zeros((R, 2))
    for r in range(first, first + n):
        acc[m[r]] += [energies_all[r], float(r)]
    total = comm.allreduce_sum_host(acc)
    try:
        comm.partition(7)
        bad = False
    except ValueError:
        bad = True
    q.put((rank, first, n, gathered, np.array(maps), total, bad))
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_partition_gather_reduce_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=240) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
    (r0, f0, n0, g0, m0, t0, b0), (r1, f1, n1, g1, m1, t1, b1) = res
    assert (f0, n0, f1, n1) == (0, 4, 4, 4)
    rng = np.random.default_rng(5)
    expect = -rng.integers(0, 200, size=8).astype(np.float64)
    assert np.array_equal(g0, expect) and np.array_equal(g1, expect)  # rank order, both processes
    assert np.array_equal(m0, m1)                                      # identical maps, no second message
    assert np.array_equal(t0, t1)
    final = m0[-1]
    full = np.zeros((8, 2))
    for r in range(8):
        full[final[r]] += [expect[r], float(r)]
    assert np.array_equal(t0, full)
    assert b0 and b1
