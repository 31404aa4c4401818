#!/usr/bin/env python
"""Temperature replica exchange of the 2D Ising model -- the experiment of the reference's
``examples/ising2D_mpi.py`` (lines 13-41) on GPUs.

    one GPU :  python examples/ising2D_mpi.py 16 1
    N GPUs  :  python -m torch.distributed.run --nproc-per-node N examples/ising2D_mpi.py 64 1

argv[1] = number of replicas (= temperatures), argv[2] is accepted for compatibility.  The reference needs
one MPI rank per replica; here the replicas are split over the GPUs that were started.
"""
import copy
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from py_mc_b200.ising2D import ising2D, ising2D_config, observer
from py_mc_b200.mc import CanonicalMonteCarlo
from py_mc_b200.mc_mpi import RX_MPI_init, TemperatureRX_MPI

if __name__ == "__main__":
    comm, nreplicas, _ = RX_MPI_init()
    size = int(os.environ.get("PYMC_SIZE", "10"))
    nspin = size * size
    eq_sweeps = int(os.environ.get("PYMC_EQ_SWEEPS", "2000"))
    mc_sweeps = int(os.environ.get("PYMC_MC_SWEEPS", "8000"))
    model = ising2D(-1.0)
    base = ising2D_config(size, size)
    base.prepare_random()
    configs = [copy.deepcopy(base) for _ in range(nreplicas)]
    kTs = np.linspace(0.01, 5.0, nreplicas)
    rx = TemperatureRX_MPI(comm, CanonicalMonteCarlo, model, configs, kTs)
    rx.run(eq_sweeps, 1)
    means = rx.run(mc_sweeps, 1, 1, observer=observer())
    if comm.Get_rank() == 0:
        for kT, row in zip(kTs, means):
            print(kT, "\t", "\t".join(str(v / nspin) for v in row))
