"""py_mc_b200 -- B200-native Metropolis sweep / replica-exchange engine behind py_mc's plug-in API.

Drop-in modules (same names, signatures and return values as the reference):

* ``py_mc_b200.mc``       <- ``py_mc/mc.py``      (model, CanonicalMonteCarlo, observer_base, binning, ...)
* ``py_mc_b200.mc_mpi``   <- ``py_mc/mc_mpi.py``  (ParallelMC, TemperatureRX_MPI; NCCL instead of MPI)
* ``py_mc_b200.ising2D``  <- ``examples/ising2D.py`` (ising2D, ising2D_config, observer)
* ``py_mc_b200.latgas``   <- ``examples/latgas.py`` / ``latgas_move.py``

All arithmetic on configurations runs in hand-written sm_100a CUDA kernels reached through the C ABI
of ``include/pymc_b200.h`` (ctypes).  There is no CPU fallback: importing the engine classes without
the built library raises.
"""
from .seeding import set_seed, get_seed  # noqa: F401

__version__ = "0.1.0"
