"""GPU: replica exchange through the C ABI against the fixture recorded from the unmodified
TemperatureRX_MPI (tests/golden/rx.npz) -- energies, temperature histories and final
configurations are compared bit-exactly."""
import numpy as np
import pytest

from oracle import ising as o_ising
from oracle import rx as o_rx

pytestmark = pytest.mark.gpu


def _setup(case, nrep=None, offset=0):
    from py_mc_b200 import _cabi
    R = case["R"]
    kTs = np.linspace(case["kT_lo"], case["kT_hi"], R)
    thr = np.array([o_ising.threshold_table(case["J"], kT)[1] for kT in kTs])
    lat = _cabi.Lattices(_cabi.MODEL_ISING, case["lenX"], case["lenY"], nreplicas=nrep or R, ntemps=R,
                         seed=case["seed"], replica_offset=offset, nreplicas_global=R)
    lat.init_random()
    lat.set_accept_tables(thr)
    lat.rx_setup(1.0 / kTs, case["J"])
    return lat, kTs


def test_rx_matches_reference_history(golden_rx):
    z, meta = golden_rx
    for case in meta:
        n, R = case["name"], case["R"]
        lat, kTs = _setup(case)
        scheme = 0
        for s in range(1, case["nsweeps"] + 1):
            lat.sweep(1)
            if s % case["rx_every"] == 0:
                lat.rx_energies()
                lat.rx_exchange(scheme)
                scheme ^= 1
            b, m = lat.tracked()
            trank = lat.get_replica_to_T()
            assert np.array_equal(case["J"] * b.astype(np.float64), z[n + "_obs_save"][:, s - 1, 0]), (n, s)
            assert np.array_equal(np.abs(m).astype(np.float64), z[n + "_obs_save"][:, s - 1, 1]), (n, s)
            assert np.array_equal(trank, z[n + "_Trank_hist"][:, s - 1]), (n, s)
        assert np.array_equal(lat.get_replica_to_T(), z[n + "_rank_to_T"])
        for r in range(R):
            assert np.array_equal(lat.download(r), z[n + "_configs"][r]), (n, r)
        tried, acc = lat.rx_stats()
        nex = case["nsweeps"] // case["rx_every"]
        assert tried.sum() == sum(len(range(k & 1, R - 1, 2)) for k in range(nex))
        assert lat.rx_round == nex
        lat.close()


def test_rx_cycle_equals_manual_loop(golden_rx):
    z, meta = golden_rx
    for case in meta:
        n = case["name"]
        lat, _ = _setup(case)
        ncycles = case["nsweeps"] // case["rx_every"]
        lat.rx_cycle(ncycles, case["rx_every"], 0)
        assert np.array_equal(lat.get_replica_to_T(), z[n + "_rank_to_T"])
        for r in range(case["R"]):
            assert np.array_equal(lat.download(r), z[n + "_configs"][r])
        lat.close()


def test_partitioned_handles_agree_with_single_handle(golden_rx):
    """Two handles that each own half of the replicas (as two GPUs would), with the energies
    gathered by hand between them, must reproduce the single-handle run: the decision kernel
    depends only on all-gathered data."""
    z, meta = golden_rx
    case = meta[0]
    n, R = case["name"], case["R"]
    half = R // 2
    a, _ = _setup(case, nrep=half, offset=0)
    b, _ = _setup(case, nrep=R - half, offset=half)
    scheme = 0
    for s in range(1, case["nsweeps"] + 1):
        a.sweep(1)
        b.sweep(1)
        if s % case["rx_every"] == 0:
            a.rx_energies()
            b.rx_energies()
            ea, eb = a.rx_get_energies(), b.rx_get_energies()
            full = np.concatenate([ea[:half], eb[half:]])
            # stand-in for the NCCL all-gather: write the gathered vector into both handles
            import ctypes
            from py_mc_b200 import _cabi
            for lat in (a, b):
                ptr, _ = lat.rx_buffers()
                _cuda_memcpy_h2d(ptr, full)
            a.rx_exchange(scheme)
            b.rx_exchange(scheme)
            scheme ^= 1
            assert np.array_equal(a.get_replica_to_T(), b.get_replica_to_T())
    assert np.array_equal(a.get_replica_to_T(), z[n + "_rank_to_T"])
    for r in range(R):
        lat, loc = (a, r) if r < half else (b, r - half)
        assert np.array_equal(lat.download(loc), z[n + "_configs"][r])


def _cuda_memcpy_h2d(dev_ptr, host_array):
    import ctypes
    rt = ctypes.CDLL("libcudart.so")
    rt.cudaMemcpy.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
    h = np.ascontiguousarray(host_array)
    assert rt.cudaMemcpy(dev_ptr, h.ctypes.data, h.nbytes, 1) == 0


def test_exchange_acceptance_statistics():
    """Property at a size without fixture: acceptance follows min(1, exp(-delta)) when energies are
    held fixed -- compare the device decision with the oracle's over many rounds."""
    from py_mc_b200 import _cabi
    R, seed = 16, 31337
    kTs = np.linspace(1.0, 4.0, R)
    lat = _cabi.Lattices(_cabi.MODEL_ISING, 8, 8, nreplicas=R, ntemps=R, seed=seed)
    lat.init_random()
    lat.set_accept_tables(np.array([o_ising.threshold_table(-1.0, kT)[1] for kT in kTs]))
    lat.rx_setup(1.0 / kTs, -1.0)
    lat.sweep(3)
    b, _ = lat.tracked()
    E = -1.0 * b.astype(np.float64)
    ref_map = np.arange(R)
    for rnd in range(200):
        lat.rx_energies()
        lat.rx_exchange(rnd & 1)
        o_rx.xtrial(ref_map, E, 1.0 / kTs, rnd & 1, seed, rnd)
    assert np.array_equal(lat.get_replica_to_T(), ref_map)
    tried, acc = lat.rx_stats()
    assert 0 < acc.sum() < tried.sum()


def test_lattice_gas_replica_exchange_matches_oracle_replay():
    """Exchange with the two-term energy J*B + (Eads-mu)*N of the lattice gas (latgas.py:17-25): the device
    forms it without FMA contraction, so the decisions equal the host's float64 arithmetic."""
    from py_mc_b200 import _cabi
    from py_mc_b200.latgas import latticegas
    from oracle import latgas as o_latgas
    R, lenX, lenY, seed = 6, 8, 12, 2025
    J, Eads, mu = -0.75, -0.5, 0.125
    kTs = np.linspace(0.6, 1.4, R)
    model = latticegas(Eads=Eads, J=J, mu=mu)
    lat = _cabi.Lattices(_cabi.MODEL_LATGAS, lenX, lenY, nreplicas=R, ntemps=R, seed=seed)
    lat.init_random()
    lat.set_accept_tables(np.array([model._pmc_thresholds(kT) for kT in kTs]))
    lat.rx_setup(1.0 / kTs, *model._pmc_energy_coeffs())
    cfgs = [o_latgas.init_config(seed, r, lenX, lenY) for r in range(R)]
    thr = [o_latgas.threshold_table(J, Eads, mu, kT) for kT in kTs]
    rank_to_T = np.arange(R)
    scheme = 0
    for s in range(1, 25):
        lat.sweep(1)
        for r in range(R):
            o_latgas.sweep(cfgs[r], thr[rank_to_T[r]], seed, r, s - 1)
        if s % 2 == 0:
            lat.rx_energies()
            lat.rx_exchange(scheme)
            E = [o_latgas.energy(c, J, Eads, mu) for c in cfgs]
            assert np.array_equal(lat.rx_get_energies(), np.array(E))  # bit-exact float64 energies
            o_rx.xtrial(rank_to_T, E, 1.0 / kTs, scheme, seed, s // 2 - 1)
            scheme ^= 1
        assert np.array_equal(lat.get_replica_to_T(), rank_to_T), s
    for r in range(R):
        assert np.array_equal(lat.download(r), cfgs[r])
    tried, acc = lat.rx_stats()
    assert acc.sum() > 0
