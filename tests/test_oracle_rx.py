"""CPU: the replica-exchange oracle (oracle/rx.py + oracle/ising.py) reproduces what the unmodified
TemperatureRX_MPI produced under the fake mpi4py with injected draws (tests/golden/rx.npz)."""
import numpy as np

from oracle import ising as o_ising
from oracle import rx as o_rx


def replay(case):
    R, lenX, lenY, J, seed = case["R"], case["lenX"], case["lenY"], case["J"], case["seed"]
    kTs = np.linspace(case["kT_lo"], case["kT_hi"], R)
    betas = 1.0 / kTs
    cfgs = [o_ising.init_config(seed, r, lenX, lenY) for r in range(R)]
    B = [o_ising.bond_sum(c) for c in cfgs]
    rank_to_T = np.arange(R)
    thr = [o_ising.threshold_table(J, kT) for kT in kTs]
    scheme, rnd = 0, 0
    obs, trank = [], []
    for s in range(1, case["nsweeps"] + 1):
        for r in range(R):
            _, dB, _ = o_ising.sweep(cfgs[r], J, kTs[rank_to_T[r]], seed, r, s - 1, thr[rank_to_T[r]])
            B[r] += dB
        if s % case["rx_every"] == 0:
            o_rx.xtrial(rank_to_T, [J * b for b in B], betas, scheme, seed, rnd)
            scheme ^= 1
            rnd += 1
        obs.append([[J * B[r], abs(o_ising.magnetization(cfgs[r]))] for r in range(R)])
        trank.append(rank_to_T.copy())
    return cfgs, rank_to_T, np.array(obs), np.array(trank), kTs


def test_rx_replay_matches_reference(golden_rx):
    z, meta = golden_rx
    for case in meta:
        n = case["name"]
        cfgs, rank_to_T, obs, trank, kTs = replay(case)
        assert np.array_equal(rank_to_T, z[n + "_rank_to_T"])
        assert np.array_equal(np.array(cfgs), z[n + "_configs"])
        assert np.array_equal(obs.transpose(1, 0, 2), z[n + "_obs_save"])      # [rank, sample, obs], bit-exact
        assert np.array_equal(trank.T, z[n + "_Trank_hist"])
        assert np.array_equal(kTs[trank.T], z[n + "_kT_hist"])
        # per-temperature means as TemperatureRX_MPI.run returns them (mc_mpi.py:172,200-209)
        R, ns = case["R"], obs.shape[0]
        acc = np.zeros((R, R, 2))  # [rank, T, obs]
        for s in range(ns):
            for r in range(R):
                acc[r, trank[s, r]] += obs[s, r]
        acc /= ns
        total = np.zeros((R, 2))
        for r in range(R):  # Allreduce in rank order
            total = total + acc[r]
        assert np.array_equal(total, z[n + "_obs_list"])
        assert case["files"] == ["Trank_hist.npy", "calc.pickle", "kT_hist.npy", "obs.dat", "obs_save.npy"]
