"""GPU: the TemperatureRX_MPI drop-in (py_mc_b200.mc_mpi) against the fixture recorded from the
reference: returned per-temperature means, on-disk files, histories, final configurations, reload."""
import os
import pickle

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture()
def workdir(tmp_path):
    old = os.getcwd()
    os.chdir(tmp_path)
    yield tmp_path
    os.chdir(old)


def _build(case, z):
    import py_mc_b200
    from py_mc_b200.ising2D import ising2D, ising2D_config
    from py_mc_b200.mc import CanonicalMonteCarlo
    from py_mc_b200.mc_mpi import TemperatureRX_MPI
    from oracle import ising as o_ising
    py_mc_b200.set_seed(case["seed"])
    R = case["R"]
    kTs = np.linspace(case["kT_lo"], case["kT_hi"], R)
    configs = []
    for r in range(R):
        cfg = ising2D_config(case["lenX"], case["lenY"])
        cfg.config[:, :] = o_ising.init_config(case["seed"], r, case["lenX"], case["lenY"])
        configs.append(cfg)
    rx = TemperatureRX_MPI(None, CanonicalMonteCarlo, ising2D(case["J"]), configs, kTs)
    return rx, configs, kTs


def test_run_matches_reference(golden_rx, workdir):
    from py_mc_b200.ising2D import observer
    z, meta = golden_rx
    for case in meta:
        sub = workdir / case["name"]
        sub.mkdir()
        os.chdir(sub)
        n, R = case["name"], case["R"]
        rx, configs, kTs = _build(case, z)
        out = rx.run(case["nsweeps"], case["rx_every"], 1, 10 ** 12, observer(), True, True)
        assert np.array_equal(np.array(out, dtype=np.float64), z[n + "_obs_list"])  # bit-exact means per T
        assert np.array_equal(rx.rank_to_T, z[n + "_rank_to_T"])
        for r in range(R):
            assert np.array_equal(configs[r].config.astype(np.int8), z[n + "_configs"][r])
            assert sorted(os.listdir(str(r))) == case["files"]
            assert np.array_equal(np.load(os.path.join(str(r), "obs_save.npy")), z[n + "_obs_save"][r])
            assert np.array_equal(np.load(os.path.join(str(r), "Trank_hist.npy")), z[n + "_Trank_hist"][r])
            assert np.array_equal(np.load(os.path.join(str(r), "kT_hist.npy")), z[n + "_kT_hist"][r])
            with open(os.path.join(str(r), "calc.pickle"), "rb") as f:
                cfg = pickle.load(f)
            assert np.array_equal(cfg.config.astype(np.int8), z[n + "_configs"][r])
        for name in case["top_files"]:
            assert os.path.exists(name)
        with open("rank_to_T.pickle", "rb") as f:
            assert np.array_equal(pickle.load(f), z[n + "_rank_to_T"])
        assert np.array_equal(np.load("kTs.npy"), kTs)
        assert rx.states[0].energy == z[n + "_energy"][0]


def test_split_run_with_reload_equals_one_run(golden_rx, workdir):
    """run(a) ; new object ; reload() ; run(b)  ==  run(a+b): restart is bit-reproducible because the
    random-stream position travels with the checkpoint."""
    from py_mc_b200.ising2D import observer
    z, meta = golden_rx
    case = meta[0]
    n, R = case["name"], case["R"]
    first = 16  # multiple of rx_every, even number of exchanges so the scheme parity restarts at 0
    rx, configs, kTs = _build(case, z)
    rx.run(first, case["rx_every"], 1, 10 ** 12, observer(), True, True)
    rx2, configs2, _ = _build(case, z)
    rx2.reload()
    rx2.run(case["nsweeps"] - first, case["rx_every"], 1, 10 ** 12, observer(), True, True)
    assert np.array_equal(rx2.rank_to_T, z[n + "_rank_to_T"])
    for r in range(R):
        assert np.array_equal(rx2.configs[r].config.astype(np.int8), z[n + "_configs"][r])
        assert np.array_equal(np.load(os.path.join(str(r), "obs_save.npy")), z[n + "_obs_save"][r])
        assert np.array_equal(np.load(os.path.join(str(r), "Trank_hist.npy")), z[n + "_Trank_hist"][r])


def test_parallel_mc_returns_per_replica_means(workdir):
    import py_mc_b200
    from py_mc_b200.ising2D import ising2D, ising2D_config, observer
    from py_mc_b200.mc import CanonicalMonteCarlo
    from py_mc_b200.mc_mpi import ParallelMC
    py_mc_b200.set_seed(3)
    kTs = [1.5, 2.5, 3.5]
    configs = []
    for _ in kTs:
        c = ising2D_config(8, 8)
        c.prepare_random()
        configs.append(c)
    pmc = ParallelMC(None, CanonicalMonteCarlo, ising2D(-1.0), configs, kTs)
    out = pmc.run(200, 2, observer())
    assert out.shape == (3, 2)
    assert out[0, 0] < out[2, 0] < 0  # colder replica has lower energy
    assert os.path.exists("0/config.pickle") and os.path.exists("2/obs.dat")


def test_fused_cycles_and_host_edits_on_a_word_aligned_ladder(workdir):
    """Eight 256 x 512 lattices (2^20 sites: where PYMC_KERNEL_AUTO takes it): the batch-resident kernel applies, so the engine runs whole stretches of exchange cycles
    as one launch (TemperatureRX_MPI._advance).  The outcome must not depend on the kernel; and a host edit of a replica whose config object was created with another ``layout=`` than the
    first one must reach the batch of the engine (it used to be diverted to a private lattice)."""
    import py_mc_b200
    from py_mc_b200.ising2D import ising2D, ising2D_config, observer
    from py_mc_b200.mc import CanonicalMonteCarlo
    from py_mc_b200.mc_mpi import TemperatureRX_MPI
    from oracle import ising as o_ising
    R, shape, seed = 8, (256, 512), 31
    kTs = np.linspace(2.2, 2.35, R)

    def engine(layouts, no_persistent=False):
        py_mc_b200.set_seed(seed)
        if no_persistent:
            os.environ["PYMC_NO_PERSISTENT"] = "1"
        cfgs = []
        for r in range(R):
            c = ising2D_config(shape[0], shape[1], layout=layouts[r % len(layouts)])
            c.config[:, :] = o_ising.init_config(seed, r, *shape)
            cfgs.append(c)
        rx = TemperatureRX_MPI(None, CanonicalMonteCarlo, ising2D(-1.0), cfgs, kTs)
        os.environ.pop("PYMC_NO_PERSISTENT", None)
        return rx, cfgs

    results = []
    for k, (layouts, cuts, nop) in enumerate(((["int8"], (24,), False), (["int8", "bits"], (24,), False), (["int8"], (24,), True))):
        sub = workdir / ("run%d" % k)
        sub.mkdir()
        os.chdir(sub)
        rx, cfgs = engine(layouts, nop)
        assert rx._fused == (not nop)
        outs = [rx.run(n, 3, 2, 10 ** 12, observer(), True, True) for n in cuts]
        # a host edit between runs: flip one replica completely, then four more sweeps with exchanges
        cfgs[1].config[:, :] = -cfgs[1].config
        rx.run(4, 2)
        results.append((np.array(rx.rank_to_T), [np.array(c.config) for c in cfgs], [s.energy for s in rx.states], outs[0][0]))
    for other in results[1:]:
        assert np.array_equal(results[0][0], other[0])
        for a, b in zip(results[0][1], other[1]):
            assert np.array_equal(a, b)
        assert results[0][2] == other[2]
        assert np.array_equal(np.array(results[0][3]), np.array(other[3]))  # per-temperature means of the first run
