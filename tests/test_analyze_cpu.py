"""CPU: host-side mirrors of the reference's post-processing -- binning (py_mc/mc.py:47-61), obs
encode/decode (mc.py:65-104) and the per-temperature analysis script
(examples/analyze_ising2D_mpi.py), pinned to files the unmodified script wrote (tests/golden/rx.npz)."""
import os

import numpy as np


def test_analyze_writes_the_reference_files(golden_rx, tmp_path):
    from py_mc_b200.analyze import analyze, regroup
    z, meta = golden_rx
    for case in meta:
        n, R = case["name"], case["R"]
        d = tmp_path / n
        for r in range(R):
            (d / str(r)).mkdir(parents=True)
            np.save(d / str(r) / "obs_save.npy", z[n + "_obs_save"][r])
            np.save(d / str(r) / "Trank_hist.npy", z[n + "_Trank_hist"][r])
        means, errors = analyze(R, str(d))
        assert open(d / "obs_mean.dat").read() == case["obs_mean_dat"]   # byte-identical
        assert open(d / "binning.dat").read() == case["binning_dat"]
        series = regroup(z[n + "_obs_save"], z[n + "_Trank_hist"])
        # every sample of every temperature comes from exactly one replica
        assert np.array_equal(series.sum(axis=0), z[n + "_obs_save"].sum(axis=0))


def test_binning_and_obs_codec_match_reference_semantics():
    from py_mc_b200.mc import args_info, binning, obs_decode, obs_encode
    rng = np.random.default_rng(0)
    x = rng.normal(size=1000 + 37)
    err = binning(x, 5)
    assert len(err) == 6
    y = x[(len(x) % 32):]
    assert err[0] == np.sqrt(np.var(y, ddof=1) / len(y))
    y2 = (y[0::2] + y[1::2]) / 2.0
    assert err[1] == np.sqrt(np.var(y2, ddof=1) / len(y2))
    enc = obs_encode(1.5, np.arange(3.0), [[1.0, 2.0], [3.0, 4.0]])
    info = args_info(1.5, np.arange(3.0), [[1.0, 2.0], [3.0, 4.0]])
    assert list(info) == [3, 1, 3, 4] and enc.shape == (8,)
    dec = obs_decode(info, enc)
    assert dec[0] == 1.5 and np.array_equal(dec[1], np.arange(3.0)) and np.array_equal(dec[2], [1, 2, 3, 4])
