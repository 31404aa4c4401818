"""GPU physics checks with stated statistical error bars (independent streams): exact 4x4 canonical
averages incl. specific heat and Binder cumulant from the device-side accumulators, Kaufman's exact
finite-size energy at the Onsager T_c, and the device observer on the replica-exchange path."""
import os

import numpy as np
import pytest

from oracle import ising as o_ising
from oracle import kaufman

pytestmark = pytest.mark.gpu

# kT: (<E>/N, <|M|>/N, C_v/N, U4) of the 4x4 periodic Ising model, J = -1
EXACT_4x4 = {
    1.0: (-1.997158440256, 0.999274583701, 0.023409397457, 0.666553756794),
    2.0: (-1.755380288777, 0.918943267402, 0.605532657210, 0.644810791188),
    2.269185314: (-1.565623787638, 0.843860444813, 0.783266825929, 0.617199318412),
    3.0: (-1.017069626955, 0.601291326039, 0.603134714254, 0.480540660059),
    5.0: (-0.456135369543, 0.342765627554, 0.119546188787, 0.207859765073),
}


def test_device_accumulators_reproduce_exact_4x4_thermodynamics():
    from py_mc_b200 import _cabi
    from py_mc_b200.observers import thermo
    kTs = np.array(sorted(EXACT_4x4))
    nT, per = len(kTs), 256
    R = nT * per
    lat = _cabi.Lattices(_cabi.MODEL_ISING, 4, 4, nreplicas=R, ntemps=nT, seed=99)
    lat.init_random()
    lat.set_accept_tables(np.array([o_ising.threshold_table(-1.0, kT)[1] for kT in kTs]))
    lat.set_replica_to_T((np.arange(R) % nT).astype(np.int32))
    lat.sweep(500)
    blocks = []
    host_sum = np.zeros((nT, 6))
    for blk in range(16):
        lat.stats_reset(-1.0)
        for _ in range(100):
            lat.sweep(4)
            lat.stats_accumulate()
            if blk == 0:  # cross-check the kernel's sums against host arithmetic on the same samples
                b, m = lat.tracked()
                e, m = -1.0 * b.astype(np.float64), m.astype(np.float64)
                for k, v in enumerate((e, e * e, np.abs(m), m * m, m ** 4, m)):
                    host_sum[:, k] += np.bincount(np.arange(R) % nT, weights=v, minlength=nT)
        sums, counts = lat.stats_get()
        assert np.all(counts == 100 * per)
        if blk == 0:
            assert np.array_equal(sums, host_sum)  # integer-valued doubles: exact
        blocks.append(thermo(sums, counts, kTs))
    for i, kT in enumerate(kTs):
        e_ex, m_ex, cv_ex, u_ex = EXACT_4x4[kT]
        for key, exact, scale in (("mean_E", e_ex, 16.0), ("mean_absM", m_ex, 16.0), ("specific_heat", cv_ex, 16.0),
                                  ("binder", u_ex, 1.0)):
            vals = np.array([b[key][i] for b in blocks]) / scale
            mean, err = vals.mean(), vals.std(ddof=1) / np.sqrt(len(vals))
            assert abs(mean - exact) < 5 * err + 2e-4, (kT, key, mean, exact, err)


def test_energy_at_onsager_tc_matches_kaufman_32x32():
    """T_c = 2/ln(1+sqrt 2); <E>/N and C_v/N of the 32x32 torus from Kaufman's exact partition function."""
    from py_mc_b200 import _cabi
    L, R = 32, 512
    kT = kaufman.onsager_tc()
    e_ex, cv_ex = kaufman.energy_and_cv_per_site(L, L, kT)
    lat = _cabi.Lattices(_cabi.MODEL_ISING, L, L, nreplicas=R, ntemps=1, seed=2718)
    lat.init_random()
    lat.set_accept_tables(o_ising.threshold_table(-1.0, kT)[1][None, :])
    lat.sweep(4000)
    es = []
    for _ in range(150):
        lat.sweep(20)
        b, _ = lat.tracked()
        es.append(-1.0 * b.astype(np.float64))
    es = np.array(es)                     # [sample, replica]; replicas are independent chains
    per_rep = es.mean(axis=0) / (L * L)
    mean, err = per_rep.mean(), per_rep.std(ddof=1) / np.sqrt(R)
    assert abs(mean - e_ex) < 5 * err, (mean, e_ex, err)
    cv_rep = es.var(axis=0, ddof=1) / (kT * kT) / (L * L)
    cv, cv_err = cv_rep.mean(), cv_rep.std(ddof=1) / np.sqrt(R)
    # the per-replica variance over a finite correlated window underestimates C_v slightly; allow that bias
    assert abs(cv - cv_ex) < 5 * cv_err + 0.08 * cv_ex, (cv, cv_ex, cv_err)


def test_device_observer_on_replica_exchange(tmp_path):
    import py_mc_b200
    from py_mc_b200.ising2D import ising2D, ising2D_config
    from py_mc_b200.mc import CanonicalMonteCarlo
    from py_mc_b200.mc_mpi import TemperatureRX_MPI
    from py_mc_b200.observers import device_observer
    old = os.getcwd()
    os.chdir(tmp_path)
    try:
        py_mc_b200.set_seed(31)
        R, L = 16, 16
        kTs = np.linspace(1.8, 3.0, R)
        configs = []
        for _ in range(R):
            c = ising2D_config(L, L)
            c.prepare_random()
            configs.append(c)
        rx = TemperatureRX_MPI(None, CanonicalMonteCarlo, ising2D(-1.0), configs, kTs)
        rx.run(1500, 5)
        obs = device_observer()
        out = rx.run(6000, 5, 5, observer=obs)
        assert len(out) == R and np.all(obs.counts == 1200)
        assert rx.accept_count > 0
        e = np.array([o[0] for o in out]) / (L * L)
        assert np.all(np.diff(e) > -0.02) and e[0] < e[-1]      # energy rises with temperature
        for i in (0, R // 2, R - 1):
            e_ex, _ = kaufman.energy_and_cv_per_site(L, L, kTs[i])
            assert abs(e[i] - e_ex) < 0.02, (kTs[i], e[i], e_ex)
        assert obs.binder[0] > obs.binder[-1]                    # ordered side -> 2/3, disordered -> 0
        assert np.argmax(obs.specific_heat) not in (0, R - 1)    # C_v peaks inside the ladder (near T_c)
    finally:
        os.chdir(old)


def test_device_observer_with_canonical_engine(tmp_path):
    """Same samples, two observers: the device accumulators must equal the host observer's means exactly
    (integer-valued sums) when both run the same random stream."""
    import py_mc_b200
    from py_mc_b200.ising2D import ising2D, ising2D_config, observer
    from py_mc_b200.mc import CanonicalMonteCarlo
    from py_mc_b200.observers import device_observer
    old = os.getcwd()
    os.chdir(tmp_path)
    try:
        init = o_ising.init_config(9, 0, 16, 16)
        results = []
        for obs in (observer(), device_observer()):
            py_mc_b200.set_seed(9)
            cfg = ising2D_config(16, 16)
            cfg.config[:, :] = init
            calc = CanonicalMonteCarlo(ising2D(-1.0), 2.4, cfg)
            results.append(calc.run(600, 3, 10 ** 9, obs))
            last = obs
        assert results[0][0] == pytest.approx(results[1][0], rel=1e-13)
        assert results[0][1] == pytest.approx(results[1][1], rel=1e-13)
        assert last.counts[0] == 200 and last.specific_heat[0] > 0 and 0 < last.binder[0] < 2.0 / 3.0 + 1e-9
    finally:
        os.chdir(old)


@pytest.mark.parametrize("layout", [0, 1])
def test_onsager_energy_and_spontaneous_magnetisation_1024(layout):
    """Large lattice, independent stream, below T_c: <E>/N against the exact value (Kaufman at L = 512 equals the
    Onsager limit to 1e-9) and <|M|>/N against Onsager-Yang m = (1 - sinh(2/kT)^-4)^(1/8).  Error bars: the sample
    standard deviation of e is sqrt(kT^2 c_v / N) = 1.7e-3, of m about 1e-3; 64 samples 20 sweeps apart."""
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D
    L, kT, nsamp = 1024, 2.0, 64
    lat = _cabi.Lattices(_cabi.MODEL_ISING, L, L, seed=777, layout=layout)
    lat.upload(0, np.ones((L, L), dtype=np.int8))          # cold start
    lat.set_accept_tables(ising2D(-1.0)._pmc_thresholds(kT)[None, :])
    lat.sweep(300)
    e, m = [], []
    for _ in range(nsamp):
        lat.sweep(20)
        b, mag = lat.tracked()
        e.append(-float(b[0]) / (L * L))
        m.append(abs(float(mag[0])) / (L * L))
    b2, m2 = lat.reduce_observables()
    assert int(b2[0]) == int(b[0]) and int(m2[0]) == int(mag[0])   # tracked sums stayed exact over 1580 sweeps
    lat.close()
    e_exact, _ = kaufman.energy_and_cv_per_site(512, 512, kT)
    m_exact = (1.0 - np.sinh(2.0 / kT) ** -4) ** 0.125
    e_err = np.std(e, ddof=1) / np.sqrt(nsamp)
    m_err = np.std(m, ddof=1) / np.sqrt(nsamp)
    assert abs(np.mean(e) - e_exact) < 6 * e_err + 2e-4, (np.mean(e), e_exact, e_err)
    assert abs(np.mean(m) - m_exact) < 6 * m_err + 2e-4, (np.mean(m), m_exact, m_err)
    assert e_err < 1e-3 and m_err < 1e-3


def test_binder_cumulants_cross_at_the_onsager_temperature(tmp_path):
    """T_c from the crossing of U4(T) for L = 16 and L = 32 (replica exchange over a ladder around T_c, device-side
    accumulators): the crossing of this pair of sizes lies within 0.02 of 2/ln(1+sqrt 2) = 2.269185 and the crossing
    value near the universal U* = 0.61 of periodic square lattices."""
    import py_mc_b200
    from py_mc_b200.ising2D import ising2D, ising2D_config
    from py_mc_b200.mc import CanonicalMonteCarlo
    from py_mc_b200.mc_mpi import TemperatureRX_MPI
    from py_mc_b200.observers import device_observer
    old = os.getcwd()
    os.chdir(tmp_path)
    try:
        R = 16
        kTs = np.linspace(2.15, 2.40, R)
        U = {}
        for L in (16, 32):
            py_mc_b200.set_seed(1000 + L)
            configs = []
            for _ in range(R):
                c = ising2D_config(L, L)
                c.prepare_random()
                configs.append(c)
            sub = tmp_path / ("L%d" % L)
            sub.mkdir()
            os.chdir(sub)
            rx = TemperatureRX_MPI(None, CanonicalMonteCarlo, ising2D(-1.0), configs, kTs)
            rx.run(3000, 5)
            obs = device_observer()
            rx.run(40000, 5, 5, observer=obs)
            assert np.all(obs.counts == 8000)
            U[L] = np.array(obs.binder)
        d = U[16] - U[32]                       # > 0 above T_c (the smaller lattice stays more ordered), < 0 below
        assert d[0] < 0 < d[-1], d
        k = int(np.where((d[:-1] <= 0) & (d[1:] > 0))[0][0])
        t_cross = kTs[k] - d[k] * (kTs[k + 1] - kTs[k]) / (d[k + 1] - d[k])
        u_cross = U[32][k] + (U[32][k + 1] - U[32][k]) * (t_cross - kTs[k]) / (kTs[k + 1] - kTs[k])
        assert abs(t_cross - 2.269185314213022) < 0.02, (t_cross, d)
        assert 0.58 < u_cross < 0.63, u_cross
    finally:
        os.chdir(old)


def test_device_series_and_binning_match_the_reference_post_processing(tmp_path):
    """The same replica-exchange run twice: with the host observer (per-replica obs_save.npy / Trank_hist.npy, then
    py_mc_b200.analyze.analyze, whose output is byte-identical to examples/analyze_ising2D_mpi.py -- tests/
    test_analyze_cpu.py) and with device_observer(series=True), where the series by temperature and the blocking
    analysis (py_mc/mc.py:47-61) stay on the device.  The series must agree exactly (they are the same exact integers
    times J), the means to rounding, the blocking errors to 1e-9 relative (summation order)."""
    import py_mc_b200
    from py_mc_b200 import analyze
    from py_mc_b200.ising2D import ising2D, ising2D_config, observer
    from py_mc_b200.mc import CanonicalMonteCarlo, binning
    from py_mc_b200.mc_mpi import TemperatureRX_MPI
    from py_mc_b200.observers import device_observer
    R, shape, seed, nsteps, every, sample = 6, (16, 16), 77, 1500, 2, 3
    kTs = np.linspace(2.0, 2.8, R)

    def engine():
        py_mc_b200.set_seed(seed)
        cfgs = []
        for r in range(R):
            c = ising2D_config(*shape)
            c.config[:, :] = o_ising.init_config(seed, r, *shape)
            cfgs.append(c)
        return TemperatureRX_MPI(None, CanonicalMonteCarlo, ising2D(-1.0), cfgs, kTs)

    old = os.getcwd()
    try:
        os.chdir(tmp_path)
        engine().run(nsteps, every, sample, 10 ** 12, observer(), True, True)
        means, errors = analyze.analyze(R, ".", write=False)
        obs = np.array([np.load(os.path.join(str(r), "obs_save.npy")) for r in range(R)])
        th = np.array([np.load(os.path.join(str(r), "Trank_hist.npy")) for r in range(R)])
        host_series = analyze.regroup(obs, th)          # [T, sample, 2]
        dev = device_observer(series=True)
        engine().run(nsteps, every, sample, observer=dev)
    finally:
        os.chdir(old)
    series = dev.series()                                # [sample, T, 2]
    assert series.shape == (nsteps // sample, R, 2)
    assert np.array_equal(series.transpose(1, 0, 2), host_series)
    dmean, derr, levels = dev.analyze()
    assert np.allclose(dmean, means, rtol=1e-13, atol=0)
    assert np.allclose(derr, errors, rtol=1e-9, atol=0)
    nlev = int(np.log2(series.shape[0] // 30))
    for t in (0, R - 1):
        for j in (0, 1):
            assert np.allclose(levels[t, j], binning(series[:, t, j], nlev), rtol=1e-9, atol=0)
