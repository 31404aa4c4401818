"""GPU: asynchronous upload / download / tracked sums (pymc_*_async) and the pipeline built on them
(py_mc_b200/pipeline.py).  A job of the pipeline -- host configuration in, sweeps, configuration and exact sums out --
must give exactly what the synchronous calls give and what the oracle computes, whichever handle ran it and whatever
was in flight beside it."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TC = 2.269185314213022
SEED = 77


def _thr(kT=TC):
    from py_mc_b200.ising2D import ising2D
    return ising2D(-1.0)._pmc_thresholds(kT)


def _sync_job(cfg, first_sweep, nsweeps, kernel):
    from py_mc_b200 import _cabi
    lat = _cabi.Lattices(_cabi.MODEL_ISING, cfg.shape[0], cfg.shape[1], seed=SEED, kernel=kernel)
    lat.set_accept_tables(_thr()[None, :])
    lat.upload(0, cfg)
    lat.sweep_counter = first_sweep
    lat.sweep(nsweeps)
    out = lat.download(0)
    b, m = lat.tracked()
    lat.close()
    return out, int(b[0]), int(m[0])


@pytest.mark.parametrize("shape,nsweeps,kernel", [((64, 256), 3, 0), ((64, 256), 1, 0), ((296, 256), 4, 5), ((32, 1024), 1, 0)])
def test_pipeline_jobs_equal_synchronous_calls_and_oracle(shape, nsweeps, kernel):
    from py_mc_b200 import _cabi
    from py_mc_b200.pipeline import SweepPipeline
    from oracle import ising as o_ising
    rng = np.random.default_rng(3)
    njobs, depth = 7, 3
    ins = [_cabi.pinned_empty(shape, np.int8) for _ in range(njobs)]
    outs = [_cabi.pinned_empty(shape, np.int8) for _ in range(njobs)]
    for a in ins:
        a[...] = rng.choice(np.array([-1, 1], dtype=np.int8), size=shape)
    firsts = [5 * j for j in range(njobs)]
    results = []
    with SweepPipeline(_cabi.MODEL_ISING, shape[0], shape[1], _thr(), nsweeps, depth=depth, seed=SEED, kernel=kernel) as pipe:
        for j in range(njobs):
            r = pipe.submit(ins[j], outs[j], first_sweep=firsts[j], tag=j)
            if r is not None:
                results.append(r)
        results += pipe.drain()
    assert [r[0] for r in results] == list(range(njobs))  # submission order
    thr = o_ising.threshold_table(-1.0, TC)
    for j, (tag, out, bond, second) in enumerate(results):
        assert out is outs[j]
        want, wb, wm = _sync_job(np.array(ins[j]), firsts[j], nsweeps, kernel)
        assert np.array_equal(out, want), j
        assert (bond, second) == (wb, wm)
        if j < 2:  # and the oracle, on the first jobs
            ref = np.array(ins[j])
            for sw in range(nsweeps):
                o_ising.sweep(ref, -1.0, TC, SEED, 0, firsts[j] + sw, thr)
            assert np.array_equal(out, ref)
            assert bond == o_ising.bond_sum(ref) and second == int(ref.astype(np.int64).sum())


def test_async_calls_reject_wrong_buffers_and_report_barrier_errors():
    from py_mc_b200 import _cabi
    lat = _cabi.Lattices(_cabi.MODEL_ISING, 64, 128, seed=1)
    lat.set_accept_tables(_thr()[None, :])
    good = _cabi.pinned_empty((64, 128), np.int8)
    good[...] = 1
    with pytest.raises(ValueError):
        lat.upload_async(0, np.ones((64, 64), dtype=np.int8))
    with pytest.raises(ValueError):
        lat.download_async(0, np.ones((64, 128), dtype=np.int16))
    with pytest.raises(_cabi.PymcError):
        lat.upload_async(3, good)  # no such replica
    lat.upload_async(0, good)
    lat.sweep(2)
    lat.debug_raise_async_error(0)
    lat.download_async(0, good)
    with pytest.raises(_cabi.PymcError, match="persistent"):
        lat.synchronize()  # the asynchronous calls report a failed barrier at the synchronisation that completes them
    lat.close()
