"""CPU: checks against the LIVE, unmodified reference -- run wherever its files are present (/root/reference in the
build container, or the git-ignored copy oracle/_ref made by oracle/make_ref.py), skipped elsewhere.

* every committed fixture under tests/golden/ is regenerated in memory by the committed generator scripts and must
  come back array for array identical (the fixtures are what pins the oracle and the CUDA path to the reference);
* ``prepare_random`` (examples/ising2D.py:64-70): the drop-in class draws the same configuration from the same
  ``random`` state, and the device hot start (stream domain 1, oracle.ising.init_config) is what the reference's own
  ``prepare_random`` produces when its ``rand.random`` is fed the stream's uniforms in its row-major draw order.
"""
import contextlib
import io
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

from oracle.make_ref import reference_root  # noqa: E402

REF = reference_root()
pytestmark = pytest.mark.skipif(REF is None, reason="the reference files are not on this machine")


@pytest.fixture(scope="module")
def generators():
    os.environ["PYMC_REFERENCE"] = REF
    sys.path.insert(0, GOLDEN)
    import make_golden as mg
    from make_golden_more import MORE
    gens = {"ising": mg.gen_ising, "ising_de": mg.gen_ising_de}
    gens.update(MORE)
    return gens


@pytest.mark.parametrize("family", ["ising", "ising_de", "rx", "latgas", "latgas_move", "discs"])
def test_committed_fixture_is_what_the_reference_produces_today(generators, family):
    out = {}
    with contextlib.redirect_stdout(io.StringIO()):
        generators[family](out)
    z = np.load(os.path.join(GOLDEN, family + ".npz"))
    assert set(out) == set(z.files)
    for k in out:
        assert np.array_equal(np.asarray(out[k]), z[k]), (family, k)


def _reference_ising():
    for p in (os.path.join(REF, "examples"), REF):
        if p not in sys.path:
            sys.path.insert(0, p)
    import ising2D as ref_ising
    return ref_ising


def test_prepare_random_draws_like_the_reference():
    import random
    ref_ising = _reference_ising()
    from py_mc_b200.ising2D import ising2D_config
    ref_ising.rand = random  # (the fixture generators above leave their draw injector in the module)
    random.seed(20260611)
    a = ref_ising.ising2D_config(6, 10)
    a.prepare_random()
    random.seed(20260611)
    b = ising2D_config(6, 10)
    b.prepare_random()
    assert np.array_equal(np.asarray(a.config), np.asarray(b.config))
    assert str(a) == str(b)


def test_device_hot_start_is_the_reference_rule_on_the_stream():
    """oracle.ising.init_config is what init_kernel is tested against on the GPU (tests/test_gpu_ising.py); here it is
    tied to the reference: prepare_random fed with the stream's uniforms, in the order in which it draws."""
    from oracle import ising as o_ising
    from oracle import philox
    ref_ising = _reference_ising()
    seed, replica, lenX, lenY = 4711, 3, 6, 12
    x32 = philox.init_x32(seed, replica, lenX * lenY)  # one draw per site, row-major
    u = iter(philox.x32_to_uniform(np.asarray(x32)).ravel().tolist())

    class Inj:
        @staticmethod
        def random():
            return next(u)

    saved = ref_ising.rand
    ref_ising.rand = Inj
    try:
        cfg = ref_ising.ising2D_config(lenX, lenY)
        cfg.prepare_random()
    finally:
        ref_ising.rand = saved
    assert np.array_equal(np.asarray(cfg.config).astype(np.int8), o_ising.init_config(seed, replica, lenX, lenY))
