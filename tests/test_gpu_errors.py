"""GPU: a barrier inside a kernel that gives up waiting must surface as an error of the next call that returns
data, not as silently wrong data (persistent sweep kernel: pmc_persist.cuh / pmc_pbatch.cuh; slab neighbour
barrier: pmc_fused_i8.cuh).  The condition is raised through the library's test hook, which sets the same flag
from a kernel on the handle's stream."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _lattice():
    from py_mc_b200 import _cabi
    from py_mc_b200.ising2D import ising2D
    lat = _cabi.Lattices(_cabi.MODEL_ISING, 64, 128, nreplicas=2, ntemps=1, seed=1)
    lat.init_random()
    lat.set_accept_tables(ising2D(-1.0)._pmc_thresholds(2.3)[None, :])
    lat.sweep(2)
    return lat


@pytest.mark.parametrize("which,word", [(0, "persistent"), (1, "slab")])
@pytest.mark.parametrize("call", ["synchronize", "download", "tracked", "reduce_observables", "reduce_observables_static",
                                  "get_replica_to_T", "timer_stop"])
def test_timed_out_barrier_is_an_error_of_the_next_data_returning_call(which, word, call):
    from py_mc_b200 import _cabi
    lat = _lattice()
    lat.timer_start()
    lat.debug_raise_async_error(which)
    fn = {"synchronize": lat.synchronize, "download": lambda: lat.download(0), "tracked": lat.tracked,
          "reduce_observables": lat.reduce_observables, "reduce_observables_static": lat.reduce_observables_static,
          "get_replica_to_T": lat.get_replica_to_T, "timer_stop": lat.timer_stop}[call]
    with pytest.raises(_cabi.PymcError, match=word) as info:
        fn()
    assert info.value.code == -2
    # reported once: after re-initialising the handle works again
    lat.init_random()
    lat.sweep(1)
    lat.synchronize()
    b, m = lat.tracked()
    rb, rm = lat.reduce_observables()
    assert np.array_equal(b, rb) and np.array_equal(m, rm)
    lat.close()
