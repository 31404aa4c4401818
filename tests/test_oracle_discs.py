"""CPU: pin oracle/harddisc.py to the fixture recorded from examples/2D_hardcore/model_setup.py."""
import numpy as np

from oracle import harddisc as o_hd


def test_trajectory_and_gr_match_reference(golden_discs):
    z, meta = golden_discs
    for m in meta:
        n = m["name"]
        c = o_hd.prepare_ordered(m["L"], m["ndisc"], m["rdisc"])
        assert np.array_equal(c, z[n + "_start"])
        assert not o_hd.overlaps(c, m["L"], m["rdisc"]) and m["e0"] == 0
        nb = len(z[n + "_gridx"])
        g0 = o_hd.gr_counts(c, m["L"], m["dr"], nb).astype(np.float64)
        g0 *= m["g_pref"] / (z[n + "_gridx"] * m["dr"])
        assert np.array_equal(g0, z[n + "_g0"])
        nacc, ids = 0, []
        for t in range(m["ntrials"]):
            i, acc = o_hd.trial(c, m["L"], m["rdisc"], m["maxstep"], m["seed"], m["replica"], t)
            nacc += acc
            ids.append(i)
            if (t + 1) % 500 == 0:
                assert np.array_equal(c, z[n + "_snaps"][(t + 1) // 500 - 1]), (n, t)  # bit-exact float64
        assert nacc == m["nacc"] and ids[:50] == list(z[n + "_ids"])
        g = o_hd.gr_counts(c, m["L"], m["dr"], nb).astype(np.float64)
        g *= m["g_pref"] / (z[n + "_gridx"] * m["dr"])
        assert np.array_equal(g, z[n + "_g_end"])
