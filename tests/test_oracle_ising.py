"""CPU: pin the oracle (oracle/ising.py, oracle/philox.py) to the golden fixtures that
tests/golden/make_golden.py recorded from the unmodified reference."""
import math

import numpy as np
import pytest

from oracle import ising as o_ising
from oracle import philox


def test_philox_known_answers():
    # Random123 KATs for Philox4x32-10
    def run(c, k):
        return [int(v) for v in philox.philox4x32_10(np.array(c, dtype=np.uint32), k)]

    assert run([0, 0, 0, 0], (0, 0)) == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    assert run([0xFFFFFFFF] * 4, (0xFFFFFFFF, 0xFFFFFFFF)) == [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]
    assert run([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], (0xA4093822, 0x299F31D0)) == [
        0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]


def test_plane_definition_of_x32():
    # x32 of a lane is assembled MSB-first from the 32 planes of its word
    seed, rep, step = 11, 3, 5
    x = philox.word_lane_x32(seed, rep, step, np.array([7]))[0]
    even, odd = philox.word_planes(seed, rep, step, np.array([7]), 0)
    for lane in (0, 1, 31, 63):
        assert (int(x[lane]) >> 31) & 1 == (int(even[0]) >> lane) & 1
        assert (int(x[lane]) >> 30) & 1 == (int(odd[0]) >> lane) & 1


def test_threshold_equivalence():
    # u <= p  <=>  x32 <= floor(p * 2**32), on both sides of each threshold
    for J, kT in [(-1.0, 5.0), (-1.0, 2.269185314213022), (-1.0, 0.01), (0.3, 0.7), (1.0, 2.0)]:
        thr = o_ising.threshold_table(J, kT)
        for si, s in enumerate((-1, 1)):
            for ni, nn in enumerate((-4, -2, 0, 2, 4)):
                dE = o_ising.delta_e(J, s, nn)
                t = int(thr[si, ni])
                for x in {0, 1, max(t - 1, 0), t, min(t + 1, 0xFFFFFFFF), 0xFFFFFFFF}:
                    u = x / 4294967296.0
                    ref = (dE < 0.0) or (u <= math.exp(-dE / kT))
                    assert ref == (x <= t), (J, kT, s, nn, x, t)
                # symmetric table: only s*nn matters (what the bit-packed kernel relies on)
                assert thr[si, ni] == thr[1 - si, 4 - ni]


def test_delta_e_and_energy_match_reference(golden_ising_de):
    z, meta = golden_ising_de
    for m in meta:
        init = z[m["name"] + "_init"]
        nn = o_ising.neighbour_sum(init)
        dE = np.array([[o_ising.delta_e(m["J"], init[x, y], nn[x, y]) for y in range(m["lenY"])]
                       for x in range(m["lenX"])])
        assert np.array_equal(dE, z[m["name"] + "_dE"])  # bit-exact float64
        assert o_ising.energy_reference_order(init, m["J"]) == float(z[m["name"] + "_energy"])
        assert o_ising.magnetization(init) == float(z[m["name"] + "_mag"])
        if float(m["J"]).is_integer():
            assert m["J"] * o_ising.bond_sum(init) == float(z[m["name"] + "_energy"])


def test_init_config_matches_fixture(golden_ising):
    z, meta = golden_ising
    for m in meta:
        init = o_ising.init_config(m["seed"], m["replica"], m["lenX"], m["lenY"])
        assert np.array_equal(init, z[m["name"] + "_init"])


def test_vectorised_sweep_matches_reference_trajectory(golden_ising):
    z, meta = golden_ising
    for m in meta:
        cfg = z[m["name"] + "_init"].copy()
        J, kT = m["J"], m["kT"]
        B = o_ising.bond_sum(cfg)
        for sw in range(m["nsweeps"]):
            nacc, dB, dM = o_ising.sweep(cfg, J, kT, m["seed"], m["replica"], sw)
            B += dB
            assert np.array_equal(cfg, z[m["name"] + "_configs"][sw]), (m["name"], sw)
            assert B == o_ising.bond_sum(cfg)
            assert o_ising.magnetization(cfg) == z[m["name"] + "_mag"][sw]
            if float(J).is_integer():
                assert J * B == z[m["name"] + "_energy"][sw]
            else:
                assert J * B == pytest.approx(z[m["name"] + "_energy"][sw], abs=1e-9)


def test_python_site_loop_matches_reference_trajectory(golden_ising):
    # the scalar float64 port (the thing timed as the CPU baseline) gives the same trajectory
    z, meta = golden_ising
    for m in meta[:6]:
        cfg = z[m["name"] + "_init"].astype(np.float64)
        e = o_ising.energy_reference_order(cfg, m["J"])
        assert e == float(z[m["name"] + "_e0"])
        for sw in range(m["nsweeps"]):
            # the reference recomputes the energy at the start of every run() (mc.py:191)
            e = o_ising.energy_reference_order(cfg, m["J"])
            e = o_ising.sweep_sites_python(cfg, m["J"], m["kT"], m["seed"], m["replica"], sw, e)
            assert np.array_equal(cfg.astype(np.int8), z[m["name"] + "_configs"][sw])
            assert e == z[m["name"] + "_energy"][sw]  # bit-exact incl. the non-dyadic J case
