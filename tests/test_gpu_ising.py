"""GPU parity tests for the Ising path: CUDA kernels (through the C ABI) vs the oracle and the
golden fixtures recorded from the reference.  Integer/byte results are compared bit-exactly."""
import numpy as np
import pytest

from oracle import ising as o_ising

pytestmark = pytest.mark.gpu

INT8, BITS = 0, 1


def _cabi():
    from py_mc_b200 import _cabi
    return _cabi


def _layouts(lenY):
    return [INT8, BITS] if lenY % 128 == 0 else [INT8]


def _make(case, layout, **kw):
    c = _cabi()
    return c.Lattices(c.MODEL_ISING, case["lenX"], case["lenY"], nreplicas=1, ntemps=1, seed=case["seed"],
                      layout=layout, replica_offset=case["replica"], nreplicas_global=case["replica"] + 1, **kw)


def test_device_present():
    assert _cabi().device_count() >= 1


@pytest.mark.parametrize("shape", [(2, 2), (10, 10), (6, 130), (4, 260), (16, 128), (8, 384), (64, 4)])
def test_upload_download_roundtrip(shape):
    c = _cabi()
    rng = np.random.default_rng(1)
    cfg = rng.choice(np.array([-1, 1], dtype=np.int8), size=shape)
    for layout in _layouts(shape[1]):
        lat = c.Lattices(c.MODEL_ISING, shape[0], shape[1], layout=layout)
        lat.upload(0, cfg)
        assert np.array_equal(lat.download(0), cfg)
        b, m = lat.reduce_observables()
        assert int(b[0]) == o_ising.bond_sum(cfg)
        assert int(m[0]) == o_ising.magnetization(cfg)
        lat.close()


def test_unsupported_shapes_fail_loudly():
    c = _cabi()
    with pytest.raises(c.PymcError, match="even"):
        c.Lattices(c.MODEL_ISING, 5, 8)
    with pytest.raises(c.PymcError, match="128"):
        c.Lattices(c.MODEL_ISING, 8, 64, layout=BITS)
    with pytest.raises(c.PymcError, match="too large"):  # 2^35 sites: past the 32-bit chunk index of the upload kernels
        c.Lattices(c.MODEL_ISING, 262144, 131072)
    lat = c.Lattices(c.MODEL_ISING, 8, 8)
    with pytest.raises(c.PymcError, match="set_accept_tables"):
        lat.sweep(1)


def test_energy_and_dE_match_reference(golden_ising_de):
    z, meta = golden_ising_de
    c = _cabi()
    for m in meta:
        init = z[m["name"] + "_init"]
        lat = c.Lattices(c.MODEL_ISING, m["lenX"], m["lenY"])
        lat.upload(0, init)
        b, mag = lat.reduce_observables()
        assert float(mag[0]) == float(z[m["name"] + "_mag"])
        if float(m["J"]).is_integer() or m["J"] == 2.5:
            assert m["J"] * float(b[0]) == float(z[m["name"] + "_energy"])  # bit-exact for dyadic J
        else:
            assert m["J"] * float(b[0]) == pytest.approx(float(z[m["name"] + "_energy"]), abs=1e-12)
        xs, ys = np.meshgrid(np.arange(m["lenX"]), np.arange(m["lenY"]), indexing="ij")
        s, nn = lat.propose(0, xs.ravel(), ys.ravel())
        dE = np.array([-2.0 * m["J"] * float(a) * float(b_) for a, b_ in zip(s, nn)]).reshape(m["lenX"], m["lenY"])
        assert np.array_equal(dE, z[m["name"] + "_dE"])  # bit-exact float64, every site
        assert np.array_equal(lat.download(0), init)  # config unchanged by propose
        lat.close()


def test_init_random_matches_stream_definition(golden_ising):
    z, meta = golden_ising
    for m in meta:
        for layout in _layouts(m["lenY"]):
            lat = _make(m, layout)
            lat.init_random()
            assert np.array_equal(lat.download(0), z[m["name"] + "_init"]), (m["name"], layout)
            lat.close()


def test_trajectories_match_reference(golden_ising):
    """Accept/reject trajectories are bit-exact with the reference fed the same uniforms."""
    z, meta = golden_ising
    for m in meta:
        thr = o_ising.threshold_table(m["J"], m["kT"])[1]  # row s=+1: index = aligned neighbours
        for layout in _layouts(m["lenY"]):
            lat = _make(m, layout)
            lat.upload(0, z[m["name"] + "_init"])
            lat.set_accept_tables(thr[None, :])
            for sw in range(m["nsweeps"]):
                lat.sweep(1)
                cfg = lat.download(0)
                assert np.array_equal(cfg, z[m["name"] + "_configs"][sw]), (m["name"], layout, sw)
                b, mag = lat.tracked()
                assert float(mag[0]) == z[m["name"] + "_mag"][sw]
                if float(m["J"]).is_integer():
                    assert m["J"] * float(b[0]) == z[m["name"] + "_energy"][sw]
                b2, mag2 = lat.reduce_observables()
                assert (b2[0], mag2[0]) == (b[0], mag[0])  # tracked == recomputed
            assert lat.sweep_counter == m["nsweeps"]
            lat.close()


@pytest.mark.parametrize("shape,kT", [((64, 256), 2.269185314213022), ((128, 1024), 2.0), ((30, 44), 3.0),
                                      ((6, 256), 2.5), ((258, 128), 1.5), ((40, 4096), 2.3)])
def test_sweeps_match_oracle_medium(shape, kT):
    c = _cabi()
    seed, nsweeps, J = 2026, 4, -1.0
    ref = o_ising.init_config(seed, 0, *shape)
    B = o_ising.bond_sum(ref)
    thr = o_ising.threshold_table(J, kT)
    lats = []
    for layout in _layouts(shape[1]):
        lat = c.Lattices(c.MODEL_ISING, shape[0], shape[1], seed=seed, layout=layout)
        lat.init_random()
        lat.set_accept_tables(thr[1][None, :])
        lats.append(lat)
    for sw in range(nsweeps):
        _, dB, _ = o_ising.sweep(ref, J, kT, seed, 0, sw, thr)
        B += dB
        for lat in lats:
            lat.sweep(1)
            assert np.array_equal(lat.download(0), ref), (shape, lat.layout, sw)
            b, m = lat.tracked()
            assert int(b[0]) == B and int(m[0]) == o_ising.magnetization(ref)


def test_int8_and_bits_agree_on_large_lattice():
    """Size-independent property at a size the oracle cannot reach: both layouts run the same
    random stream, so their configurations must stay identical; tracked sums must equal recomputed."""
    c = _cabi()
    L, seed = 2048, 99
    kT = 2.269185314213022
    thr = o_ising.threshold_table(-1.0, kT)[1][None, :]
    a = c.Lattices(c.MODEL_ISING, L, L, seed=seed, layout=INT8)
    b = c.Lattices(c.MODEL_ISING, L, L, seed=seed, layout=BITS)
    for lat in (a, b):
        lat.init_random()
        lat.set_accept_tables(thr)
        lat.sweep(20)
    ca, cb = a.download(0), b.download(0)
    assert np.array_equal(ca, cb)
    assert set(np.unique(ca)) <= {-1, 1}
    ta, tb = a.tracked(), b.tracked()
    assert ta[0][0] == tb[0][0] and ta[1][0] == tb[1][0]
    ra = a.reduce_observables()
    assert ra[0][0] == ta[0][0] and ra[1][0] == ta[1][0]
    assert ra[0][0] == o_ising.bond_sum(ca)


def test_batched_replicas_are_independent_streams():
    """Replica r of a batch must evolve exactly like a single-replica run with replica index r."""
    c = _cabi()
    lenX, lenY, seed, R = 12, 20, 5, 7
    kTs = np.linspace(1.0, 4.0, R)
    thr = np.array([o_ising.threshold_table(-1.0, kT)[1] for kT in kTs])
    lat = c.Lattices(c.MODEL_ISING, lenX, lenY, nreplicas=R, ntemps=R, seed=seed)
    lat.init_random()
    lat.set_accept_tables(thr)
    lat.sweep(6)
    for r in range(R):
        ref = o_ising.init_config(seed, r, lenX, lenY)
        for sw in range(6):
            o_ising.sweep(ref, -1.0, kTs[r], seed, r, sw)
        assert np.array_equal(lat.download(r), ref), r
    # permuting the temperature map permutes which table row each replica uses
    perm = np.roll(np.arange(R), 1).astype(np.int32)
    lat.set_replica_to_T(perm)
    assert np.array_equal(lat.get_replica_to_T(), perm)
    before = [lat.download(r) for r in range(R)]
    lat.sweep(1)
    for r in range(R):
        ref = before[r].copy()
        o_ising.sweep(ref, -1.0, kTs[perm[r]], seed, r, 6)
        assert np.array_equal(lat.download(r), ref), r


@pytest.mark.parametrize("shape,nrep", [((32, 1024), 1), ((64, 2048), 3), ((96, 1024), 2)])
def test_fused_two_pass_and_generic_kernels_agree(shape, nrep):
    """The fused two-colour int8 kernel recomputes a halo ring per tile; it must follow the same
    trajectory as the two-pass and the generic kernels (and the oracle), including lattices that are a
    single tile wide or high, where the ring wraps onto the tile itself."""
    c = _cabi()
    seed, kT = 4711, 2.3
    thr = o_ising.threshold_table(-1.0, kT)[1][None, :]
    lats = {}
    for name, kernel, layout in (("fused", c.KERNEL_AUTO, INT8), ("two_pass", c.KERNEL_TWO_PASS, INT8),
                                 ("generic", c.KERNEL_GENERIC, INT8), ("bits", c.KERNEL_AUTO, BITS)):
        lat = c.Lattices(c.MODEL_ISING, shape[0], shape[1], nreplicas=nrep, seed=seed, layout=layout, kernel=kernel)
        lat.init_random()
        lat.set_accept_tables(thr)
        lats[name] = lat
    # the kernels that serve calls of a single sweep (longer calls of the AUTO handles stay on chip: test_zx_gpu_pbatch.py)
    assert lats["fused"].sweep_kernel_for(1) == (1, "sweep_fused_i8_kernel")
    assert lats["two_pass"].sweep_kernel_for(1) == (2, "sweep_tiled_i8_kernel")
    assert lats["generic"].sweep_kernel_for(1) == (2, "sweep_word_kernel")
    assert lats["bits"].sweep_kernel_for(1) == (2, "sweep_bits_rows_kernel")
    ref = o_ising.init_config(seed, nrep - 1, *shape)
    for sw in range(5):  # odd number of sweeps: the fused kernel ends in its second buffer
        for lat in lats.values():
            lat.sweep(1)
        o_ising.sweep(ref, -1.0, kT, seed, nrep - 1, sw)
    want = {r: lats["generic"].download(r) for r in range(nrep)}
    assert np.array_equal(want[nrep - 1], ref)
    for name, lat in lats.items():
        for r in range(nrep):
            assert np.array_equal(lat.download(r), want[r]), (name, r)
        assert np.array_equal(lat.tracked()[0], lats["generic"].tracked()[0]), name
        assert np.array_equal(lat.tracked()[1], lats["generic"].tracked()[1]), name
        b, m = lat.reduce_observables()
        assert np.array_equal(b, lats["generic"].tracked()[0]) and np.array_equal(m, lats["generic"].tracked()[1])


def test_large_lattice_against_scalar_c_oracle():
    """4096 x 4096 (256 fused tiles per colour-row block, 8 tiles wide): all three int8/bit kernels against the
    independent scalar C restatement, configuration and tracked sums."""
    from oracle import cbuild
    c = _cabi()
    co = cbuild.load()
    L, seed, kT, nsweeps = 4096, 808, 2.269185314213022, 3
    thr = o_ising.threshold_table(-1.0, kT)[1]
    ref = co.init(L, L, seed, 0)
    B, M = co.observables(ref)
    lats = []
    for layout, kernel in ((INT8, c.KERNEL_AUTO), (INT8, c.KERNEL_TWO_PASS), (BITS, c.KERNEL_AUTO)):
        lat = c.Lattices(c.MODEL_ISING, L, L, seed=seed, layout=layout, kernel=kernel)
        lat.init_random()
        lat.set_accept_tables(thr[None, :])
        lat.sweep(nsweeps)
        lats.append(lat)
    for sw in range(nsweeps):
        _, dB, dM = co.sweep(ref, thr, seed, 0, sw)
        B, M = B + dB, M + dM
    for lat in lats:
        assert np.array_equal(lat.download(0), ref), lat.sweep_kernel_info()
        b, m = lat.tracked()
        assert (int(b[0]), int(m[0])) == (B, M)
