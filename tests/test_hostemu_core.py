"""CPU: the kernels' per-word core (py_mc_b200/csrc/pmc_stream.cuh, pmc_core.cuh, pmc_pair.cuh) compiled for the
host with a thin shim (tests/hostemu/) and checked against the fixtures recorded from the unmodified reference and
against the numpy specification -- the same source the GPU executes, tested without a GPU.  The kernels' tiling and
addressing are not covered here (tests/test_gpu_*.py do that)."""
import numpy as np
import pytest

from oracle import ising as o_ising
from oracle import latgas as o_latgas
from oracle import pairmove as o_pair
from py_mc_b200 import accept
from hostemu import HostCore  # tests/hostemu (tests/ is on sys.path through conftest.py)


@pytest.fixture(scope="module")
def core():
    return HostCore()


def ising_table(J, kT):
    """Thresholds by number of aligned neighbours a = 0..4, as py_mc_b200.ising2D builds them."""
    return accept.table([-2.0 * J * 1.0 * float(2 * a - 4) for a in range(5)], kT)


def test_ising_core_reproduces_reference_trajectories(core, golden_ising):
    z, meta = golden_ising
    for m in meta:
        cfg = np.ascontiguousarray(z[m["name"] + "_init"], dtype=np.int8).copy()
        thr = ising_table(m["J"], m["kT"])
        B = o_ising.bond_sum(cfg)
        for sw in range(m["nsweeps"]):
            dB, dM = core.flip_sweep(core.MODEL_ISING, cfg, thr, m["seed"], m["replica"], sw)
            B += dB
            assert np.array_equal(cfg, z[m["name"] + "_configs"][sw]), (m["name"], sw)
            assert B == o_ising.bond_sum(cfg)                       # the tracked bond sum the kernels keep
            assert int(cfg.astype(np.int64).sum()) == z[m["name"] + "_mag"][sw]


def test_latgas_core_reproduces_reference_trajectories(core, golden_latgas):
    z, meta = golden_latgas
    for m in meta:                                                   # includes the hard-core case lg_hardcore
        cfg = np.ascontiguousarray(z[m["name"] + "_init"], dtype=np.int8).copy()
        thr = o_latgas.threshold_table(m["J"], m["Eads"], m["mu"], m["kT"])
        B, N = o_latgas.bond_sum(cfg), int(cfg.sum())
        for sw in range(m["nsweeps"]):
            dB, dN = core.flip_sweep(core.MODEL_LATGAS, cfg, thr, m["seed"], m["replica"], sw)
            B, N = B + dB, N + dN
            assert np.array_equal(cfg, z[m["name"] + "_configs"][sw]), (m["name"], sw)
            assert (B, N) == (o_latgas.bond_sum(cfg), int(cfg.sum()))


@pytest.mark.parametrize("shape,J,kT", [((6, 130), 1.0, 1.7), ((4, 256), -1.0, 2.269185314213022), ((8, 36), -0.37, 0.9),
                                        ((2, 2), -1.0, 1.0), ((4, 128), -1.0, 0.05)])
def test_ising_core_matches_specification_on_word_boundaries(core, shape, J, kT):
    """Partial last words, exactly one / two words per row, both signs of J, thresholds that underflow to 0."""
    seed, replica = 424242, 3
    ref = o_ising.init_config(seed, replica, *shape)
    cfg = ref.copy()
    thr = ising_table(J, kT)
    for sw in range(6):
        o_ising.sweep(ref, J, kT, seed, replica, sw)
        core.flip_sweep(core.MODEL_ISING, cfg, thr, seed, replica, sw)
        assert np.array_equal(cfg, ref), (shape, sw)


@pytest.mark.parametrize("shape,J,kT", [((8, 12), -1.0, 1.0), ((4, 260), 1.0, 0.7), ((12, 132), -0.5, 0.6), ((4, 4), 2.0, 1.5)])
def test_pair_core_matches_specification(core, shape, J, kT):
    seed = 909
    thr = o_pair.threshold_table(J, kT)
    ref = o_latgas.init_config(seed, 0, *shape)
    cfg = ref.copy()
    B = o_latgas.bond_sum(cfg)
    for sw in range(4):
        _, nacc, dB_ref = o_pair.sweep_classes(ref, thr, seed, 0, sw)
        dB = core.pair_sweep(cfg, thr, seed, 0, sw)
        B += dB
        assert nacc > 0 and dB == dB_ref
        assert np.array_equal(cfg, ref), (shape, sw)
        assert B == o_latgas.bond_sum(cfg) and int(cfg.sum()) == int(ref.sum())
    # a subset of the classes
    o_pair.sweep_classes(ref, thr, seed, 0, 4, classes=[1, 6, 9, 14])
    core.pair_sweep(cfg, thr, seed, 0, 4, class_mask=(1 << 1) | (1 << 6) | (1 << 9) | (1 << 14))
    assert np.array_equal(cfg, ref)


@pytest.mark.parametrize("shape,nblocks,nthreads", [((8, 128), 1, 32), ((8, 128), 8, 32), ((10, 256), 3, 64),
                                                    ((22, 384), 7, 32), ((16, 128), 5, 96)])
def test_persistent_kernel_slab_logic_matches_specification(core, shape, nblocks, nthreads):
    """pmc_persist.cuh block after block on the host: uneven slabs, one-row slabs, a single block that is its own
    neighbour, more threads than words; Ising against the numpy specification, lattice gas against it too."""
    seed, replica, J, kT = 99, 2, -1.0, 2.3
    ref = o_ising.init_config(seed, replica, *shape)
    cfg = ref.copy()
    B = o_ising.bond_sum(cfg)
    dB, dM = core.persist_sweeps(core.MODEL_ISING, cfg, ising_table(J, kT), seed, replica, 3, 4, nblocks, nthreads)
    for sw in range(3, 7):
        o_ising.sweep(ref, J, kT, seed, replica, sw)
    assert np.array_equal(cfg, ref)
    assert B + dB == o_ising.bond_sum(cfg) and dM == int(cfg.astype(np.int64).sum()) - int(o_ising.init_config(seed, replica, *shape).astype(np.int64).sum())
    thr = o_latgas.threshold_table(-0.7, -0.2, 0.1, 0.8)
    ref = o_latgas.init_config(seed, replica, *shape)
    cfg = ref.copy()
    core.persist_sweeps(core.MODEL_LATGAS, cfg, thr, seed, replica, 0, 3, nblocks, nthreads)
    for sw in range(3):
        o_latgas.sweep(ref, thr, seed, replica, sw)
    assert np.array_equal(cfg, ref)


@pytest.mark.parametrize("nrep,shape,nblocks,nthreads", [
    (1, (8, 128), 1, 32),        # one block holds everything: no remote rows at all
    (1, (8, 128), 8, 32),        # one row per block: the row above and the row below are both remote
    (1, (22, 384), 7, 32),       # uneven pieces, nw = 3 (not a power of two)
    (3, (10, 256), 4, 64),       # replica boundaries inside blocks: rows 0 / lenX-1 of a replica in another block
    (5, (6, 128), 7, 32),        # a block with pieces of three replicas
    (4, (8, 128), 4, 32),        # blocks == replicas: every lattice wraps inside its own block
    (2, (16, 128), 5, 96),       # more threads than items in a pass
    (6, (4, 256), 24, 32),       # exactly one row per block
])
def test_batch_resident_kernel_logic_matches_specification(core, nrep, shape, nblocks, nthreads):
    """pmc_pbatch.cuh block after block on the host: the rows of ALL replicas cut into nblocks pieces regardless of
    the replica boundaries, export rows first, rows of other blocks through the exchange buffer, one plan per
    replica (every replica at its own temperature), exact sums from the kernel's own reduction.  Ising against the
    numpy specification, the lattice gas too."""
    seed, offset, J = 77, 3, -1.0
    kTs = np.linspace(1.5, 3.5, nrep)
    ref = np.stack([o_ising.init_config(seed, offset + r, *shape) for r in range(nrep)])
    cfg = ref.copy()
    thr = np.stack([ising_table(J, kT) for kT in kTs])
    sums = core.pbatch_sweeps(core.MODEL_ISING, cfg, thr, seed, offset, 2, 3, nblocks, nthreads)
    for r in range(nrep):
        for sw in range(2, 5):
            o_ising.sweep(ref[r], J, kTs[r], seed, offset + r, sw)
        assert np.array_equal(cfg[r], ref[r]), r
        assert int(sums[r, 0]) == o_ising.bond_sum(ref[r]) and int(sums[r, 1]) == int(ref[r].astype(np.int64).sum())
    ref = np.stack([o_latgas.init_config(seed, offset + r, *shape) for r in range(nrep)])
    cfg = ref.copy()
    thr = np.stack([o_latgas.threshold_table(-0.7, -0.2, 0.1, kT) for kT in kTs])
    sums = core.pbatch_sweeps(core.MODEL_LATGAS, cfg, thr, seed, offset, 0, 2, nblocks, nthreads)
    for r in range(nrep):
        for sw in range(2):
            o_latgas.sweep(ref[r], thr[r], seed, offset + r, sw)
        assert np.array_equal(cfg[r], ref[r]), r
        assert int(sums[r, 0]) == o_latgas.bond_sum(ref[r]) and int(sums[r, 1]) == int(ref[r].sum())


@pytest.mark.parametrize("shape,pieces", [
    ((260, 1024), 1),    # 8 words per row, 64 row tasks per iteration: the smallest piece the task order allows
    ((526, 1024), 2),    # two ragged pieces; the halo rows of one are the first / last rows of the other
    ((132, 2048), 1),
    ((68, 4096), 1),
    ((110, 8192), 3),
    ((20, 16384), 1),    # 128 words per row, 4 row tasks per iteration (the headline geometry), shortest piece
    ((62, 16384), 3),
    ((46, 16384), 2),
])
def test_row_streaming_kernel_logic_matches_specification(core, shape, pieces):
    """pmc_rowstream.cuh piece after piece on the host: the task order (colour 1 of row j a fixed distance behind
    colour 0 of row j+1), the packing order running ahead of it, the ring of bit rows, the staging slots and the
    per-iteration tables.  The warps of a block may be a whole iteration apart on the GPU, so the emulation checks
    every read of a ring slot against what the slot holds at that time and refuses to read anything written, or to
    overwrite anything read, less than two iterations earlier (tests/hostemu/hostemu.cpp); the result is compared
    with the numpy specification, sums included."""
    seed, replica, J, kT = 91, 2, -1.0, 2.3
    ref = o_ising.init_config(seed, replica, *shape)
    cfg = ref.copy()
    B0 = o_ising.bond_sum(ref)
    dB = dS = 0
    for sw in range(2):
        b, s = core.rowstream_sweep(core.MODEL_ISING, cfg, ising_table(J, kT), seed, replica, sw, pieces)
        dB += b
        dS += s
        o_ising.sweep(ref, J, kT, seed, replica, sw)
        assert np.array_equal(cfg, ref), sw
    assert B0 + dB == o_ising.bond_sum(ref)
    ref = o_latgas.init_config(seed, replica, *shape)
    cfg = ref.copy()
    thr = o_latgas.threshold_table(-0.7, -0.2, 0.1, 0.8)
    core.rowstream_sweep(core.MODEL_LATGAS, cfg, thr, seed, replica, 0, pieces)
    o_latgas.sweep(ref, thr, seed, replica, 0)
    assert np.array_equal(cfg, ref)


@pytest.mark.parametrize("shape,nslabs,nblocks,nthreads", [
    ((16, 128), 2, 2, 32),      # two slabs: the slab above is the slab below
    ((24, 256), 3, 4, 32),      # three slabs of 8 rows in 4 blocks: every block of two rows
    ((32, 128), 4, 2, 64),
    ((40, 384), 1, 5, 32),      # a single slab: its neighbours are itself (the periodic lattice through the exchange buffer)
])
def test_batch_resident_kernel_on_row_slabs_matches_specification(core, shape, nslabs, nblocks, nthreads):
    """pmc_pbatch.cuh, slab instantiation, with several emulated "GPUs": each slab's first / last block exports row 0 / the
    last row into the exchange buffer of the slab above / below, imports from its own; the random stream is indexed by
    whole-lattice rows.  Must equal the plain sweep of the whole lattice (numpy specification), sums included."""
    seed, replica, J, kT = 123, 1, -1.0, 2.4
    ref = o_ising.init_config(seed, replica, *shape)
    cfg = ref.copy()
    b, m = core.pbatch_slab_sweeps(core.MODEL_ISING, cfg, ising_table(J, kT), seed, replica, 3, 4, nslabs, nblocks, nthreads)
    for sw in range(3, 7):
        o_ising.sweep(ref, J, kT, seed, replica, sw)
    assert np.array_equal(cfg, ref)
    assert b == o_ising.bond_sum(ref) and m == int(ref.astype(np.int64).sum())
