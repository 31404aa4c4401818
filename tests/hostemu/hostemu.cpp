// hostemu.cpp -- drives the device core (update_word, pair_update: py_mc_b200/csrc/pmc_core.cuh, pmc_pair.cuh)
// on the host, one 64-site word at a time, over a natural [lenX][lenY] int8 image.
//
// TEST INFRASTRUCTURE ONLY (see cuda_host_shim.h).  The neighbour words are gathered site by site from the
// image, so the kernels' own tiling and addressing are NOT exercised here (the GPU tests do that); what is
// exercised is everything a kernel calls per word: class masks, plan rows, Philox bit planes, the bit-serial
// comparison, the per-lane tail and the tracked-observable deltas.
#include "cuda_host_shim.h"

#define PMC_WARP 1  // a "warp" of one thread (see cuda_host_shim.h)
#include "../../py_mc_b200/csrc/pmc_pair.cuh"
#include "../../py_mc_b200/csrc/pmc_persist.cuh"
#include "../../py_mc_b200/csrc/pmc_pbatch.cuh"
#include "../../py_mc_b200/csrc/pmc_rowstream.cuh"

#include <cstring>
#include <vector>

using namespace pmc;

namespace {

struct Image {
    int8_t *p;
    int lenX, lenY;
    int at(int i, int j) const {
        i = ((i % lenX) + lenX) % lenX;
        j = ((j % lenY) + lenY) % lenY;
        return p[(size_t)i * lenY + j] > 0;
    }
    int8_t *site(int i, int j) const { return p + (size_t)(((i % lenX) + lenX) % lenX) * lenY + (((j % lenY) + lenY) % lenY); }
};

// word of the values at offset (di, dj) from the sites of word w of `colour` in row i; lane l = site k = 64w + l
uint64_t gather(const Image &im, int colour, int i, int w, int di, int dj) {
    const int nk = im.lenY / 2, q = (i + colour) & 1;
    uint64_t v = 0;
    for (int l = 0; l < 64 && 64 * w + l < nk; ++l) v |= (uint64_t)im.at(i + di, 2 * (64 * w + l) + q + dj) << l;
    return v;
}

uint64_t valid_mask(const Image &im, int w) {
    const int n = im.lenY / 2 - 64 * w;
    return n >= 64 ? ~0ull : ((1ull << n) - 1);
}

template <int MODEL>
void flip_sweep(const Image &im, const uint32_t *thr, uint64_t seed, uint32_t replica, uint64_t sweep, long long *dB,
                long long *dS) {
    uint32_t row[PLAN_WORDS];
    Plan<MODEL>::build_row(thr, row);
    const uint32_t(&rowref)[PLAN_WORDS] = row;
    Plan<MODEL> plan;
    plan.init(ParamRow{rowref});
    const int nw = (im.lenY / 2 + 63) / 64;
    const PhiloxKey key = make_key(seed);
    for (int colour = 0; colour < 2; ++colour)
        for (int i = 0; i < im.lenX; ++i)
            for (int w = 0; w < nw; ++w) {
                const uint64_t S = gather(im, colour, i, w, 0, 0), valid = valid_mask(im, w);
                int b = 0, s2 = 0;
                const uint64_t Snew =
                    update_word<MODEL>(S, gather(im, colour, i, w, -1, 0), gather(im, colour, i, w, 1, 0),
                                       gather(im, colour, i, w, 0, -1), gather(im, colour, i, w, 0, 1), valid, plan,
                                       (uint32_t)(i * nw + w), replica, 2 * sweep + colour, key, b, s2);
                *dB += b;
                *dS += s2;
                const uint64_t flip = (S ^ Snew) & valid;
                const int q = (i + colour) & 1;
                for (int l = 0; l < 64; ++l)
                    if ((flip >> l) & 1ull) {
                        int8_t *s = im.site(i, 2 * (64 * w + l) + q);
                        *s = MODEL == MODEL_ISING ? (int8_t)-*s : (int8_t)(1 - *s);
                    }
            }
}

}  // namespace

extern "C" {

// thr: 5 (Ising, by aligned neighbours) or 10 (lattice gas, 5*c + nn) class thresholds.  In place.
int hostemu_flip_sweep(int model, int8_t *img, int lenX, int lenY, const uint32_t *thr, uint64_t seed, uint32_t replica,
                       uint64_t sweep, long long *dB, long long *dS) {
    if (lenX < 2 || lenY < 2 || (lenX & 1) || (lenY & 1)) return -1;
    Image im = {img, lenX, lenY};
    *dB = *dS = 0;
    if (model == MODEL_ISING)
        flip_sweep<MODEL_ISING>(im, thr, seed, replica, sweep, dB, dS);
    else if (model == MODEL_LATGAS)
        flip_sweep<MODEL_LATGAS>(im, thr, seed, replica, sweep, dB, dS);
    else
        return -1;
    return 0;
}

// thr7[d + 3], d = -3..3.  Bond classes as in pmc_pair.cuh.  In place.
int hostemu_pair_sweep(int8_t *img, int lenX, int lenY, const uint32_t *thr7, uint64_t seed, uint32_t replica,
                       uint64_t sweep, uint32_t class_mask, long long *dB) {
    if (lenX % 4 || lenY % 4) return -1;
    Image im = {img, lenX, lenY};
    uint32_t row[PLAN_WORDS];
    if (!PairPlan::build_row(thr7, row)) return -2;
    const uint32_t(&rowref)[PLAN_WORDS] = row;
    PairPlan plan;
    plan.init(ParamRow{rowref});
    const int nw = (lenY / 2 + 63) / 64;
    const PhiloxKey key = make_key(seed);
    *dB = 0;
    for (int cls = 0; cls < 16; ++cls) {
        if (!((class_mask >> cls) & 1u)) continue;
        const bool horiz = cls < 8;
        const int ri = horiz ? cls >> 2 : (cls - 8) >> 1, cj = horiz ? cls & 3 : (cls - 8) & 1;
        for (int i = ri; i < lenX; i += horiz ? 2 : 4) {
            const int ca = (i + cj) & 1, q = (i + ca) & 1;
            for (int w = 0; w < nw; ++w) {
                uint64_t active = valid_mask(im, w);
                if (horiz)
                    for (int l = 0; l < 64; ++l)
                        if ((2 * (64 * w + l) + q) % 4 != cj) active &= ~(1ull << l);
                const int bi = horiz ? 0 : 1, bj = horiz ? 1 : 0;  // offset of the partner
                const uint64_t S = gather(im, ca, i, w, 0, 0), Bw = gather(im, ca, i, w, bi, bj);
                uint64_t A0, A1, B0, B1;
                if (horiz) {
                    add3(gather(im, ca, i, w, -1, 0), gather(im, ca, i, w, 1, 0), gather(im, ca, i, w, 0, -1), A0, A1);
                    add3(gather(im, ca, i, w, -1, 1), gather(im, ca, i, w, 1, 1), gather(im, ca, i, w, 0, 2), B0, B1);
                } else {
                    add3(gather(im, ca, i, w, -1, 0), gather(im, ca, i, w, 0, -1), gather(im, ca, i, w, 0, 1), A0, A1);
                    add3(gather(im, ca, i, w, 2, 0), gather(im, ca, i, w, 1, -1), gather(im, ca, i, w, 1, 1), B0, B1);
                }
                int b = 0;
                const uint64_t acc = pair_update(S, Bw, A0, A1, B0, B1, active, plan, (uint32_t)(i * nw + w), replica,
                                                 16 * sweep + cls, key, b);
                *dB += b;
                for (int l = 0; l < 64; ++l)
                    if ((acc >> l) & 1ull) {
                        const int j = 2 * (64 * w + l) + q;
                        int8_t *a = im.site(i, j), *p = im.site(i + bi, j + bj);
                        const int8_t t = *a;
                        *a = *p;
                        *p = t;
                    }
            }
        }
    }
    return 0;
}

}  // extern "C"

namespace {

// The persistent kernel's per-thread pieces (pmc_persist.cuh), block after block and thread after thread: checks the
// slab geometry, the halo rows and their exchange buffer -- everything but the barriers themselves.
template <int MODEL>
void persist_sweeps(const Image &im, const uint32_t *thr, uint64_t seed, uint32_t replica, uint64_t sweep0, int nsweeps,
                    int nblocks, int T, long long *dB, long long *dS) {
    uint32_t row[PLAN_WORDS];
    Plan<MODEL>::build_row(thr, row);
    const uint32_t(&rowref)[PLAN_WORDS] = row;
    Plan<MODEL> plan;
    plan.init(ParamRow{rowref});
    const int lenX = im.lenX, nw = im.lenY / 128;
    const PhiloxKey key = make_key(seed);
    std::vector<std::vector<uint64_t>> slab(nblocks);
    std::vector<uint64_t> halo((size_t)2 * nblocks * 2 * nw, 0xDEADBEEFDEADBEEFull);
    auto r0 = [&](int b) { return persist_row0(b, lenX, nblocks); };
    for (int b = 0; b < nblocks; ++b) {  // load: own rows and one row either side, both colours
        const int nrows = r0(b + 1) - r0(b);
        slab[b].assign((size_t)2 * (nrows + 2) * nw, 0);
        for (int colour = 0; colour < 2; ++colour)
            for (int lr = -1; lr <= nrows; ++lr)
                for (int w = 0; w < nw; ++w)
                    persist_word(slab[b].data(), nrows, nw, colour, lr, w) =
                        gather(im, colour, ((r0(b) + lr) % lenX + lenX) % lenX, w, 0, 0);
    }
    for (int sw = 0; sw < nsweeps; ++sw)
        for (int colour = 0; colour < 2; ++colour) {
            for (int b = 0; b < nblocks; ++b)
                for (int t = 0; t < T; ++t)
                    persist_pass<MODEL>(slab[b].data(), r0(b + 1) - r0(b), nw, r0(b), colour, plan, replica,
                                        2 * (sweep0 + sw) + colour, key, t, T, *dB, *dS);
            for (int b = 0; b < nblocks; ++b)
                for (int t = 0; t < T; ++t)
                    persist_publish(slab[b].data(), r0(b + 1) - r0(b), nw, colour, b, nblocks, halo.data(), t, T);
            for (int b = 0; b < nblocks; ++b)
                for (int t = 0; t < T; ++t)
                    persist_collect(slab[b].data(), r0(b + 1) - r0(b), nw, colour, b, nblocks, halo.data(), t, T);
        }
    for (int b = 0; b < nblocks; ++b) {  // store
        const int nrows = r0(b + 1) - r0(b);
        for (int colour = 0; colour < 2; ++colour)
            for (int lr = 0; lr < nrows; ++lr)
                for (int w = 0; w < nw; ++w) {
                    const uint64_t v = persist_word(slab[b].data(), nrows, nw, colour, lr, w);
                    const int i = r0(b) + lr, q = (i + colour) & 1;
                    for (int l = 0; l < 64; ++l) {
                        const int bit = (int)((v >> l) & 1ull);
                        *im.site(i, 2 * (64 * w + l) + q) = MODEL == MODEL_ISING ? (int8_t)(2 * bit - 1) : (int8_t)bit;
                    }
                }
    }
}

}  // namespace

extern "C" int hostemu_persist_sweeps(int model, int8_t *img, int lenX, int lenY, const uint32_t *thr, uint64_t seed,
                                      uint32_t replica, uint64_t sweep0, int nsweeps, int nblocks, int nthreads,
                                      long long *dB, long long *dS) {
    if (lenY % 128 || (lenX & 1) || nblocks < 1 || nblocks > lenX || nthreads < 1) return -1;
    Image im = {img, lenX, lenY};
    *dB = *dS = 0;
    if (model == MODEL_ISING)
        persist_sweeps<MODEL_ISING>(im, thr, seed, replica, sweep0, nsweeps, nblocks, nthreads, dB, dS);
    else if (model == MODEL_LATGAS)
        persist_sweeps<MODEL_LATGAS>(im, thr, seed, replica, sweep0, nsweeps, nblocks, nthreads, dB, dS);
    else
        return -1;
    return 0;
}

namespace {

// The batch-resident kernel's per-thread pieces (pmc_pbatch.cuh) on the host: B emulated blocks in lock step, thread
// after thread.  Checks the geometry (which rows a block takes from / gives to which other block), the export-first
// item order, the flag bookkeeping (a wait that the lock-step order cannot satisfy is reported as an error), the
// per-replica plans and the exact sums of pb_reduce.  imgs: [nrep][lenX][lenY]; thr: [nrep][nclass] (each replica at
// its own temperature); sums: [nrep][2] out.
template <int MODEL>
int pbatch_sweeps(int8_t *imgs, int nrep, int lenX, int lenY, const uint32_t *thr, int nclass, uint64_t seed, uint32_t rep_offset,
                  uint64_t sweep0, int nsweeps, int B, int T, long long *sums) {
    const int nw = lenY / 128;
    const size_t plane = (size_t)lenX * lenY;
    std::vector<PBGeom> geo(B);
    std::vector<std::vector<uint64_t>> rows(B);
    std::vector<std::vector<int>> segoff(B), segstart(B);
    std::vector<std::vector<uint32_t>> splan(B);
    const int nslots = pb_nslots(B, nrep);
    std::vector<PBPacket> xbuf((size_t)2 * nslots * nw, PBPacket{0, 0});
    int err = 0;
    unsigned int abort_word = 0;
    const PhiloxKey keys = make_key(seed);
    std::vector<PBCtx> ctx(B);
    for (int b = 0; b < B; ++b) {
        PBGeom &g = geo[b];
        pb_geometry(b, B, nrep, lenX, g);
        if (g.nseg > PB_MAXSEG || g.nxrow * nw > T) return -2;
        rows[b].assign((size_t)2 * g.nrows * nw, 0);
        segoff[b].resize(g.nseg);
        segstart[b].resize(g.nseg);
        splan[b].resize((size_t)g.nseg * PLAN_WORDS);
        for (int k = 0; k < g.nseg; ++k) {
            segstart[b][k] = k == 0 ? 0 : (g.rep_first + k) * lenX - g.R0;
            segoff[b][k] = k == 0 ? g.i0 : -((g.rep_first + k) * lenX - g.R0);
            Plan<MODEL>::build_row(thr + (size_t)(g.rep_first + k) * nclass, splan[b].data() + (size_t)k * PLAN_WORDS);
        }
        for (int colour = 0; colour < 2; ++colour)
            for (int lr = 0; lr < g.nrows; ++lr) {
                const int gr = g.R0 + lr, rep = gr / lenX, i = gr - rep * lenX;
                Image im = {imgs + (size_t)rep * plane, lenX, lenY};
                for (int w = 0; w < nw; ++w) rows[b][((size_t)colour * g.nrows + lr) * nw + w] = gather(im, colour, i, w, 0, 0);
            }
        PBCtx &c = ctx[b];
        c.s = rows[b].data();
        c.segoff = segoff[b].data();
        c.segstart = segstart[b].data();
        c.splan = splan[b].data();
        c.xbuf = c.xbuf_up = c.xbuf_dn = xbuf.data();
        c.epoch = 4096u * 7u;  // any base: the pass numbers of a launch are relative to it
        c.row0 = 0;
        c.wait_cycles = 0;
        c.nblocks = B;
        c.abort_word = &abort_word;
        c.err = &err;
        c.lenX = lenX;
        c.nw = nw;
        c.nw_shift = -1;
        for (int k = 0; k < 20; ++k)
            if ((1 << k) == nw) c.nw_shift = k;
        c.nslots = nslots;
        c.b = b;
        c.rep_offset = rep_offset;
        c.key = &keys;
    }
    std::vector<std::vector<PBI4>> cache_a(B, std::vector<PBI4>(T)), cache_b(B, std::vector<PBI4>(T));
    for (int b = 0; b < B; ++b)
        for (int t = 0; t < T; ++t) {
            pb_publish_loaded(geo[b], ctx[b], t, T);
            PBItems f;
            pb_items(geo[b], ctx[b], t, T, f);
            pb_items_store(f, cache_a[b].data(), cache_b[b].data(), t);
        }
    unsigned int events = 1;
    for (int sw = 0; sw < nsweeps; ++sw)
        for (int colour = 0; colour < 2; ++colour) {
            for (int b = 0; b < B; ++b)
                for (int t = 0; t < T; ++t) {
                    // with the per-thread item data cached (as the kernel does when shared memory has room) in even
                    // blocks, recomputed in odd ones; the plan is set up afresh for every call here
                    Plan<MODEL> plan;
                    int plan_seg = -1;
                    if ((b & 1) == 0)
                        pb_pass<MODEL, true, false>(geo[b], ctx[b], colour, 2 * (sweep0 + sw) + colour, events, t, T, plan, plan_seg,
                                             cache_a[b].data(), cache_b[b].data());
                    else
                        pb_pass<MODEL, false, false>(geo[b], ctx[b], colour, 2 * (sweep0 + sw) + colour, events, t, T, plan, plan_seg,
                                              nullptr, nullptr);
                }
            ++events;
        }
    for (int r = 0; r < 2 * nrep; ++r) sums[r] = 0;
    for (int b = 0; b < B; ++b) {
        long long acc[2 * PB_MAXSEG] = {0};
        for (int t = 0; t < T; ++t) pb_reduce<MODEL>(geo[b], ctx[b], events, acc, t, T);
        for (int k = 0; k < geo[b].nseg; ++k) {
            sums[2 * (geo[b].rep_first + k)] += acc[2 * k];
            sums[2 * (geo[b].rep_first + k) + 1] += acc[2 * k + 1];
        }
    }
    if (err) return -4;
    for (int b = 0; b < B; ++b) {  // store
        const PBGeom &g = geo[b];
        for (int colour = 0; colour < 2; ++colour)
            for (int lr = 0; lr < g.nrows; ++lr) {
                const int gr = g.R0 + lr, rep = gr / lenX, i = gr - rep * lenX, q = (i + colour) & 1;
                Image im = {imgs + (size_t)rep * plane, lenX, lenY};
                for (int w = 0; w < nw; ++w) {
                    const uint64_t v = rows[b][((size_t)colour * g.nrows + lr) * nw + w];
                    for (int l = 0; l < 64; ++l) {
                        const int bit = (int)((v >> l) & 1ull);
                        *im.site(i, 2 * (64 * w + l) + q) = MODEL == MODEL_ISING ? (int8_t)(2 * bit - 1) : (int8_t)bit;
                    }
                }
            }
    }
    return 0;
}

}  // namespace

extern "C" int hostemu_pbatch_sweeps(int model, int8_t *imgs, int nrep, int lenX, int lenY, const uint32_t *thr, uint64_t seed,
                                     uint32_t rep_offset, uint64_t sweep0, int nsweeps, int nblocks, int nthreads,
                                     long long *sums) {
    if (lenY % 128 || (lenX & 1) || nblocks < 1 || (long long)nblocks > (long long)nrep * lenX || nthreads < 1) return -1;
    if (model == MODEL_ISING)
        return pbatch_sweeps<MODEL_ISING>(imgs, nrep, lenX, lenY, thr, 5, seed, rep_offset, sweep0, nsweeps, nblocks, nthreads, sums);
    if (model == MODEL_LATGAS)
        return pbatch_sweeps<MODEL_LATGAS>(imgs, nrep, lenX, lenY, thr, 10, seed, rep_offset, sweep0, nsweeps, nblocks, nthreads, sums);
    return -1;
}

namespace {

// The row-streaming kernel's schedule and per-thread phases (pmc_rowstream.cuh) on the host: piece after piece,
// iteration after iteration, thread after thread, with the bulk copies replaced by memcpy at the point where the kernel
// issues them.  The warps of a block may be a whole iteration apart on the GPU (a warp starting iteration `it` has only
// waited for every warp to finish iteration it - 2), so besides the result (compared by the caller with the plain sweep)
// the emulation enforces the rule that makes that safe: every ring slot read holds the row the reader means, in the state
// it needs (old / new), written at least TWO iterations earlier; no slot is written that was read or written less than two
// iterations earlier; a row is stored exactly two iterations after it was computed and not overwritten before; a staging
// slot is refilled two iterations after it was packed from.
template <int MODEL, int NWL2>
int rowstream_sweep(Image &im, const uint32_t *thr, uint64_t seed, uint32_t replica, uint64_t sweep, int pieces, long long *dB,
                    long long *dS) {
    const int lenX = im.lenX, nk = im.lenY / 2;
    RSParams p;
    if (!rs_make_params(lenX, nk, pieces, p)) return -2;
    if (p.nw_log2 != NWL2) return -2;
    const int nw = p.nw, RW = p.rw, T = RS_THREADS;
    // colour-split int8 source and destination [2][lenX][nk]
    std::vector<int8_t> src((size_t)2 * lenX * nk), dst((size_t)2 * lenX * nk, 77);
    for (int i = 0; i < lenX; ++i)
        for (int j = 0; j < im.lenY; ++j) src[((size_t)((i + j) & 1) * lenX + i) * nk + (j >> 1)] = im.p[(size_t)i * im.lenY + j];
    uint32_t row[PLAN_WORDS];
    Plan<MODEL>::build_row(thr, row);
    const uint32_t(&rowref)[PLAN_WORDS] = row;
    Plan<MODEL> plan;
    plan.init(ParamRow{rowref});
    const PhiloxKey key = make_key(seed);
    int tdB = 0, tdS = 0;
    for (int piece = 0; piece < pieces; ++piece) {
        int a, n;
        rs_piece(p, piece, a, n);
        if (n < rs_min_rows(p.Q)) return -3;
        const int ntasks = 2 * n + 2, npacks = 2 * n + 6;
        const int niter = p.prologue + ((ntasks + p.Q - 1) >> p.q_log2) + RS_LAG;
        std::vector<uint4> stage((size_t)RS_STAGES * T * 4);
        std::vector<int> stage_group(RS_STAGES, -1), stage_packed_at(RS_STAGES, -100);
        std::vector<uint64_t> W((size_t)(2 * RW + 1) * nw, 0xAAAAAAAAAAAAAAAAull);
        // per ring slot: index held, state (0 old, 1 new), iteration of the last write / of the last read by another task
        struct Slot {
            int idx = -1000, state = -1, wr = -100, rd = -100;
        };
        std::vector<Slot> r0(RW), r1(RW);
        RSCompute ct[RS_MAXQ];
        RSMove pt[RS_MAXQ], ut[RS_MAXQ];
        auto issue_group = [&](int g, int it) -> int {
            const int first = g << p.q_log2;
            if (first >= npacks) return 0;
            const int slot = g % RS_STAGES;
            if (stage_group[slot] >= 0 && stage_packed_at[slot] > it - 2) return -10;  // refilled too early
            const int nvalid = npacks - first < p.Q ? npacks - first : p.Q;
            for (int k = 0; k < nvalid; ++k) {
                const RSTask task = rs_pack_task(p, a, n, first + k);
                const int colour = (task.info >> 16) & 1;
                memcpy(reinterpret_cast<int8_t *>(stage.data()) + (size_t)slot * T * 64 + (size_t)k * nk,
                       src.data() + ((size_t)colour * lenX + task.row) * nk, nk);
            }
            stage_group[slot] = g;
            stage_packed_at[slot] = 1 << 30;  // not packed yet
            return 0;
        };
        for (int g = 0; g < RS_STAGES; ++g)
            if (int rc = issue_group(g, -100)) return rc;
        for (int it = 0; it < niter; ++it) {
            const int kc = it - p.prologue;
            for (int k = 0; k < p.Q; ++k) {
                ct[k] = rs_build_compute(p, a, n, (kc << p.q_log2) + k);
                pt[k] = rs_build_pack(p, n, (it << p.q_log2) + k);
                ut[k] = rs_build_unpack(p, a, n, ((kc - RS_LAG) << p.q_log2) + k);
            }
            if (it >= 2)
                if (int rc = issue_group(it, it)) return rc;
            // ---- store what iteration it - RS_LAG computed
            if (kc >= RS_LAG) {
                for (int q = 0; q < p.Q; ++q) {
                    int colour, idx;
                    if (!rs_task_index(p, n, ((kc - RS_LAG) << p.q_log2) + q, colour, idx)) continue;
                    if (colour == 0 && (idx < 1 || idx > n)) continue;
                    Slot &sl = colour ? r1[(idx + 2) % p.rw] : r0[idx % p.rw];
                    if (sl.idx != (colour ? idx + 2 : idx) || sl.state != 1 || sl.wr != it - RS_LAG) return -20;
                    sl.rd = it;
                }
            }
            if (kc >= 0)
                for (int t = 0; t < T; ++t) rs_unpack<MODEL, NWL2>(ut, W.data(), reinterpret_cast<uint4 *>(dst.data()), t);
            // ---- the tasks of this iteration
            if (kc >= 0 && (kc << p.q_log2) < ntasks) {
                for (int q = 0; q < p.Q; ++q) {  // reads first
                    int colour, idx;
                    if (!rs_task_index(p, n, (kc << p.q_log2) + q, colour, idx)) continue;
                    for (int d = 0; d < 3; ++d) {
                        // c0[m] reads OLD c1h[m .. m+2]; c1[j] reads NEW c0[j .. j+2]
                        Slot &sl = colour ? r0[(idx + d) % p.rw] : r1[(idx + d) % p.rw];
                        if (sl.idx != idx + d || sl.state != (colour ? 1 : 0) || sl.wr > it - 2) return -30 - colour;
                        sl.rd = it;
                    }
                }
                for (int q = 0; q < p.Q; ++q) {  // then the writes
                    int colour, idx;
                    if (!rs_task_index(p, n, (kc << p.q_log2) + q, colour, idx)) continue;
                    Slot &sl = colour ? r1[(idx + 2) % p.rw] : r0[idx % p.rw];
                    if (sl.idx != (colour ? idx + 2 : idx) || sl.state != 0 || sl.wr > it - 2 || sl.rd > it - 2) return -40 - colour;
                    sl.state = 1;
                    sl.wr = it;
                }
                for (int t = 0; t < T; ++t) rs_compute<MODEL, NWL2>(ct, W.data(), plan, replica, sweep, key, t, tdB, tdS);
            }
            // ---- pack
            if ((it << p.q_log2) < npacks) {
                const int slot = it % RS_STAGES;
                if (stage_group[slot] != it) return -50;  // the copy for this group was never issued
                for (int q = 0; q < p.Q; ++q) {
                    int colour, idx;
                    if (!rs_pack_index(n, (it << p.q_log2) + q, colour, idx)) continue;
                    Slot &sl = colour ? r1[idx % p.rw] : r0[idx % p.rw];
                    if (sl.wr > it - 2 || sl.rd > it - 2) return -60;  // still in use by a warp one iteration behind
                    // what it held must be finished: a stored new row, a halo row, or nothing yet
                    if (sl.idx >= 0 && sl.state == 0 && !(colour == 1 && (sl.idx < 2 || sl.idx > n + 1))) return -61;
                    if (sl.idx >= 0 && sl.state == 1 && sl.rd < sl.wr + RS_LAG) return -62;  // ... and stored
                    sl.idx = idx;
                    sl.state = 0;
                    sl.wr = it;
                    sl.rd = -100;
                }
                for (int t = 0; t < T; ++t) rs_pack<MODEL, NWL2>(pt, stage.data() + (size_t)slot * T * 4, W.data(), t);
                stage_packed_at[slot] = it;
            }
        }
    }
    for (size_t k = 0; k < dst.size(); ++k)
        if (dst[k] == 77) return -70;  // a site was never stored
    for (int i = 0; i < lenX; ++i)
        for (int j = 0; j < im.lenY; ++j) im.p[(size_t)i * im.lenY + j] = dst[((size_t)((i + j) & 1) * lenX + i) * nk + (j >> 1)];
    *dB = tdB;
    *dS = tdS;
    return 0;
}

}  // namespace

extern "C" int hostemu_rowstream_sweep(int model, int8_t *img, int lenX, int lenY, const uint32_t *thr, uint64_t seed,
                                       uint32_t replica, uint64_t sweep, int pieces, long long *dB, long long *dS) {
    if (lenY % 128 || (lenX & 1) || pieces < 1) return -1;
    Image im = {img, lenX, lenY};
    *dB = *dS = 0;
    if (model != MODEL_ISING && model != MODEL_LATGAS) return -1;
    int nwl2 = 0;
    while ((1 << nwl2) < lenY / 128) ++nwl2;
    switch (nwl2) {
#define RS_CASE(K)                                                                                                       \
    case K:                                                                                                              \
        if constexpr ((RS_THREADS >> K) >= 1 && (RS_THREADS >> K) <= RS_MAXQ)                                            \
            return model == MODEL_ISING ? rowstream_sweep<MODEL_ISING, K>(im, thr, seed, replica, sweep, pieces, dB, dS) \
                                        : rowstream_sweep<MODEL_LATGAS, K>(im, thr, seed, replica, sweep, pieces, dB, dS); \
        break;
        RS_CASE(2) RS_CASE(3) RS_CASE(4) RS_CASE(5) RS_CASE(6) RS_CASE(7) RS_CASE(8) RS_CASE(9)
#undef RS_CASE
    }
    return -2;
}

extern "C" int hostemu_rowstream_threads(void) { return RS_THREADS; }

namespace {

// One lattice in row slabs over several "GPUs", each running the batch-resident kernel on its owned rows (pmc_pbatch.cuh,
// SLAB instantiation): G slabs of lenX / G rows, B blocks each, every slab with its own exchange buffer; what a slab's
// first / last block exports of row 0 / the last row goes into the exchange buffer of the slab above / below (on the GPUs:
// peer memory), the random stream is indexed by whole-lattice rows.  All blocks of all slabs in lock step.  The caller
// compares the result with the plain sweep of the whole lattice; sums[0..1] = bond sum and second observable added up over
// the slabs (a slab's share of the boundary bonds is its own business).
template <int MODEL>
int pbatch_slab_sweeps(int8_t *img, int lenX, int lenY, const uint32_t *thr, uint64_t seed, uint32_t replica, uint64_t sweep0,
                       int nsweeps, int G, int B, int T, long long *sums) {
    const int nw = lenY / 128, nown = lenX / G, nslots = pb_nslots(B, 1);
    Image im = {img, lenX, lenY};
    struct Slab {
        std::vector<PBGeom> geo;
        std::vector<std::vector<uint64_t>> rows;
        std::vector<int> segoff, segstart;
        std::vector<uint32_t> splan;
        std::vector<PBPacket> xbuf;
        std::vector<PBCtx> ctx;
        std::vector<std::vector<PBI4>> ca, cb;
    };
    std::vector<Slab> sl(G);
    int err = 0;
    unsigned int abort_word = 0;
    const PhiloxKey keys = make_key(seed);
    for (int s = 0; s < G; ++s) {
        Slab &S = sl[s];
        S.geo.resize(B);
        S.rows.resize(B);
        S.segoff.assign(1, 0);
        S.segstart.assign(1, 0);
        S.splan.resize(PLAN_WORDS);
        Plan<MODEL>::build_row(thr, S.splan.data());
        S.xbuf.assign((size_t)2 * nslots * nw, PBPacket{0, 0});
        S.ctx.resize(B);
        S.ca.assign(B, std::vector<PBI4>(T));
        S.cb.assign(B, std::vector<PBI4>(T));
    }
    // per block: segoff[0] = row-in-replica of local row 0 -- differs per block, so every block gets its own copy
    std::vector<std::vector<std::vector<int>>> bsegoff(G, std::vector<std::vector<int>>(B, std::vector<int>(1)));
    for (int s = 0; s < G; ++s)
        for (int b = 0; b < B; ++b) {
            Slab &S = sl[s];
            PBGeom &g = S.geo[b];
            pb_geometry(b, B, 1, nown, g);
            if (g.nxrow * nw > T || B < 2) return -2;
            S.rows[b].assign((size_t)2 * g.nrows * nw, 0);
            bsegoff[s][b][0] = g.i0;
            for (int colour = 0; colour < 2; ++colour)
                for (int lr = 0; lr < g.nrows; ++lr)
                    for (int w = 0; w < nw; ++w)
                        S.rows[b][((size_t)colour * g.nrows + lr) * nw + w] = gather(im, colour, s * nown + g.R0 + lr, w, 0, 0);
            PBCtx &c = S.ctx[b];
            c.s = S.rows[b].data();
            c.segoff = bsegoff[s][b].data();
            c.segstart = S.segstart.data();
            c.splan = S.splan.data();
            c.xbuf = S.xbuf.data();
            c.xbuf_up = sl[(s + G - 1) % G].xbuf.data();
            c.xbuf_dn = sl[(s + 1) % G].xbuf.data();
            c.epoch = 16384u * 3u;
            c.row0 = s * nown;
            c.wait_cycles = 0;
            c.abort_word = &abort_word;
            c.err = &err;
            c.lenX = nown;
            c.nw = nw;
            c.nw_shift = -1;
            for (int k = 0; k < 20; ++k)
                if ((1 << k) == nw) c.nw_shift = k;
            c.nslots = nslots;
            c.b = b;
            c.nblocks = B;
            c.rep_offset = replica;
            c.key = &keys;
        }
    for (int s = 0; s < G; ++s)
        for (int b = 0; b < B; ++b)
            for (int t = 0; t < T; ++t) {
                // the slab instantiation of the publishing step: through pb_export_buffer
                const PBGeom &g = sl[s].geo[b];
                const PBCtx &c = sl[s].ctx[b];
                pb_publish_loaded(g, c, t, T);
                PBItems f;
                pb_items(g, c, t, T, f);
                pb_items_store(f, sl[s].ca[b].data(), sl[s].cb[b].data(), t);
            }
    unsigned int events = 1;
    for (int sw = 0; sw < nsweeps; ++sw)
        for (int colour = 0; colour < 2; ++colour) {
            for (int s = 0; s < G; ++s)
                for (int b = 0; b < B; ++b)
                    for (int t = 0; t < T; ++t) {
                        Plan<MODEL> plan;
                        int plan_seg = -1;
                        if ((b + s) & 1)
                            pb_pass<MODEL, true, true>(sl[s].geo[b], sl[s].ctx[b], colour, 2 * (sweep0 + sw) + colour, events, t, T, plan,
                                                       plan_seg, sl[s].ca[b].data(), sl[s].cb[b].data());
                        else
                            pb_pass<MODEL, false, true>(sl[s].geo[b], sl[s].ctx[b], colour, 2 * (sweep0 + sw) + colour, events, t, T, plan,
                                                        plan_seg, nullptr, nullptr);
                    }
            ++events;
        }
    sums[0] = sums[1] = 0;
    for (int s = 0; s < G; ++s)
        for (int b = 0; b < B; ++b) {
            long long acc[2 * PB_MAXSEG] = {0};
            for (int t = 0; t < T; ++t) pb_reduce<MODEL>(sl[s].geo[b], sl[s].ctx[b], events, acc, t, T);
            sums[0] += acc[0];
            sums[1] += acc[1];
        }
    if (err) return -4;
    for (int s = 0; s < G; ++s)
        for (int b = 0; b < B; ++b) {
            const PBGeom &g = sl[s].geo[b];
            for (int colour = 0; colour < 2; ++colour)
                for (int lr = 0; lr < g.nrows; ++lr) {
                    const int i = s * nown + g.R0 + lr, q = (i + colour) & 1;
                    for (int w = 0; w < nw; ++w) {
                        const uint64_t v = sl[s].rows[b][((size_t)colour * g.nrows + lr) * nw + w];
                        for (int l = 0; l < 64; ++l) {
                            const int bit = (int)((v >> l) & 1ull);
                            *im.site(i, 2 * (64 * w + l) + q) = MODEL == MODEL_ISING ? (int8_t)(2 * bit - 1) : (int8_t)bit;
                        }
                    }
                }
        }
    return 0;
}

}  // namespace

extern "C" int hostemu_pbatch_slab_sweeps(int model, int8_t *img, int lenX, int lenY, const uint32_t *thr, uint64_t seed,
                                          uint32_t replica, uint64_t sweep0, int nsweeps, int nslabs, int nblocks, int nthreads,
                                          long long *sums) {
    if (lenY % 128 || nslabs < 1 || lenX % nslabs || ((lenX / nslabs) & 1) || nblocks < 2 || nblocks > lenX / nslabs || nthreads < 1)
        return -1;
    if (model == MODEL_ISING)
        return pbatch_slab_sweeps<MODEL_ISING>(img, lenX, lenY, thr, seed, replica, sweep0, nsweeps, nslabs, nblocks, nthreads, sums);
    if (model == MODEL_LATGAS)
        return pbatch_slab_sweeps<MODEL_LATGAS>(img, lenX, lenY, thr, seed, replica, sweep0, nsweeps, nslabs, nblocks, nthreads, sums);
    return -1;
}
