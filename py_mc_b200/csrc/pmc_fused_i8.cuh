// pmc_fused_i8.cuh -- K1f: one full checkerboard sweep (both colours) of the int8 layout per launch.
//
// The two-pass kernel (pmc_tiled_i8.cuh) reads the other colour again in every pass: 3 bytes of HBM
// traffic per attempt, two load phases and two store phases per sweep.  This kernel works out of
// place (src -> dst, the handle ping-pongs two buffers): a block loads a tile of both colours with
// its halo once, updates colour 0 on the tile plus the ring the second colour needs, updates colour 1
// on the tile from the *new* colour-0 values, and stores both -- every site read once and written
// once per sweep (2 bytes per attempt), half the byte<->bit packing.  The ring is recomputed by the
// neighbouring tile with the same counter-based random numbers, so both tiles agree on it and the
// trajectory is identical to the two-pass kernels'.
//
// Tile = 32 rows x 8 words (512 sites wide) per colour, 256 threads.
//   colour 0 (updated first): rows i0-1 .. i0+32 (34 x 8 = 272 words) + one site per tile row just
//     outside the tile in x (the x-neighbour of a colour-1 edge site), handled as a word with a single
//     valid lane;
//   colour 1: rows i0 .. i0+31 (256 words); its inputs are rows i0-2 .. i0+33 of old colour 1 for the
//     colour-0 update.
// Needs lenX % 32 == 0 and (lenY/128) % 8 == 0; other shapes use the two-pass kernel.
//
// Slab mode (one lattice split by rows over several GPUs, pymc_create_slab): the local arrays carry two
// halo rows on either side, the tiles cover the owned rows only, the random stream is indexed by the row of
// the whole lattice, and phase C also writes the first and last two owned rows into the neighbours' halo
// rows through peer pointers -- compute and halo exchange in one kernel.
#pragma once
#include "pmc_tiled_i8.cuh"

namespace pmc {

constexpr int FUSED_TW = 8, FUSED_TH = 32, FUSED_THREADS = 256;

// Neighbour synchronisation of the slab mode, carried by the sweep kernel itself (null `mine` = off).
// Every rank counts collective steps; after finishing step n it writes n into the flag words its two
// neighbours poll.  The tiles of step n+1 that touch a neighbour (first / last tile row: they read halo
// rows that neighbour wrote during step n and write halo rows it read during step n) wait until that
// neighbour's flag says n; all other tiles start at once.  The last block of the grid to finish publishes
// this rank's step number.  Waits are bounded; a time-out raises *err instead of hanging the GPU.
struct SlabSync {
    unsigned long long *mine;  // [2]: written by the up / down neighbour
    unsigned long long *up, *dn;  // the neighbours' flag words
    unsigned long long wait_epoch, signal_epoch;
    unsigned int *done;  // [2] counters of finished top-row / bottom-row tiles of this launch
    int *err;
};
#ifndef PMC_FUSED_MINBLOCKS
#define PMC_FUSED_MINBLOCKS 4
#endif

// grid = (nw / 8, lenX / 32, nrep)
template <int MODEL, bool SLAB, class LatT>
__global__ void __launch_bounds__(FUSED_THREADS, PMC_FUSED_MINBLOCKS)
    sweep_fused_i8_kernel(LatT L, const int8_t *__restrict__ src, int8_t *__restrict__ dst, uint64_t sweep,
                          int8_t *__restrict__ peer_up, int8_t *__restrict__ peer_dn, SlabSync sync) {
    constexpr int TW = FUSED_TW, TH = FUSED_TH, CPR = TW * 4;  // 32 chunks of 16 bytes per tile row
    constexpr int NB = (TH + 2) * TW, NW = (TH + 4) * TW;      // 272 / 288 words
    constexpr int NBC = NB * 4, NWC = NW * 4;                  // 1088 / 1152 chunks
    __shared__ __align__(16) uint64_t s_b[NB];  // colour 0, rows i0-1 .. i0+TH   (old, then new)
    __shared__ __align__(16) uint64_t s_w[NW];  // colour 1, rows i0-2 .. i0+TH+1 (old, then new on the tile)
    __shared__ uint32_t s_wedge[TH + 2];        // old colour-1 site beside the tile, per colour-0 row
    __shared__ uint32_t s_eitem[TH];            // 5 bits: the colour-0 edge site of each tile row and its neighbours
    __shared__ uint32_t s_bedge[TH];            // new colour-0 edge site of each tile row
    __shared__ int s_red[2][8];
    uint16_t *s_b16 = reinterpret_cast<uint16_t *>(s_b), *s_w16 = reinterpret_cast<uint16_t *>(s_w);

    const int t = threadIdx.x;
    // tile rows are issued boundary first (top, bottom, then the interior) so that in slab mode the rows the
    // neighbours are waiting for leave early
    const int ntr = gridDim.y;
    const int tile_row = !(SLAB && sync.mine) ? (int)blockIdx.y : (blockIdx.y == 0 ? 0 : (blockIdx.y == 1 ? ntr - 1 : (int)blockIdx.y - 1));
    const int rep = blockIdx.z, w0 = blockIdx.x * TW, i0 = (SLAB ? L.own_lo : 0) + tile_row * TH;
    const int lenX = L.lenX, nk = L.nk;
    // row of the whole lattice (random-stream index); differs from the local row only in slab mode
    const int row_shift = L.row_shift, lenXg = L.lenX_global;
    auto grow = [row_shift, lenXg](int i) {
        if (!SLAB) return i;
        int g = i + row_shift;
        return g < 0 ? g + lenXg : (g >= lenXg ? g - lenXg : g);
    };
    Plan<MODEL> plan;
    fetch_plan(L, rep, plan);
    const int8_t *sb = src + (size_t)(rep * 2 + 0) * lenX * nk;
    const int8_t *sw = src + (size_t)(rep * 2 + 1) * lenX * nk;
    auto wrap_row = [lenX](int i) { return i < 0 ? i + lenX : (i >= lenX ? i - lenX : i); };
    auto wrap_k = [nk](int k) { return k < 0 ? k + nk : (k >= nk ? k - nk : k); };

    if (SLAB && sync.mine) {  // slab mode: boundary tiles wait for the neighbour they share rows with
        const bool top = tile_row == 0, bottom = tile_row == ntr - 1;
        if ((top || bottom) && t == 0) {
            volatile unsigned long long *f = sync.mine;
            const long long t0 = clock64();
            while ((top && f[0] < sync.wait_epoch) || (bottom && f[1] < sync.wait_epoch)) {
                if (clock64() - t0 > 4000000000LL) {
                    *sync.err = 1;
                    break;
                }
            }
            __threadfence_system();
        }
        if (top || bottom) __syncthreads();
    }

    // ---- phase A: load both colours with halo (128-bit, coalesced), pack, park --------------------
#pragma unroll
    for (int g = 0; g < 3; ++g) {
        uint4 v[3];
        int id[3];
        bool isb[3], on[3];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            const int idx = t + FUSED_THREADS * (3 * g + j);
            on[j] = idx < NBC + NWC;
            isb[j] = idx < NBC;
            id[j] = isb[j] ? idx : idx - NBC;
            const int r = id[j] >> 5, c = id[j] & (CPR - 1);
            const int i = wrap_row(i0 - (isb[j] ? 1 : 2) + r);
            if (on[j])
                v[j] = __ldg(reinterpret_cast<const uint4 *>((isb[j] ? sb : sw) + (size_t)i * nk + (size_t)w0 * 64) + c);
        }
#pragma unroll
        for (int j = 0; j < 3; ++j)
            if (on[j]) (isb[j] ? s_b16 : s_w16)[id[j]] = (uint16_t)pack16<MODEL>(v[j]);
    }
    if (t < TH + 2) {  // colour-1 site next to the tile in each colour-0 row
        const int i = wrap_row(i0 - 1 + t);
        const int k = wrap_k((i & 1) ? 64 * (w0 + TW) : 64 * w0 - 1);  // colour 0: q = i & 1
        s_wedge[t] = sw[(size_t)i * nk + k] > 0;
    } else if (t >= 64 && t < 64 + TH) {  // the colour-0 site just outside the tile that colour 1 of row i needs
        const int r = t - 64, i = i0 + r;
        const int qw = (i + 1) & 1, qb = i & 1;
        const int ke = wrap_k(qw ? 64 * (w0 + TW) : 64 * w0 - 1);
        const int ks = wrap_k(ke - 1 + 2 * qb);
        const int im = wrap_row(i - 1), ip = wrap_row(i + 1);
        uint32_t bits = (uint32_t)(sb[(size_t)i * nk + ke] > 0);
        bits |= (uint32_t)(sw[(size_t)im * nk + ke] > 0) << 1;
        bits |= (uint32_t)(sw[(size_t)ip * nk + ke] > 0) << 2;
        bits |= (uint32_t)(sw[(size_t)i * nk + ke] > 0) << 3;
        bits |= (uint32_t)(sw[(size_t)i * nk + ks] > 0) << 4;
        s_eitem[r] = bits;
    }
    __syncthreads();

    int dB = 0, dS = 0;
    const uint32_t greplica = (uint32_t)(L.rep_offset + rep);

    // ---- phase B1: colour 0 on rows i0-1 .. i0+TH, then the TH edge sites ---------------------------
#pragma unroll 1
    for (int round = 0; round < 2; ++round) {
        const int item = t + FUSED_THREADS * round;
        if (round == 1 && t >= 64) break;  // items 256..303 live in warps 0 and 1
        uint64_t S = 0, U = 0, D = 0, C = 0, Sd = 0, valid = 0;
        uint32_t word_idx = 0;
        bool owned = false;
        int lane = 0;
        if (item < NB) {
            const int r = item >> 3, w = item & (TW - 1);
            const int i = wrap_row(i0 - 1 + r);
            S = s_b[item];
            U = s_w[r * TW + w];
            C = s_w[(r + 1) * TW + w];
            D = s_w[(r + 2) * TW + w];
            if (!(i & 1)) {
                const uint64_t in = w > 0 ? (s_w[(r + 1) * TW + w - 1] >> 63) : (uint64_t)s_wedge[r];
                Sd = (C << 1) | in;
            } else {
                const uint64_t in = w < TW - 1 ? (s_w[(r + 1) * TW + w + 1] & 1ull) : (uint64_t)s_wedge[r];
                Sd = (C >> 1) | (in << 63);
            }
            valid = ~0ull;
            word_idx = (uint32_t)(grow(i) * L.nw + w0 + w);
            owned = r >= 1 && r <= TH;
        } else if (item < NB + TH) {
            const int r = item - NB, i = i0 + r;
            const int qw = (i + 1) & 1;
            lane = qw ? 0 : 63;
            int we = qw ? w0 + TW : w0 - 1;
            we = we < 0 ? we + L.nw : (we >= L.nw ? we - L.nw : we);
            const uint64_t bits = s_eitem[r];
            S = (bits & 1) << lane;
            U = ((bits >> 1) & 1) << lane;
            D = ((bits >> 2) & 1) << lane;
            C = ((bits >> 3) & 1) << lane;
            Sd = ((bits >> 4) & 1) << lane;
            valid = 1ull << lane;
            word_idx = (uint32_t)(grow(i) * L.nw + we);
        }
        int b, s2;
        const uint64_t Snew = update_word<MODEL>(S, U, D, C, Sd, valid, plan, word_idx, greplica, 2 * sweep, L.key, b, s2);
        if (item < NB) {
            s_b[item] = Snew;
            if (owned) {
                dB += b;
                dS += s2;
            }
        } else if (item < NB + TH) {
            s_bedge[item - NB] = (uint32_t)((Snew >> lane) & 1ull);
        }
    }
    __syncthreads();

    // ---- phase B2: colour 1 on the tile, from the new colour-0 words ------------------------------------
    {
        const int r = t >> 3, w = t & (TW - 1);
        const int i = i0 + r;
        const uint64_t S = s_w[(r + 2) * TW + w];
        const uint64_t U = s_b[r * TW + w], C = s_b[(r + 1) * TW + w], D = s_b[(r + 2) * TW + w];
        uint64_t Sd;
        if (!((i + 1) & 1)) {
            const uint64_t in = w > 0 ? (s_b[(r + 1) * TW + w - 1] >> 63) : (uint64_t)s_bedge[r];
            Sd = (C << 1) | in;
        } else {
            const uint64_t in = w < TW - 1 ? (s_b[(r + 1) * TW + w + 1] & 1ull) : (uint64_t)s_bedge[r];
            Sd = (C >> 1) | (in << 63);
        }
        int b, s2;
        const uint64_t Snew = update_word<MODEL>(S, U, D, C, Sd, ~0ull, plan, (uint32_t)(grow(i) * L.nw + w0 + w), greplica,
                                                 2 * sweep + 1, L.key, b, s2);
        s_w[(r + 2) * TW + w] = Snew;
        dB += b;
        dS += s2;
    }
    __syncthreads();

    // ---- phase C: unpack the tile of both colours, 128-bit stores ------------------------------------------
    int8_t *db = dst + (size_t)(rep * 2 + 0) * lenX * nk;
    int8_t *dw = dst + (size_t)(rep * 2 + 1) * lenX * nk;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int idx = t + FUSED_THREADS * j;
        const int r = idx >> 5, c = idx & (CPR - 1);
        const size_t off = (size_t)(i0 + r) * nk + (size_t)w0 * 64;
        const uint4 vb = unpack16<MODEL>(s_b16[(r + 1) * CPR + c]), vw = unpack16<MODEL>(s_w16[(r + 2) * CPR + c]);
        reinterpret_cast<uint4 *>(db + off)[c] = vb;
        reinterpret_cast<uint4 *>(dw + off)[c] = vw;
        // Slab mode: my first / last two rows are the neighbours' halo rows.  Store them straight into the
        // neighbour's destination buffer over NVLink (peer mapping), so no separate exchange step exists.
        const int ro = i0 + r - L.own_lo, nown = L.own_hi - L.own_lo;
        int8_t *peer = nullptr;
        int prow = 0;
        if (!SLAB) {
        } else if (peer_up && ro < 2) {
            peer = peer_up;
            prow = L.own_hi + ro;  // the up neighbour's lower halo
        } else if (peer_dn && ro >= nown - 2) {
            peer = peer_dn;
            prow = ro - (nown - 2);  // the down neighbour's upper halo
        }
        if (peer) {
            const size_t poff = (size_t)prow * nk + (size_t)w0 * 64;
            reinterpret_cast<uint4 *>(peer + (size_t)(rep * 2 + 0) * lenX * nk + poff)[c] = vb;
            reinterpret_cast<uint4 *>(peer + (size_t)(rep * 2 + 1) * lenX * nk + poff)[c] = vw;
        }
    }

    dB = __reduce_add_sync(0xFFFFFFFFu, dB);
    dS = __reduce_add_sync(0xFFFFFFFFu, dS);
    if ((t & 31) == 0) {
        s_red[0][t >> 5] = dB;
        s_red[1][t >> 5] = dS;
    }
    __syncthreads();
    if (t < 2) {
        int v = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) v += s_red[t][k];
        if (v) atomicAdd(L.tracked + 2 * (size_t)rep + t, (unsigned long long)(long long)v);
    }
    if (SLAB && sync.mine && (tile_row == 0 || tile_row == ntr - 1)) {
        // Only the boundary tiles read my halo rows and write the neighbours': when the last tile of the top
        // (bottom) row is done, the up (down) neighbour may go on -- usually long before this kernel ends.
        __threadfence_system();
        __syncthreads();
        if (t == 0) {
            const unsigned int per_row = gridDim.x * gridDim.z;
            if (tile_row == 0 && atomicAdd(sync.done + 0, 1u) == per_row - 1) {
                sync.done[0] = 0;
                __threadfence_system();
                *(volatile unsigned long long *)(sync.up + 1) = sync.signal_epoch;
            }
            if (tile_row == ntr - 1 && atomicAdd(sync.done + 1, 1u) == per_row - 1) {
                sync.done[1] = 0;
                __threadfence_system();
                *(volatile unsigned long long *)(sync.dn + 0) = sync.signal_epoch;
            }
        }
    }
}

}  // namespace pmc
