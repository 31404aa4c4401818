// pmc_tiled_i8.cuh -- K1: checkerboard Metropolis pass on the int8 layout, tiled through shared memory.
//
// A block of 256 threads updates a tile of TH rows x TW words (TH*TW = 256, one 64-site word per
// thread) of one colour.  HBM is touched only with coalesced 128-bit accesses: every 16-byte chunk
// of the tile (target colour) and of the TH+2 neighbour rows (other colour) is loaded once by one
// thread, packed to 16 bits and parked in shared memory; after the update the new words are
// unpacked back to bytes and stored with 128-bit stores.  Between the two the update itself is the
// bit-sliced word routine of pmc_core.cuh, so this kernel follows the same random stream -- and so
// the same trajectory -- as the generic and the bit-packed kernels.
//
// Algorithmic traffic per pass: the colour being updated is read and written once (2 x N/2 bytes).
// Actual traffic adds one read of the other colour (N/2 bytes, plus 2/TH halo rows, mostly L2 hits).
#pragma once
#include "pmc_core.cuh"
#include "pmc_pack.cuh"

namespace pmc {

constexpr int TILE_THREADS = 256;
#ifndef PMC_TILED_MINBLOCKS
#define PMC_TILED_MINBLOCKS 5
#endif

// grid = (nw / TW, ceil(lenX / TH), nrep)
template <int MODEL, class LatT>
__global__ void __launch_bounds__(TILE_THREADS, PMC_TILED_MINBLOCKS) sweep_tiled_i8_kernel(LatT L, int colour, uint64_t step, int tw_log2) {
    // packed bits: [0, 256) target tile, [256, 256 + (TH+2)*TW) neighbour rows; as u16 pieces for the
    // chunk-granular phases, as u64 words for the update phase
    __shared__ __align__(16) uint64_t s_words[256 + 288];
    __shared__ uint32_t s_edge[256];
    __shared__ int s_red[2][8];
    uint16_t *s_pieces = reinterpret_cast<uint16_t *>(s_words);

    const int TW = 1 << tw_log2, TH = TILE_THREADS >> tw_log2;
    const int rep = blockIdx.z;
    const int w0 = blockIdx.x << tw_log2;  // first word column of the tile
    const int i0 = blockIdx.y * TH;        // first row of the tile
    const int oc = colour ^ 1;
    const int t = threadIdx.x;
    const int chunks_per_row = TW * 4;     // 16-byte chunks per tile row
    const int cpr_log2 = tw_log2 + 2;

    Plan<MODEL> plan;
    fetch_plan(L, rep, plan);

    const int8_t *tgt = L.i8 + (size_t)(rep * 2 + colour) * L.lenX * L.nk;
    const int8_t *oth = L.i8 + (size_t)(rep * 2 + oc) * L.lenX * L.nk;

    // ---- phase A: coalesced 128-bit loads, pack, park in shared memory ----------------------
    {
        uint4 vt[4];
        uint4 vo[5];
        bool ot[4], oo[5];
#pragma unroll
        for (int j = 0; j < 4; ++j) {  // target tile: TH*TW*4 = 1024 chunks
            const int idx = t + TILE_THREADS * j;
            const int r = idx >> cpr_log2, c = idx & (chunks_per_row - 1);
            const int i = i0 + r;
            ot[j] = i < L.lenX;
            if (ot[j]) vt[j] = __ldg(reinterpret_cast<const uint4 *>(tgt + (size_t)i * L.nk + (size_t)w0 * 64) + c);
        }
        const int nochunks = (TH + 2) * chunks_per_row;  // <= 1152
#pragma unroll
        for (int j = 0; j < 5; ++j) {  // neighbour rows i0-1 .. i0+TH
            const int idx = t + TILE_THREADS * j;
            const int r = idx >> cpr_log2, c = idx & (chunks_per_row - 1);
            int i = i0 - 1 + r;
            oo[j] = idx < nochunks && i <= L.lenX;  // row lenX wraps to 0 (needed when the tile ends the lattice)
            if (i < 0) i += L.lenX;
            if (i >= L.lenX) i -= L.lenX;
            if (oo[j]) vo[j] = __ldg(reinterpret_cast<const uint4 *>(oth + (size_t)i * L.nk + (size_t)w0 * 64) + c);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
            s_pieces[t + TILE_THREADS * j] = ot[j] ? (uint16_t)pack16<MODEL>(vt[j]) : (uint16_t)0;
#pragma unroll
        for (int j = 0; j < 5; ++j)
            if (t + TILE_THREADS * j < nochunks)
                s_pieces[1024 + t + TILE_THREADS * j] = oo[j] ? (uint16_t)pack16<MODEL>(vo[j]) : (uint16_t)0;
        // one extra site per row beyond the tile edge: k = 64*w0 - 1 (rows whose colour starts at
        // column 0) or k = 64*(w0+TW) (rows whose colour starts at column 1), periodic
        if (t < TH) {
            const int i = i0 + t;
            uint32_t e = 0;
            if (i < L.lenX) {
                const int q = (i + colour) & 1;
                int k = q ? 64 * (w0 + TW) : 64 * w0 - 1;
                if (k < 0) k = L.nk - 1;
                if (k >= L.nk) k = 0;
                e = oth[(size_t)i * L.nk + k] > 0;
            }
            s_edge[t] = e;
        }
    }
    __syncthreads();

    // ---- phase B: one word per thread ---------------------------------------------------------
    const int r = t >> tw_log2, w = t & (TW - 1);
    const int i = i0 + r;
    const bool active = i < L.lenX;
    const uint64_t *s_tgt = s_words, *s_oth = s_words + 256;
    const bool q = (i + colour) & 1;
    const uint64_t S = s_tgt[t];
    const uint64_t U = s_oth[r * TW + w], C = s_oth[(r + 1) * TW + w], D = s_oth[(r + 2) * TW + w];
    uint64_t Sd;
    if (!q) {
        const uint64_t in = w > 0 ? (s_oth[(r + 1) * TW + w - 1] >> 63) : (uint64_t)s_edge[r];
        Sd = (C << 1) | in;
    } else {
        const uint64_t in = w < TW - 1 ? (s_oth[(r + 1) * TW + w + 1] & 1ull) : (uint64_t)s_edge[r];
        Sd = (C >> 1) | (in << 63);
    }
    int dB = 0, dS = 0;
    const uint32_t word_idx = (uint32_t)(i * L.nw + w0 + w);
    const uint64_t Snew = update_word<MODEL>(S, U, D, C, Sd, active ? ~0ull : 0ull, plan, word_idx,
                                             (uint32_t)(L.rep_offset + rep), step, L.key, dB, dS);
    s_words[t] = Snew;  // each thread overwrites only the word it read itself
    __syncthreads();

    // ---- phase C: unpack and store with 128-bit stores -----------------------------------------
    int8_t *tgt_w = L.i8 + (size_t)(rep * 2 + colour) * L.lenX * L.nk;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int idx = t + TILE_THREADS * j;
        const int rr = idx >> cpr_log2, c = idx & (chunks_per_row - 1);
        const int ii = i0 + rr;
        if (ii < L.lenX)
            reinterpret_cast<uint4 *>(tgt_w + (size_t)ii * L.nk + (size_t)w0 * 64)[c] = unpack16<MODEL>(s_pieces[idx]);
    }

    // ---- tracked observables: block sum, one atomic pair ---------------------------------------
    dB = __reduce_add_sync(0xFFFFFFFFu, active ? dB : 0);
    dS = __reduce_add_sync(0xFFFFFFFFu, active ? dS : 0);
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
}

}  // namespace pmc
