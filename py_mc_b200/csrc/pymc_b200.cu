// pymc_b200.cu -- kernels and C ABI of the B200-native Metropolis sweep engine (sm_100a).
//
// Data layout in HBM (per handle = per GPU):
//   int8 layout : spins[replica][colour][lenX][nk]   int8,  nk = lenY/2
//   bits layout : bits [replica][colour][lenX][nw]   u64,   nw = nk/64
// colour = (x + y) & 1; within a row the sites of one colour are numbered k = y >> 1, and 64
// consecutive k form a *word* (the unit of the random stream, pmc_stream.cuh).
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

#include "../../include/pymc_b200.h"
#include "pmc_core.cuh"
#include "pmc_tiled_i8.cuh"
#include "pmc_rx.cuh"
#include "pmc_layout.cuh"
#include "pmc_fused_i8.cuh"
#include "pmc_pair.cuh"
#include "pmc_resident.cuh"
#include "pmc_persist.cuh"
#include "pmc_pbatch.cuh"
#include "pmc_rowstream.cuh"
#include "pmc_discs.cuh"

// largest bit image (bytes) for which PYMC_KERNEL_AUTO takes the resident kernel: several blocks per SM
static const size_t RESIDENT_AUTO_MAX_IMAGE = 32 * 1024;
// smallest single lattice (sites) for which PYMC_KERNEL_AUTO takes the persistent kernel K11 on calls of >= 2 sweeps.
// Measured (profiles/r2_persist_probe.json): 6.2 / 7.2 / 11.7 us per sweep at 1024^2 / 2048^2 / 4096^2 against 8.2 /
// 8.3 / 16.5 for the fused kernel; its two grid barriers cost 2.9 us per colour pass.  From PBATCH_SINGLE_MIN_SITES
// on the batch-resident kernel K12 (no grid barrier) is faster still and takes over.
static const size_t PERSIST_AUTO_MIN_SITES = (size_t)1024 * 1024;
// smallest batch (sites of all replicas together) for which PYMC_KERNEL_AUTO takes the batch-resident kernel K12
// on calls of >= 2 sweeps.  Since its rows travel between blocks as {data, pass number} packets it is the fastest
// kernel at every size measured (profiles/r2_pbatch_probe.json, us per sweep K12 / K11 / per-sweep kernels: 128^2
// 3.7 / 5.7 / 12.1, 1024^2 4.8 / 6.4 / 8.2, 4096^2 8.0 / 11.6 / 17.2, 16384^2 103.7 / 104.3 / 170.0), so the
// threshold only keeps it away from lattices of a few rows, where a launch has nothing to amortise.
static const size_t PBATCH_AUTO_MIN_SITES = (size_t)128 * 128;
static const size_t PBATCH_SINGLE_MIN_SITES = (size_t)128 * 128;
// PYMC_KERNEL_AUTO runs the sweeps the resident kernels do not take (single-sweep calls, lattices beyond the chip) on the
// row-streaming kernel K13 instead of the tiled one from this many sites (all replicas together) and this many words per
// colour row on.  Measured (profiles/r2_rowstream_probe.json, ms per sweep K13 / tiled): 16384^2 0.1592 / 0.1761,
// 16 x 4096^2 0.1653 / 0.1761; below that the pieces (at least 4Q + 4 rows each) no longer fill two blocks per SM (8192^2
// 0.059 / 0.050, 4096^2 0.034 / 0.017) and narrow rows put too many row tasks into one iteration (256 x 1024^2: 0.28 / 0.17).
static const size_t ROWSTREAM_AUTO_MIN_SITES = (size_t)1 << 27;
static const int ROWSTREAM_AUTO_MIN_NW = 32;
#ifndef PMC_PERSIST_AUTO
#define PMC_PERSIST_AUTO 1  // 0: only PYMC_KERNEL_PERSISTENT selects the persistent kernel
#endif

using namespace pmc;

// ------------------------------------------------------------------------------------------
// error plumbing
// ------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

static int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

#define CUDA_TRY(expr)                                                                                  \
    do {                                                                                                \
        cudaError_t _e = (expr);                                                                        \
        if (_e != cudaSuccess)                                                                          \
            return fail(PYMC_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                        __LINE__);                                                                      \
    } while (0)

#define REQUIRE(cond, code, ...) \
    do {                         \
        if (!(cond)) return fail(code, __VA_ARGS__); \
    } while (0)

// ------------------------------------------------------------------------------------------
// lattice descriptor passed by value to kernels
// ------------------------------------------------------------------------------------------
struct Lat {
    int8_t *i8;
    uint64_t *bits;
    int lenX, lenY, nk, nw;
    int nrep, rep_offset;
    const uint32_t *plans;     // [ntemps][PLAN_WORDS] per-temperature plan rows (pmc_core.cuh)
    const uint32_t *pair_plans;  // [ntemps][PLAN_WORDS] rows for the pair-exchange kernel (pmc_pair.cuh)
    const int32_t *rep_to_T;   // [nrep_global]
    unsigned long long *tracked;  // [nrep][2]
    PhiloxKey key;
    // Row-slab mode (one lattice split over GPUs): the local arrays hold the owned rows [own_lo, own_hi)
    // plus two halo rows on either side; global row = local row + row_shift (mod lenX_global).
    // Whole-lattice mode: own_lo = 0, own_hi = lenX, row_shift = 0, lenX_global = lenX.
    int own_lo, own_hi, row_shift, lenX_global;
    // Single-temperature handles: the plan row travels in the kernel parameters (constant memory).
    int single_plan;
    uint32_t plan0[PLAN_WORDS];
};

// Plan of the replica this block works on.
template <class PlanT>
__device__ __forceinline__ void fetch_plan(const Lat &L, int rep, PlanT &plan) {
    if (L.single_plan) {
        plan.init(ParamRow{L.plan0});
    } else {
        plan.init(GlobalRow{L.plans + (size_t)L.rep_to_T[L.rep_offset + rep] * PLAN_WORDS});
    }
}

template <int LAYOUT>
struct Loader;

// one byte per site
template <>
struct Loader<PYMC_LAYOUT_INT8> {
    static __device__ __forceinline__ const int8_t *row(const Lat &L, int rep, int colour, int i) {
        return L.i8 + ((size_t)(rep * 2 + colour) * L.lenX + i) * L.nk;
    }
    static __device__ __forceinline__ uint64_t word(const Lat &L, int rep, int colour, int i, int w) {
        const int8_t *p = row(L, rep, colour, i) + 64 * w;
        if ((L.nk & 63) == 0) {  // whole words, 16-byte aligned: four 128-bit loads, multiply-gather packing
            const uint4 *q = reinterpret_cast<const uint4 *>(p);
            uint64_t m = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint4 v = __ldg(q + j);
                // sign bit set <=> -1 (Ising); 0/1 bytes of the lattice gas have bit 0 set <=> occupied.
                // "p[l] > 0" for both models: byte is 0x01.
                const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
                uint32_t r = 0;
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const uint32_t pos = w4[e] & ~(w4[e] >> 7) & 0x01010101u;  // bit 0 set and sign clear
                    r |= ((pos * 0x10204080u) >> 28) << (4 * e);
                }
                m |= (uint64_t)r << (16 * j);
            }
            return m;
        }
        int n = min(64, L.nk - 64 * w);
        uint64_t m = 0;
        for (int l = 0; l < n; ++l) m |= (uint64_t)(p[l] > 0) << l;
        return m;
    }
    static __device__ __forceinline__ int site(const Lat &L, int rep, int colour, int i, int k) {
        return row(L, rep, colour, i)[k] > 0;
    }
    // flip the occupancy (0 <-> 1) of the lanes in mask / of one site
    static __device__ __forceinline__ void toggle(const Lat &L, int rep, int colour, int i, int w, uint64_t mask) {
        int8_t *p = const_cast<int8_t *>(row(L, rep, colour, i)) + 64 * w;
        while (mask) {
            const int l = __ffsll((long long)mask) - 1;
            mask &= mask - 1;
            p[l] = (int8_t)(1 - p[l]);
        }
    }
    static __device__ __forceinline__ void toggle_site(const Lat &L, int rep, int colour, int i, int k) {
        int8_t *p = const_cast<int8_t *>(row(L, rep, colour, i)) + k;
        *p = (int8_t)(1 - *p);
    }
};

// one bit per site
template <>
struct Loader<PYMC_LAYOUT_BITS> {
    static __device__ __forceinline__ uint64_t word(const Lat &L, int rep, int colour, int i, int w) {
        return L.bits[((size_t)(rep * 2 + colour) * L.lenX + i) * L.nw + w];
    }
    static __device__ __forceinline__ int site(const Lat &L, int rep, int colour, int i, int k) {
        return (int)((word(L, rep, colour, i, k >> 6) >> (k & 63)) & 1u);
    }
    // atomic because the carry lane of a horizontal bond class lands in a word another thread also updates
    static __device__ __forceinline__ void toggle(const Lat &L, int rep, int colour, int i, int w, uint64_t mask) {
        if (mask) atomicXor((unsigned long long *)(L.bits + ((size_t)(rep * 2 + colour) * L.lenX + i) * L.nw + w), mask);
    }
    static __device__ __forceinline__ void toggle_site(const Lat &L, int rep, int colour, int i, int k) {
        toggle(L, rep, colour, i, k >> 6, 1ull << (k & 63));
    }
};

// word of `colour`, row i, whose lane l holds site k+1 (dir = +1) or k-1 (dir = -1) of word w, periodic
template <int LAYOUT>
__device__ __forceinline__ uint64_t shifted_word(const Lat &L, int rep, int colour, int i, int w, int nvalid, int dir) {
    const uint64_t W = Loader<LAYOUT>::word(L, rep, colour, i, w);
    if (dir > 0) {
        int knext = 64 * w + nvalid;
        if (knext >= L.nk) knext = 0;
        return (W >> 1) | ((uint64_t)Loader<LAYOUT>::site(L, rep, colour, i, knext) << (nvalid - 1));
    }
    int kprev = 64 * w - 1;
    if (kprev < 0) kprev = L.nk - 1;
    return (W << 1) | (uint64_t)Loader<LAYOUT>::site(L, rep, colour, i, kprev);
}

// Neighbour words of word (colour, i, w): U, D, C and the two shifted same-row words
// Lf (site k-1) and Rt (site k+1) of the other colour, periodic in both directions.
template <int LAYOUT>
__device__ __forceinline__ void load_neighbours(const Lat &L, int rep, int colour, int i, int w, int nvalid,
                                                bool needLf, bool needRt, uint64_t &U, uint64_t &D, uint64_t &C,
                                                uint64_t &Lf, uint64_t &Rt) {
    const int oc = colour ^ 1;
    const int im = (i == 0) ? L.lenX - 1 : i - 1;
    const int ip = (i == L.lenX - 1) ? 0 : i + 1;
    U = Loader<LAYOUT>::word(L, rep, oc, im, w);
    D = Loader<LAYOUT>::word(L, rep, oc, ip, w);
    C = Loader<LAYOUT>::word(L, rep, oc, i, w);
    Lf = Rt = 0;
    if (needLf) {
        int kprev = 64 * w - 1;
        if (kprev < 0) kprev = L.nk - 1;
        Lf = (C << 1) | (uint64_t)Loader<LAYOUT>::site(L, rep, oc, i, kprev);
    }
    if (needRt) {
        int knext = 64 * w + nvalid;
        if (knext >= L.nk) knext = 0;
        Rt = (C >> 1) | ((uint64_t)Loader<LAYOUT>::site(L, rep, oc, i, knext) << (nvalid - 1));
    }
}

__device__ __forceinline__ int warp_sum(int v) { return __reduce_add_sync(0xFFFFFFFFu, v); }

// Block-level sum of two ints, one 64-bit atomic pair per block.
__device__ __forceinline__ void block_accumulate(int a, int b, unsigned long long *dst) {
    __shared__ int sa[32], sb[32];
    a = warp_sum(a);
    b = warp_sum(b);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) {
        sa[wid] = a;
        sb[wid] = b;
    }
    __syncthreads();
    if (wid == 0) {
        const int nwarp = (blockDim.x + 31) >> 5;
        a = lane < nwarp ? sa[lane] : 0;
        b = lane < nwarp ? sb[lane] : 0;
        a = warp_sum(a);
        b = warp_sum(b);
        if (lane == 0) {
            if (a) atomicAdd(dst, (unsigned long long)(long long)a);
            if (b) atomicAdd(dst + 1, (unsigned long long)(long long)b);
        }
    }
}

template <int MODEL>
__device__ __forceinline__ void load_plan(const Lat &L, int rep, Plan<MODEL> &plan) {
    fetch_plan(L, rep, plan);
}

// ------------------------------------------------------------------------------------------
// K1/K2 (first form): one thread per word, any even lenX/lenY (int8) or lenY % 128 == 0 (bits).
// grid = (ceil(lenX*nw / block), nrep)
// ------------------------------------------------------------------------------------------
template <int LAYOUT, int MODEL>
__global__ void __launch_bounds__(128) sweep_word_kernel(Lat L, int colour, uint64_t step) {
    const int rep = blockIdx.y;
    const int word = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = word < L.lenX * L.nw;
    Plan<MODEL> plan;
    load_plan<MODEL>(L, rep, plan);

    int i = 0, w = 0, nvalid = 0;
    uint64_t S = 0, U = 0, D = 0, C = 0, Lf = 0, Rt = 0, valid = 0;
    bool q = false;
    if (active) {
        i = word / L.nw;
        w = word - i * L.nw;
        nvalid = min(64, L.nk - 64 * w);
        valid = nvalid == 64 ? ~0ull : ((1ull << nvalid) - 1);
        q = (i + colour) & 1;  // first column of this colour in row i
        S = Loader<LAYOUT>::word(L, rep, colour, i, w);
        load_neighbours<LAYOUT>(L, rep, colour, i, w, nvalid, !q, q, U, D, C, Lf, Rt);
    }
    int dB = 0, dS = 0;
    uint64_t Snew = update_word<MODEL>(S, U, D, C, q ? Rt : Lf, valid, plan, (uint32_t)word,
                                       (uint32_t)(L.rep_offset + rep), step, L.key, dB, dS);
    if (active) {
        uint64_t flip = (S ^ Snew) & valid;
        if (LAYOUT == PYMC_LAYOUT_BITS) {
            if (flip) L.bits[((size_t)(rep * 2 + colour) * L.lenX + i) * L.nw + w] = Snew;
        } else {
            int8_t *p = L.i8 + ((size_t)(rep * 2 + colour) * L.lenX + i) * L.nk + 64 * w;
            while (flip) {
                int l = __ffsll((long long)flip) - 1;
                flip &= flip - 1;
                p[l] = (MODEL == MODEL_ISING) ? (int8_t)-p[l] : (int8_t)(1 - p[l]);
            }
        }
    } else {
        dB = dS = 0;
    }
    block_accumulate(dB, dS, L.tracked + 2 * (size_t)rep);
}

// ------------------------------------------------------------------------------------------
// K2: bit-packed layout, one thread walks RPT consecutive rows of one word column so that the
// neighbour rows are loaded once and the per-thread set-up (plan, index arithmetic) is shared.
// All loads of a thread are issued before the first update.  grid = (ceil(lenX/RPT*nw / block), nrep)
// ------------------------------------------------------------------------------------------
template <int MODEL, int RPT>
__global__ void __launch_bounds__(128) sweep_bits_rows_kernel(Lat L, int colour, uint64_t step) {
    const int rep = blockIdx.y;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int ngroups = L.lenX / RPT;
    const bool active = idx < ngroups * L.nw;
    Plan<MODEL> plan;
    load_plan<MODEL>(L, rep, plan);
    const int g = active ? idx / L.nw : 0;
    const int w = active ? idx - g * L.nw : 0;
    const int i0 = g * RPT;
    const int wl = w == 0 ? L.nw - 1 : w - 1, wr = w == L.nw - 1 ? 0 : w + 1;
    uint64_t *tgt = L.bits + (size_t)(rep * 2 + colour) * L.lenX * L.nw;
    const uint64_t *oth = L.bits + (size_t)(rep * 2 + (colour ^ 1)) * L.lenX * L.nw;

    uint64_t S[RPT], O[RPT + 2], A[RPT];
    {
        const int im = i0 == 0 ? L.lenX - 1 : i0 - 1;
        O[0] = oth[(size_t)im * L.nw + w];
#pragma unroll
        for (int r = 0; r < RPT; ++r) {
            const int i = i0 + r;
            const bool q = (i + colour) & 1;
            O[r + 1] = oth[(size_t)i * L.nw + w];
            A[r] = oth[(size_t)i * L.nw + (q ? wr : wl)];
            S[r] = tgt[(size_t)i * L.nw + w];
        }
        const int ip = i0 + RPT == L.lenX ? 0 : i0 + RPT;
        O[RPT + 1] = oth[(size_t)ip * L.nw + w];
    }
    int dB = 0, dS = 0;
#pragma unroll
    for (int r = 0; r < RPT; ++r) {
        const int i = i0 + r;
        const bool q = (i + colour) & 1;
        const uint64_t C = O[r + 1];
        const uint64_t Sd = q ? ((C >> 1) | (A[r] << 63)) : ((C << 1) | (A[r] >> 63));
        int b, s2;
        const uint64_t Snew = update_word<MODEL>(S[r], O[r], O[r + 2], C, Sd, active ? ~0ull : 0ull, plan,
                                                 (uint32_t)(i * L.nw + w), (uint32_t)(L.rep_offset + rep), step, L.key,
                                                 b, s2);
        if (active && Snew != S[r]) tgt[(size_t)i * L.nw + w] = Snew;
        dB += b;
        dS += s2;
    }
    block_accumulate(active ? dB : 0, active ? dS : 0, L.tracked + 2 * (size_t)rep);
}

// ------------------------------------------------------------------------------------------
// K6: one bond class of the particle-hole exchange sweep (see pmc_pair.cuh).  One thread per word of
// first sites.  grid = (ceil(rows_in_class * nw / block), nrep)
// ------------------------------------------------------------------------------------------
template <int LAYOUT>
__global__ void __launch_bounds__(128) pair_class_kernel(Lat L, int cls, uint64_t step) {
    const int rep = blockIdx.y;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const bool horiz = cls < 8;
    const int ri = horiz ? cls >> 2 : (cls - 8) >> 1, cj = horiz ? cls & 3 : (cls - 8) & 1;
    const int rstride = horiz ? 2 : 4;
    const int nrows = L.lenX / rstride;
    const bool on = idx < nrows * L.nw;
    PairPlan plan;
    plan.init(GlobalRow{L.pair_plans + (size_t)L.rep_to_T[L.rep_offset + rep] * PLAN_WORDS});
    uint64_t S = 0, Bw = 0, A0 = 0, A1 = 0, B0 = 0, B1 = 0, active = 0;
    int i = 0, w = 0, nvalid = 64, ca = 0, oc = 1, q = 0, i1 = 0;
    if (on) {
        const int ridx = idx / L.nw;
        w = idx - ridx * L.nw;
        i = ri + rstride * ridx;
        nvalid = min(64, L.nk - 64 * w);
        const uint64_t valid = nvalid == 64 ? ~0ull : ((1ull << nvalid) - 1);
        ca = (i + cj) & 1;
        oc = ca ^ 1;
        q = cj & 1;  // column parity of the first site = offset of its colour in this row
        const int im = i == 0 ? L.lenX - 1 : i - 1, ip = i == L.lenX - 1 ? 0 : i + 1;
        S = Loader<LAYOUT>::word(L, rep, ca, i, w);
        if (horiz) {
            active = valid & (((cj >> 1) & 1) ? 0xAAAAAAAAAAAAAAAAull : 0x5555555555555555ull);
            const uint64_t Coc = Loader<LAYOUT>::word(L, rep, oc, i, w);
            const uint64_t Ua = Loader<LAYOUT>::word(L, rep, oc, im, w), Da = Loader<LAYOUT>::word(L, rep, oc, ip, w);
            uint64_t La, Ub, Db;
            if (!q) {  // second site at the same k of the other colour; left neighbour of the first at k-1
                Bw = Coc;
                La = shifted_word<LAYOUT>(L, rep, oc, i, w, nvalid, -1);
                Ub = Loader<LAYOUT>::word(L, rep, ca, im, w);
                Db = Loader<LAYOUT>::word(L, rep, ca, ip, w);
            } else {  // second site at k+1; left neighbour of the first at k
                Bw = shifted_word<LAYOUT>(L, rep, oc, i, w, nvalid, +1);
                La = Coc;
                Ub = shifted_word<LAYOUT>(L, rep, ca, im, w, nvalid, +1);
                Db = shifted_word<LAYOUT>(L, rep, ca, ip, w, nvalid, +1);
            }
            const uint64_t Rb = shifted_word<LAYOUT>(L, rep, ca, i, w, nvalid, +1);  // (i, j+2)
            add3(Ua, Da, La, A0, A1);
            add3(Ub, Db, Rb, B0, B1);
        } else {
            active = valid;
            i1 = ip;
            const int i2 = i1 == L.lenX - 1 ? 0 : i1 + 1;
            Bw = Loader<LAYOUT>::word(L, rep, oc, i1, w);
            const uint64_t Ua = Loader<LAYOUT>::word(L, rep, oc, im, w);
            const uint64_t Ca = Loader<LAYOUT>::word(L, rep, oc, i, w);
            const uint64_t Sa = shifted_word<LAYOUT>(L, rep, oc, i, w, nvalid, q ? +1 : -1);
            const uint64_t Db = Loader<LAYOUT>::word(L, rep, ca, i2, w);
            const uint64_t Cb = Loader<LAYOUT>::word(L, rep, ca, i1, w);
            const uint64_t Sb = shifted_word<LAYOUT>(L, rep, ca, i1, w, nvalid, q ? +1 : -1);
            add3(Ua, Ca, Sa, A0, A1);
            add3(Db, Cb, Sb, B0, B1);
        }
    }
    int dB = 0;
    const uint64_t acc = pair_update(S, Bw, A0, A1, B0, B1, active, plan, (uint32_t)(i * L.nw + w),
                                     (uint32_t)(L.rep_offset + rep), step, L.key, dB);
    if (on && acc) {
        Loader<LAYOUT>::toggle(L, rep, ca, i, w, acc);
        if (!horiz) {
            Loader<LAYOUT>::toggle(L, rep, oc, i1, w, acc);
        } else if (!q) {
            Loader<LAYOUT>::toggle(L, rep, oc, i, w, acc);
        } else {
            const uint64_t top = 1ull << (nvalid - 1);
            const uint64_t valid = nvalid == 64 ? ~0ull : ((1ull << nvalid) - 1);
            Loader<LAYOUT>::toggle(L, rep, oc, i, w, (acc << 1) & valid);
            if (acc & top) {
                int knext = 64 * w + nvalid;
                if (knext >= L.nk) knext = 0;
                Loader<LAYOUT>::toggle_site(L, rep, oc, i, knext);
            }
        }
    }
    block_accumulate(on ? dB : 0, 0, L.tracked + 2 * (size_t)rep);
}

// ------------------------------------------------------------------------------------------
// K3: exact integer observables per replica (bond sum and magnetisation / particle number).
// Each word of either colour contributes its bonds to the row above and to the site on its left.
// grid = (ceil(2*lenX*nw / block), nrep)
// ------------------------------------------------------------------------------------------
template <int LAYOUT, int MODEL>
__global__ void __launch_bounds__(256) reduce_kernel(Lat L, unsigned long long *out) {
    const int rep = blockIdx.y;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int per_colour = (L.own_hi - L.own_lo) * L.nw;
    int dB = 0, dS = 0;
    if (idx < 2 * per_colour) {
        const int colour = idx / per_colour;
        const int word = idx - colour * per_colour;
        const int i = L.own_lo + word / L.nw, w = word % L.nw;
        const int nvalid = min(64, L.nk - 64 * w);
        const uint64_t valid = nvalid == 64 ? ~0ull : ((1ull << nvalid) - 1);
        const bool q = (i + colour) & 1;
        uint64_t S = Loader<LAYOUT>::word(L, rep, colour, i, w), U, D, C, Lf, Rt;
        load_neighbours<LAYOUT>(L, rep, colour, i, w, nvalid, !q, false, U, D, C, Lf, Rt);
        const uint64_t Lw = q ? C : Lf;  // the site at column y-1
        if (MODEL == MODEL_ISING) {
            dB = 2 * nvalid - 2 * (__popcll((S ^ U) & valid) + __popcll((S ^ Lw) & valid));
            dS = 2 * __popcll(S & valid) - nvalid;
        } else {
            dB = __popcll(S & U & valid) + __popcll(S & Lw & valid);
            dS = __popcll(S & valid);
        }
    }
    block_accumulate(dB, dS, out + 2 * (size_t)rep);
}

// ------------------------------------------------------------------------------------------
// layout conversion: natural int8 [lenX][lenY]  <->  colour-split storage of one replica
// ------------------------------------------------------------------------------------------
template <int LAYOUT>
__global__ void pack_kernel(Lat L, int rep, const int8_t *natural) {
    const int word = blockIdx.x * blockDim.x + threadIdx.x;
    const int colour = blockIdx.y;
    if (word >= L.lenX * L.nw) return;
    const int i = word / L.nw, w = word - i * L.nw;
    const int nvalid = min(64, L.nk - 64 * w);
    const int q = (i + colour) & 1;
    const int8_t *src = natural + (size_t)i * L.lenY + 2 * (64 * w) + q;
    if (LAYOUT == PYMC_LAYOUT_BITS) {
        uint64_t m = 0;
        for (int l = 0; l < nvalid; ++l) m |= (uint64_t)(src[2 * l] > 0) << l;
        L.bits[((size_t)(rep * 2 + colour) * L.lenX + i) * L.nw + w] = m;
    } else {
        int8_t *dst = L.i8 + ((size_t)(rep * 2 + colour) * L.lenX + i) * L.nk + 64 * w;
        for (int l = 0; l < nvalid; ++l) dst[l] = src[2 * l];
    }
}

template <int LAYOUT, int MODEL>
__global__ void unpack_kernel(Lat L, int rep, int8_t *natural) {
    const int word = blockIdx.x * blockDim.x + threadIdx.x;
    const int colour = blockIdx.y;
    if (word >= L.lenX * L.nw) return;
    const int i = word / L.nw, w = word - i * L.nw;
    const int nvalid = min(64, L.nk - 64 * w);
    const int q = (i + colour) & 1;
    int8_t *dst = natural + (size_t)i * L.lenY + 2 * (64 * w) + q;
    if (LAYOUT == PYMC_LAYOUT_BITS) {
        uint64_t m = L.bits[((size_t)(rep * 2 + colour) * L.lenX + i) * L.nw + w];
        for (int l = 0; l < nvalid; ++l) {
            int b = (int)((m >> l) & 1u);
            dst[2 * l] = (MODEL == MODEL_ISING) ? (int8_t)(2 * b - 1) : (int8_t)b;
        }
    } else {
        const int8_t *src = L.i8 + ((size_t)(rep * 2 + colour) * L.lenX + i) * L.nk + 64 * w;
        for (int l = 0; l < nvalid; ++l) dst[2 * l] = src[l];
    }
}

// Hot start from stream domain 1: thread per Philox call = 4 consecutive natural sites.
template <int LAYOUT, int MODEL>
__global__ void init_kernel(Lat L) {
    const int rep = blockIdx.y;
    const size_t call = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t nsites = (size_t)L.lenX * L.lenY;
    if (call * 4 >= nsites) return;
    uint4 o = philox4x32_10((uint32_t)call, (uint32_t)(L.rep_offset + rep), 0u, ctr3(DOMAIN_INIT, 0, 0), L.key);
    const uint32_t x[4] = {o.x, o.y, o.z, o.w};
    for (int e = 0; e < 4; ++e) {
        const size_t site = call * 4 + e;
        if (site >= nsites) break;
        const int i = (int)(site / L.lenY), j = (int)(site - (size_t)i * L.lenY);
        const int colour = (i + j) & 1, k = j >> 1;
        const int up = x[e] >= 0x80000000u;
        if (LAYOUT == PYMC_LAYOUT_BITS) {
            uint64_t *p = L.bits + ((size_t)(rep * 2 + colour) * L.lenX + i) * L.nw + (k >> 6);
            if (up)
                atomicOr((unsigned long long *)p, 1ull << (k & 63));
        } else {
            L.i8[((size_t)(rep * 2 + colour) * L.lenX + i) * L.nk + k] =
                (MODEL == MODEL_ISING) ? (int8_t)(2 * up - 1) : (int8_t)up;
        }
    }
}

// K7: site value and neighbour sum for a list of sites (parity helper for trialstep's dE).
template <int LAYOUT, int MODEL>
__global__ void propose_kernel(Lat L, int rep, const int32_t *xs, const int32_t *ys, int n, int32_t *site_out,
                               int32_t *nn_out) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const int x = xs[t], y = ys[t];
    auto val = [&](int i, int j) {
        int b = Loader<LAYOUT>::site(L, rep, (i + j) & 1, i, j >> 1);
        return (MODEL == MODEL_ISING) ? 2 * b - 1 : b;
    };
    const int xm = x == 0 ? L.lenX - 1 : x - 1, xp = x == L.lenX - 1 ? 0 : x + 1;
    const int ym = y == 0 ? L.lenY - 1 : y - 1, yp = y == L.lenY - 1 ? 0 : y + 1;
    site_out[t] = val(x, y);
    nn_out[t] = val(xm, y) + val(xp, y) + val(x, ym) + val(x, yp);
}

// ------------------------------------------------------------------------------------------
// row-slab mode helpers
// ------------------------------------------------------------------------------------------
// Hot start of a slab including its halo rows: every row is a pure function of the whole-lattice row
// index (stream domain 1), so no communication is needed.  Thread per Philox call = 4 sites of one row.
template <int MODEL>
__global__ void init_slab_kernel(Lat L) {
    const int rep = blockIdx.y;
    const int per_row = L.lenY >> 2;
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)L.lenX * per_row) return;
    const int il = (int)(idx / per_row), j4 = (int)(idx - (size_t)il * per_row);
    int g = il + L.row_shift;
    g = g < 0 ? g + L.lenX_global : (g >= L.lenX_global ? g - L.lenX_global : g);
    const size_t call = (size_t)g * per_row + j4;
    uint4 o = philox4x32_10((uint32_t)call, (uint32_t)(L.rep_offset + rep), 0u, ctr3(DOMAIN_INIT, 0, 0), L.key);
    const uint32_t x[4] = {o.x, o.y, o.z, o.w};
    for (int e = 0; e < 4; ++e) {
        const int j = 4 * j4 + e;
        const int colour = (il + j) & 1, k = j >> 1;
        const int up = x[e] >= 0x80000000u;
        L.i8[((size_t)(rep * 2 + colour) * L.lenX + il) * L.nk + k] = (MODEL == MODEL_ISING) ? (int8_t)(2 * up - 1) : (int8_t)up;
    }
}

// Copy my first / last two owned rows (both colours) into the neighbours' halo rows of the buffer with the
// same index.  grid = (ceil(nk/16 / block), 4 rows, 2*nrep colours)
__global__ void slab_push_halos_kernel(Lat L, const int8_t *mine, int8_t *peer_up, int8_t *peer_dn) {
    const int c16 = blockIdx.x * blockDim.x + threadIdx.x;
    if (c16 >= L.nk / 16) return;
    const int which = blockIdx.y;  // 0,1: first two owned rows -> up neighbour; 2,3: last two -> down neighbour
    const size_t plane = (size_t)blockIdx.z * L.lenX * L.nk;
    const int nown = L.own_hi - L.own_lo;
    int src_row, dst_row;
    int8_t *peer;
    if (which < 2) {
        src_row = L.own_lo + which;
        dst_row = L.own_hi + which;
        peer = peer_up;
    } else {
        src_row = L.own_lo + nown - 2 + (which - 2);
        dst_row = which - 2;
        peer = peer_dn;
    }
    const uint4 v = reinterpret_cast<const uint4 *>(mine + plane + (size_t)src_row * L.nk)[c16];
    reinterpret_cast<uint4 *>(peer + plane + (size_t)dst_row * L.nk)[c16] = v;
}

// Neighbour barrier over peer memory: publish my epoch to both neighbours, wait for theirs.  One thread.
// The wait is bounded: if a neighbour does not arrive, an error flag is raised instead of hanging the GPU.
__global__ void slab_sync_kernel(volatile unsigned long long *my_flags, unsigned long long *up_flags,
                                 unsigned long long *dn_flags, unsigned long long epoch, int *err,
                                 long long timeout_cycles) {
    __threadfence_system();  // my halo stores (previous kernel) are visible before the epoch is
    *(volatile unsigned long long *)(up_flags + 1) = epoch;  // I am the "down" neighbour of the rank above
    *(volatile unsigned long long *)(dn_flags + 0) = epoch;  // and the "up" neighbour of the rank below
    __threadfence_system();
    const long long t0 = clock64();
    while (my_flags[0] < epoch || my_flags[1] < epoch) {
        if (clock64() - t0 > timeout_cycles) {
            *err = 1;
            break;
        }
    }
    __threadfence_system();
}

// K13: pieces per replica.  Every block recomputes two rows of colour 0 and spends about one iteration on its
// prologue, so long pieces are cheap; but the grid should fill whole waves of resident blocks.  Score = wave
// efficiency x 2n / (2n + Q + 2) (n rows per piece, Q row tasks per iteration); 0 if the shape cannot be streamed.
static int rs_choose_pieces(int lenX, int nk, int nrep, int sms, int smem_per_block_optin) {
    RSParams p;
    if (!rs_make_params(lenX, nk, 1, p)) return 0;
    const size_t smem = rs_smem_bytes(p) + 4096;  // + static shared memory and the per-block reserve
    if (smem > (size_t)smem_per_block_optin) return 0;
    int per_sm = (int)((size_t)228 * 1024 / smem);
    if (per_sm > 1024 / RS_THREADS) per_sm = 1024 / RS_THREADS;  // 64 registers per thread
    if (per_sm < 1) per_sm = 1;
    const long slots = (long)sms * per_sm;
    const int maxP = lenX / rs_min_rows(p.Q);
    int best = 0;
    double best_score = 0.0;
    for (int P = 1; P <= maxP && (long)P * nrep <= 64 * slots; ++P) {
        const long blocks = (long)P * nrep, waves = (blocks + slots - 1) / slots;
        const double n = (double)lenX / P;
        const double score = (double)blocks / (double)(waves * slots) * (2.0 * n) / (2.0 * n + p.Q + 2.0);
        if (score > best_score + 1e-9) {
            best_score = score;
            best = P;
        }
    }
    const char *force = getenv("PYMC_ROWSTREAM_PIECES");  // experiments
    if (force && atoi(force) >= 1 && atoi(force) <= maxP) best = atoi(force);
    return best;
}

// ------------------------------------------------------------------------------------------
// handle
// ------------------------------------------------------------------------------------------
struct pymc_handle {
    pymc_params p;
    int nk, nw, nclass;
    cudaStream_t own_stream, stream;
    cudaEvent_t ev0, ev1;
    int8_t *i8;
    int8_t *i8_alt;                // second int8 buffer for the out-of-place fused sweep (allocated on first use)
    uint64_t *bits;
    uint32_t *plans;               // [ntemps][PLAN_WORDS]
    uint32_t plan0[PLAN_WORDS];    // host copy of row 0 (passed by value to kernels when ntemps == 1)
    uint32_t *pair_plans;          // [ntemps][PLAN_WORDS], allocated by pymc_set_pair_tables
    uint64_t pair_counter;         // pair-exchange sweeps done (random-stream position)
    int32_t *rep_to_T;
    unsigned long long *tracked;   // [nrep][2]
    unsigned long long *reduced;   // [nrep][2] output of reduce_kernel
    int8_t *staging;               // natural int8 image of one replica (upload/download)
    int32_t *qbuf;                 // propose scratch: x, y, site, nn
    int qcap;
    // replica exchange (allocated by pymc_rx_setup)
    double *betas;                 // [ntemps]
    double *energies;              // [nrep_global]  all-gathered energies, indexed by global replica
    int32_t *T_to_rep;             // [ntemps]
    unsigned long long *rx_tried, *rx_accepted;  // [ntemps] per lower temperature index of a pair
    double c_bond, c_second;
    double *stats;                 // [ntemps][NSTAT] device-side observable sums (pymc_stats_*)
    unsigned long long *stats_count;  // [ntemps]
    double *series;                // [series_cap][ntemps][2] (E, |M|) by temperature index per sample (pymc_series_*)
    double *series_work;           // [ntemps*2][2][series_cap] ping-pong buffers of the blocking analysis
    double *series_out;            // [ntemps*2] means + [ntemps*2][32] errors
    int series_cap, series_count;
    bool rx_ready;
    uint64_t rx_round;
    // row-slab mode (pymc_create_slab): see Lat
    bool slab;
    int own_lo, own_hi, row_shift, lenX_global, slab_index, nslabs;
    int8_t *peer_up[2], *peer_dn[2];      // neighbours' two lattice buffers (CUDA IPC mappings); [k] pairs with my buffer k
    unsigned long long *flags;            // [2] epochs written by the up / down neighbour
    unsigned long long *peer_up_flags, *peer_dn_flags;
    int *slab_error;                      // device view of err_host[1]: set if a neighbour did not arrive in time
    int8_t *buf[2];                       // the two ping-pong buffers by fixed index (i8 / i8_alt swap)
    uint64_t slab_epoch;
    unsigned int *slab_done;              // block counter used by the in-kernel barrier
    bool dirty_halo;                      // neighbours may still be writing my halo rows (wait before reading them)
    bool local_peers;                     // neighbours are handles of this process (pymc_slab_connect_local)
    bool resident;                 // sweeps run on the whole-lattice-resident kernel (pmc_resident.cuh)
    bool resident_big;             // ... only calls of >= 2 sweeps do (large int8 images: packing must be amortised)
    int resident_threads;
    bool resident_attr[2];         // dynamic shared memory limit raised for this handle's instantiation (flip model)
    // K11 (pmc_persist.cuh): one lattice resident in the shared memory of all SMs
    bool persist;                  // calls of >= 2 sweeps run on the persistent kernel
    int persist_blocks, persist_threads;
    size_t persist_smem;
    unsigned int *persist_counter; // grid barrier arrivals (zeroed before every launch)
    int *persist_err;              // device view of err_host[0]: a block timed out at the barrier
    uint64_t *persist_halo;        // [2][blocks][2][nw]
    bool persist_attr[2];
    // K12 (pmc_pbatch.cuh): all replicas of the handle resident in the shared memory of all SMs
    bool pbatch;                   // calls of >= 2 sweeps (and single-GPU exchange cycles) run on the batch kernel
    int pb_blocks, pb_threads;
    size_t pb_smem;
    int pb_cache_off;              // byte offset of the per-thread item data behind the rows in shared memory (0: none)
    unsigned int *pb_sync;         // [2] grid-barrier counter + abort word (zeroed per launch)
    PBPacket *pb_xbuf;             // [2][2*blocks + 2*nrep][nw] rows exchanged between blocks, as packets (zeroed per launch)
    long long *pb_esum;            // [2 parities][ntemps][2] exact sums per GLOBAL replica at an exchange
    unsigned int *pb_xflags;       // [PB_MAXRANKS] exchange rounds whose sums each other GPU has delivered here
    int rx_nranks, rx_rank;        // > 1 after pymc_rx_connect: the ladder is spread over that many GPUs
    PBPacket *pb_xpk;              // [2 parities][ntemps][2] packets: the sums of the other GPUs' replicas arrive here
    PBPacket *peer_xpk[PB_MAXRANKS];         // CUDA IPC mappings of the other GPUs' pb_xpk
    unsigned int *peer_xflags[PB_MAXRANKS];  // ... and pb_xflags (kept for the layout of the exported handles; unused)
    unsigned int pb_epoch;         // pass-number base of the last launch (grows by PB_EPOCH_STEP per launch)
    PBPacket *peer_xbuf_up, *peer_xbuf_dn;   // slab mode: the neighbours' exchange buffers (CUDA IPC mappings, or mine)
    bool slab_pbatch;              // slab handle whose calls of >= 2 sweeps run on K12 (connected through IPC, or a single slab)
    int32_t *pb_maps;              // [2 maps][2 parities][ntemps] replica <-> temperature maps of the kernel
    bool pb_attr[2];
    // K13 (pmc_rowstream.cuh): one sweep per launch with the rows streamed through shared memory (int8 layout)
    bool rowstream;                // single sweeps (and sweeps the resident kernels do not take) run on the row-streaming kernel
    RSParams rs;
    size_t rs_smem;
    bool rs_attr[2];
    // Time-outs of the in-kernel barriers are raised by the kernels in two ints of mapped pinned host memory, so
    // that every entry point which synchronises and returns data can look at them for free (check_async_errors).
    volatile int *err_host;        // [0] persistent kernel, [1] slab neighbour barrier
    bool tables_set;
    bool dirty;                    // configurations changed since the tracked observables were based
    uint64_t sweep_counter;
};

// K12 launch shape for `nrep` replicas of `lenX` rows each (lenX: the rows a replica has HERE -- the owned rows of a slab):
// blocks, threads and dynamic shared memory into the handle; false if the kernel cannot take the shape.
static bool plan_pbatch(pymc_handle *h, int nrep, int lenX, int sms, int smem, size_t *rows_bytes_out) {
    int coop = 0;
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, h->p.device);
    const long long NR = (long long)nrep * lenX;
    const int blocks = NR < sms ? (int)NR : sms;
    bool ok = coop && h->p.lenY % 128 == 0 && NR < (1ll << 31) / (h->nw > 0 ? h->nw : 1);
    int max_rows = 0, max_x = 0;
    for (int b = 0; ok && b < blocks; ++b) {
        PBGeom g;
        pb_geometry(b, blocks, nrep, lenX, g);
        ok = g.nseg <= PB_MAXSEG;
        if (g.nrows > max_rows) max_rows = g.nrows;
        if (g.nxrow > max_x) max_x = g.nxrow;
    }
    const size_t rows_bytes = (size_t)16 * max_rows * h->nw;
    ok = ok && rows_bytes + 2048 <= (size_t)smem;  // ~1.2 KiB of static shared memory ride along
    const long long items = (long long)max_rows * h->nw;
    const long long npass = (items + PB_MAX_THREADS - 1) / PB_MAX_THREADS;
    long long threads = npass ? ((items + npass - 1) / npass + 31) / 32 * 32 : 32;
    const long long need = ((long long)max_x * h->nw + 31) / 32 * 32;  // all export items sit in the first round
    if (threads < need) threads = need;
    if (threads < 64) threads = 64;
    ok = ok && threads <= PB_MAX_THREADS;
    h->pb_blocks = blocks;
    h->pb_threads = (int)threads;
    h->pb_smem = rows_bytes;
    // 32 bytes per thread of pass-invariant item data behind the rows, where the rows leave room (pmc_pbatch.cuh: PBItems)
    h->pb_cache_off = 0;
    if (ok && rows_bytes > 0 && rows_bytes + (size_t)threads * 32 + 2048 <= (size_t)smem && max_rows < (1 << 29) / (h->nw > 0 ? h->nw : 1) &&
        h->nw < 65536) {
        h->pb_cache_off = (int)rows_bytes;
        h->pb_smem = rows_bytes + (size_t)threads * 32;
    }
    if (rows_bytes_out) *rows_bytes_out = rows_bytes;
    return ok;
}

static Lat make_lat(const pymc_handle *h) {
    Lat L;
    L.i8 = h->i8;
    L.bits = h->bits;
    L.lenX = h->p.lenX;
    L.lenY = h->p.lenY;
    L.nk = h->nk;
    L.nw = h->nw;
    L.nrep = h->p.nreplicas;
    L.rep_offset = h->p.replica_offset;
    L.plans = h->plans;
    L.pair_plans = h->pair_plans;
    L.rep_to_T = h->rep_to_T;
    L.tracked = h->tracked;
    L.key = make_key(h->p.seed);
    L.own_lo = h->slab ? h->own_lo : 0;
    L.own_hi = h->slab ? h->own_hi : h->p.lenX;
    L.row_shift = h->slab ? h->row_shift : 0;
    L.lenX_global = h->slab ? h->lenX_global : h->p.lenX;
    L.single_plan = h->p.ntemps == 1 && h->tables_set;
    memcpy(L.plan0, h->plan0, sizeof(L.plan0));
    return L;
}

struct DeviceGuard {
    int prev;
    bool ok;
    explicit DeviceGuard(int dev) : prev(-1), ok(false) {
        if (cudaGetDevice(&prev) != cudaSuccess) return;
        ok = cudaSetDevice(dev) == cudaSuccess;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

#define ENTER(h)                                                         \
    REQUIRE((h) != nullptr, PYMC_ERR_BADARG, "%s: null handle", __func__); \
    DeviceGuard _guard((h)->p.device);                                   \
    REQUIRE(_guard.ok, PYMC_ERR_CUDA, "%s: cannot select device %d", __func__, (h)->p.device)

// dispatch on (layout, model)
#define DISPATCH(h, CALL)                                                                            \
    do {                                                                                             \
        if ((h)->p.layout == PYMC_LAYOUT_INT8 && (h)->p.model == PYMC_MODEL_ISING) {                 \
            CALL(PYMC_LAYOUT_INT8, MODEL_ISING);                                                     \
        } else if ((h)->p.layout == PYMC_LAYOUT_INT8 && (h)->p.model == PYMC_MODEL_LATGAS) {         \
            CALL(PYMC_LAYOUT_INT8, MODEL_LATGAS);                                                    \
        } else if ((h)->p.layout == PYMC_LAYOUT_BITS && (h)->p.model == PYMC_MODEL_ISING) {          \
            CALL(PYMC_LAYOUT_BITS, MODEL_ISING);                                                     \
        } else {                                                                                     \
            CALL(PYMC_LAYOUT_BITS, MODEL_LATGAS);                                                    \
        }                                                                                            \
    } while (0)

// Call after the handle's stream has been synchronised.  A barrier inside a kernel that gave up waiting leaves
// the lattices of the handle in an undefined state: report it (once) instead of returning their contents.
static int check_async_errors(pymc_handle *h, const char *where) {
    if (!h->err_host) return PYMC_OK;
    if (h->err_host[0]) {
        h->err_host[0] = 0;
        return fail(PYMC_ERR_CUDA,
                    "%s: a block of the persistent sweep kernel timed out at a barrier; the configurations and "
                    "tracked sums of this handle are invalid (upload or initialise them again)", where);
    }
    if (h->err_host[1]) {
        h->err_host[1] = 0;
        return fail(PYMC_ERR_CUDA,
                    "%s: a neighbouring slab did not reach a barrier in time; the configurations and tracked sums of "
                    "this handle are invalid", where);
    }
    return PYMC_OK;
}

#define CHECK_ASYNC(h)                                   \
    do {                                                 \
        int _rc = check_async_errors((h), __func__);     \
        if (_rc) return _rc;                             \
    } while (0)

// K13: one launch per sweep, out of place (the two int8 buffers swap after every sweep).
template <int MODEL, int NWL2>
static int rowstream_launch(pymc_handle *h, const Lat &L, const dim3 &grid, bool set_attr) {
    if (set_attr)
        CUDA_TRY(cudaFuncSetAttribute(rowstream_kernel<MODEL, NWL2, Lat>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->rs_smem));
    rowstream_kernel<MODEL, NWL2, Lat><<<grid, RS_THREADS, h->rs_smem, h->stream>>>(L, h->rs, h->i8, h->i8_alt, h->sweep_counter,
                                                                                  h->persist_err);
    return PYMC_OK;
}

template <int MODEL>
static int rowstream_dispatch(pymc_handle *h, const Lat &L, const dim3 &grid, bool set_attr) {
    switch (h->rs.nw_log2) {
#define RS_CASE(K) \
    case K:        \
        if constexpr ((RS_THREADS >> K) >= 1 && (RS_THREADS >> K) <= RS_MAXQ) return rowstream_launch<MODEL, K>(h, L, grid, set_attr); \
        break;
        RS_CASE(2) RS_CASE(3) RS_CASE(4) RS_CASE(5) RS_CASE(6) RS_CASE(7) RS_CASE(8) RS_CASE(9)
#undef RS_CASE
    }
    return fail(PYMC_ERR_UNSUPPORTED, "row-streaming kernel: no instantiation for %d words per row", h->rs.nw);
}

static int launch_rowstream(pymc_handle *h, const Lat &L, int64_t nsweeps) {
    if (nsweeps <= 0) return PYMC_OK;
    const int m = h->p.model == PYMC_MODEL_ISING ? 0 : 1;
    if (!h->i8_alt) CUDA_TRY(cudaMalloc(&h->i8_alt, (size_t)h->p.nreplicas * 2 * h->p.lenX * h->nk));
    const dim3 grid(h->rs.pieces, h->p.nreplicas);
    for (int64_t s = 0; s < nsweeps; ++s) {
        const int rc = m ? rowstream_dispatch<MODEL_LATGAS>(h, L, grid, !h->rs_attr[m]) : rowstream_dispatch<MODEL_ISING>(h, L, grid, !h->rs_attr[m]);
        if (rc) return rc;
        h->rs_attr[m] = true;
        int8_t *tmp = h->i8;
        h->i8 = h->i8_alt;
        h->i8_alt = tmp;
        h->sweep_counter++;
    }
    CUDA_TRY(cudaGetLastError());
    return PYMC_OK;
}

extern "C" {

int pymc_abi_version(void) { return PYMC_ABI_VERSION; }

const char *pymc_last_error(void) { return g_err; }

int pymc_device_count(int *count) {
    REQUIRE(count, PYMC_ERR_BADARG, "pymc_device_count: null pointer");
    CUDA_TRY(cudaGetDeviceCount(count));
    return PYMC_OK;
}

int pymc_destroy(pymc_handle *h) {
    if (!h) return PYMC_OK;
    DeviceGuard guard(h->p.device);
    if (guard.ok) {
        if (h->own_stream) cudaStreamSynchronize(h->own_stream);
        cudaFree(h->i8);
        cudaFree(h->i8_alt);
        cudaFree(h->bits);
        cudaFree(h->plans);
        cudaFree(h->pair_plans);
        cudaFree(h->rep_to_T);
        cudaFree(h->tracked);
        cudaFree(h->reduced);
        cudaFree(h->persist_counter);
        cudaFree(h->persist_halo);
        cudaFree(h->pb_sync);
        cudaFree(h->pb_xbuf);
        for (int g = 0; g < h->rx_nranks; ++g)
            if (g != h->rx_rank) {
                if (h->peer_xpk[g]) cudaIpcCloseMemHandle(h->peer_xpk[g]);
                if (h->peer_xflags[g]) cudaIpcCloseMemHandle(h->peer_xflags[g]);
            }
        cudaFree(h->pb_xpk);
        cudaFree(h->pb_esum);
        cudaFree(h->pb_xflags);
        cudaFree(h->pb_maps);
        if (h->err_host) cudaFreeHost((void *)h->err_host);
        cudaFree(h->staging);
        cudaFree(h->qbuf);
        cudaFree(h->betas);
        cudaFree(h->energies);
        cudaFree(h->T_to_rep);
        cudaFree(h->rx_tried);
        cudaFree(h->rx_accepted);
        cudaFree(h->stats);
        cudaFree(h->stats_count);
        cudaFree(h->series);
        cudaFree(h->series_work);
        cudaFree(h->series_out);
        cudaFree(h->flags);
        cudaFree(h->slab_done);
        if (h->slab && h->nslabs > 1 && h->peer_up[0] && !h->local_peers) {
            cudaIpcCloseMemHandle(h->peer_up[0]);
            cudaIpcCloseMemHandle(h->peer_up[1]);
            cudaIpcCloseMemHandle(h->peer_up_flags);
            if (h->peer_xbuf_up) cudaIpcCloseMemHandle(h->peer_xbuf_up);
            if (h->peer_dn[0] != h->peer_up[0]) {
                cudaIpcCloseMemHandle(h->peer_dn[0]);
                cudaIpcCloseMemHandle(h->peer_dn[1]);
                cudaIpcCloseMemHandle(h->peer_dn_flags);
                if (h->peer_xbuf_dn) cudaIpcCloseMemHandle(h->peer_xbuf_dn);
            }
        }
        if (h->ev0) cudaEventDestroy(h->ev0);
        if (h->ev1) cudaEventDestroy(h->ev1);
        if (h->own_stream) cudaStreamDestroy(h->own_stream);
    }
    delete h;
    return PYMC_OK;
}

int pymc_create(const pymc_params *p, pymc_handle **out) {
    REQUIRE(p && out, PYMC_ERR_BADARG, "pymc_create: null pointer");
    *out = nullptr;
    REQUIRE(p->abi_version == PYMC_ABI_VERSION, PYMC_ERR_BADARG, "pymc_create: ABI version %d, library has %d",
            p->abi_version, PYMC_ABI_VERSION);
    REQUIRE(p->model == PYMC_MODEL_ISING || p->model == PYMC_MODEL_LATGAS, PYMC_ERR_BADARG, "pymc_create: model %d",
            p->model);
    REQUIRE(p->layout == PYMC_LAYOUT_INT8 || p->layout == PYMC_LAYOUT_BITS, PYMC_ERR_BADARG,
            "pymc_create: layout %d", p->layout);
    REQUIRE(p->lenX >= 2 && p->lenY >= 2, PYMC_ERR_BADARG, "pymc_create: lattice %dx%d too small", p->lenX, p->lenY);
    REQUIRE(p->lenX % 2 == 0 && p->lenY % 2 == 0, PYMC_ERR_UNSUPPORTED,
            "pymc_create: checkerboard sweeps need even lenX and lenY (got %dx%d); there is no CPU fallback",
            p->lenX, p->lenY);
    REQUIRE(p->layout != PYMC_LAYOUT_BITS || p->lenY % 128 == 0, PYMC_ERR_UNSUPPORTED,
            "pymc_create: the bit-packed layout needs lenY %% 128 == 0 (got %d)", p->lenY);
    REQUIRE(p->nreplicas >= 1 && p->nreplicas <= 65535, PYMC_ERR_BADARG, "pymc_create: nreplicas %d not in 1..65535",
            p->nreplicas);
    REQUIRE(p->nreplicas_global >= p->nreplicas && p->replica_offset >= 0 &&
                p->replica_offset + p->nreplicas <= p->nreplicas_global,
            PYMC_ERR_BADARG, "pymc_create: replica partition [%d,+%d) outside 0..%d", p->replica_offset, p->nreplicas,
            p->nreplicas_global);
    REQUIRE(p->ntemps >= 1, PYMC_ERR_BADARG, "pymc_create: ntemps %d", p->ntemps);
    int ndev = 0;
    CUDA_TRY(cudaGetDeviceCount(&ndev));
    REQUIRE(p->device >= 0 && p->device < ndev, PYMC_ERR_BADARG, "pymc_create: device %d of %d", p->device, ndev);

    // 32-bit indices: words of one colour of one replica (all sweep kernels) and 16-byte chunks of the natural image (the
    // upload / download kernels count lenX * lenY / 16 of them in an int): 2^34 sites per replica, 131072 x 131072
    REQUIRE((uint64_t)p->lenX * (uint64_t)((p->lenY / 2 + 63) / 64) <= (1ull << 27) && (uint64_t)p->lenX * p->lenY <= (1ull << 34),
            PYMC_ERR_UNSUPPORTED, "pymc_create: lattice too large (%d x %d; at most 2^34 sites per replica)", p->lenX, p->lenY);
    REQUIRE(p->kernel >= PYMC_KERNEL_AUTO && p->kernel <= PYMC_KERNEL_ROWSTREAM, PYMC_ERR_BADARG, "pymc_create: kernel %d",
            p->kernel);

    pymc_handle *h = new (std::nothrow) pymc_handle();
    REQUIRE(h, PYMC_ERR_BADARG, "pymc_create: out of host memory");
    memset(h, 0, sizeof(*h));
    h->p = *p;
    h->nk = p->lenY / 2;
    h->nw = (h->nk + 63) / 64;
    h->nclass = p->model == PYMC_MODEL_ISING ? PYMC_NCLASS_ISING : PYMC_NCLASS_LATGAS;

    DeviceGuard guard(p->device);
    int rc = PYMC_OK;
    if (guard.ok) {
        // K10 (pmc_resident.cuh): one block keeps one replica in shared memory for a whole call.  Chosen for batches
        // that give every SM a replica and whose bit image is small; PYMC_KERNEL_RESIDENT forces it wherever it fits.
        int sms = 0, smem = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, p->device);
        cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, p->device);
        const size_t image = (size_t)16 * p->lenX * h->nw;
        const bool fits = p->lenY % 128 == 0 && image + 2048 <= (size_t)smem;
        if (p->kernel == PYMC_KERNEL_RESIDENT && !fits) {
            delete h;
            return fail(PYMC_ERR_UNSUPPORTED,
                        "pymc_create: the resident kernel needs lenY %% 128 == 0 and a bit image (%zu bytes) within the "
                        "%d bytes of shared memory of a block", image, smem);
        }
        h->resident = p->kernel == PYMC_KERNEL_RESIDENT ||
                      (p->kernel == PYMC_KERNEL_AUTO && fits && image <= RESIDENT_AUTO_MAX_IMAGE && p->nreplicas >= sms);
        // Larger int8 images (config 3: 1024 x 1024 = 128 KiB, one block per SM): the fused kernel packs and unpacks
        // bytes every sweep, the resident one once per call -- it wins from 2 sweeps per call on (measured: 0.134 vs
        // 0.173 ms per sweep of 256 replicas at 10 sweeps per call), provided the replicas fill whole waves of blocks.
        if (!h->resident && p->kernel == PYMC_KERNEL_AUTO && fits && p->layout == PYMC_LAYOUT_INT8 && p->nreplicas >= sms) {
            int smem_sm = 0;
            cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, p->device);
            long per_sm = (long)(smem_sm / (image + 2048));
            if (per_sm > 2048 / RESIDENT_MAX_THREADS) per_sm = 2048 / RESIDENT_MAX_THREADS;
            if (per_sm >= 1) {
                const long slots = per_sm * sms, waves = (p->nreplicas + slots - 1) / slots;
                h->resident_big = (double)p->nreplicas / (double)(waves * slots) >= 0.8;
            }
        }
        // K11 (pmc_persist.cuh): one lattice over the shared memory of all SMs, one row slab per block
        {
            int coop = 0;
            cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, p->device);
            const int blocks = p->lenX < sms ? p->lenX : sms;
            const int max_rows = (p->lenX + blocks - 1) / blocks;
            const size_t slab = (size_t)16 * (max_rows + 2) * h->nw;
            // 512 bytes of static shared memory ride along; 16384 x 16384 on 148 SMs needs 231 424 of the 232 448 bytes
            const bool pfits = coop && p->lenY % 128 == 0 && slab + 1024 <= (size_t)smem;
            if (p->kernel == PYMC_KERNEL_PERSISTENT && !pfits) {
                delete h;
                return fail(PYMC_ERR_UNSUPPORTED,
                            "pymc_create: the persistent kernel needs lenY %% 128 == 0, cooperative launches and row slabs "
                            "(%zu bytes) within the %d bytes of shared memory of a block", slab, smem);
            }
            const char *off = getenv("PYMC_NO_PERSISTENT");  // escape hatch: keep AUTO on the per-sweep kernels
            h->persist = p->kernel == PYMC_KERNEL_PERSISTENT ||
                         (PMC_PERSIST_AUTO && !(off && off[0] == '1') && p->kernel == PYMC_KERNEL_AUTO && pfits && p->nreplicas == 1 &&
                          (size_t)p->lenX * p->lenY >= PERSIST_AUTO_MIN_SITES);
            h->persist_blocks = blocks;
            h->persist_smem = slab;
            int pthreads = (int)(((size_t)max_rows * h->nw + 31) / 32 * 32);
            h->persist_threads = pthreads < 128 ? 128 : (pthreads > PERSIST_MAX_THREADS ? PERSIST_MAX_THREADS : pthreads);
        }
        // K12 (pmc_pbatch.cuh): the rows of ALL replicas cut into one piece per SM
        {
            size_t rows_bytes = 0;
            const bool ok = plan_pbatch(h, p->nreplicas, p->lenX, sms, smem, &rows_bytes);
            if (p->kernel == PYMC_KERNEL_PBATCH && !ok) {
                delete h;
                return fail(PYMC_ERR_UNSUPPORTED,
                            "pymc_create: the batch-resident kernel needs lenY %% 128 == 0, cooperative launches, at most %d "
                            "replicas per block and row pieces (%zu bytes) within the %d bytes of shared memory of a block",
                            PB_MAXSEG, rows_bytes, smem);
            }
            const char *off = getenv("PYMC_NO_PERSISTENT");
            const bool allowed = !(off && off[0] == '1');
            // AUTO: from a few rows on, except batches that the one-block-per-replica kernel already serves with several
            // blocks per SM (small images, >= one replica per SM)
            const size_t sites = (size_t)p->nreplicas * p->lenX * p->lenY;
            const bool auto_ok = p->kernel == PYMC_KERNEL_AUTO && allowed && ok && !h->resident &&
                                 sites >= (p->nreplicas == 1 ? PBATCH_SINGLE_MIN_SITES : PBATCH_AUTO_MIN_SITES);
            h->pbatch = p->kernel == PYMC_KERNEL_PBATCH || auto_ok;
            if (h->pbatch && p->kernel == PYMC_KERNEL_AUTO) h->persist = h->resident_big = false;
        }
        // K13 (pmc_rowstream.cuh): int8 sweeps one at a time, rows streamed through shared memory
        {
            const int pieces = p->layout == PYMC_LAYOUT_INT8 ? rs_choose_pieces(p->lenX, h->nk, p->nreplicas, sms, smem) : 0;
            const bool rs_ok = pieces > 0 && rs_make_params(p->lenX, h->nk, pieces, h->rs);
            if (p->kernel == PYMC_KERNEL_ROWSTREAM && !rs_ok) {
                delete h;
                return fail(PYMC_ERR_UNSUPPORTED,
                            "pymc_create: the row-streaming kernel needs the int8 layout, lenY = 128 * 2^k with %d <= lenY/128 <= %d "
                            "and at least 2 * %d / (lenY/128) + 2 rows", RS_THREADS / RS_MAXQ, RS_THREADS, RS_THREADS);
            }
            const char *off = getenv("PYMC_NO_ROWSTREAM");  // escape hatch: keep AUTO on the tiled per-sweep kernel
            h->rowstream = rs_ok && (p->kernel == PYMC_KERNEL_ROWSTREAM ||
                                     (p->kernel == PYMC_KERNEL_AUTO && !(off && off[0] == '1') && !h->resident &&
                                      (size_t)p->nreplicas * p->lenX * p->lenY >= ROWSTREAM_AUTO_MIN_SITES && h->nw >= ROWSTREAM_AUTO_MIN_NW));
            h->rs_smem = rs_ok ? rs_smem_bytes(h->rs) : 0;
        }
        int threads = (int)(((size_t)p->lenX * h->nw / 2 + 31) / 32 * 32);
        h->resident_threads = threads < 128 ? 128 : (threads > RESIDENT_MAX_THREADS ? RESIDENT_MAX_THREADS : threads);
    }
    auto check = [&](cudaError_t e, const char *what) {
        if (e != cudaSuccess && rc == PYMC_OK)
            rc = fail(PYMC_ERR_CUDA, "pymc_create: %s failed: %s", what, cudaGetErrorString(e));
    };
    if (!guard.ok) rc = fail(PYMC_ERR_CUDA, "pymc_create: cannot select device %d", p->device);
    if (rc == PYMC_OK) {
        check(cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking), "cudaStreamCreate");
        h->stream = h->own_stream;
        check(cudaEventCreate(&h->ev0), "cudaEventCreate");
        check(cudaEventCreate(&h->ev1), "cudaEventCreate");
        {
            void *eh = nullptr, *ed = nullptr;
            check(cudaHostAlloc(&eh, 2 * sizeof(int), cudaHostAllocMapped), "cudaHostAlloc(error flags)");
            if (eh) {
                memset(eh, 0, 2 * sizeof(int));
                check(cudaHostGetDevicePointer(&ed, eh, 0), "cudaHostGetDevicePointer(error flags)");
            }
            h->err_host = (volatile int *)eh;
            h->persist_err = (int *)ed;
            h->slab_error = ed ? (int *)ed + 1 : nullptr;
        }
        const size_t nrep = p->nreplicas;
        if (p->layout == PYMC_LAYOUT_INT8) {
            check(cudaMalloc(&h->i8, nrep * 2 * p->lenX * (size_t)h->nk), "cudaMalloc(spins)");
        } else {
            check(cudaMalloc(&h->bits, nrep * 2 * p->lenX * (size_t)h->nw * 8), "cudaMalloc(bits)");
        }
        check(cudaMalloc(&h->plans, (size_t)p->ntemps * PLAN_WORDS * 4), "cudaMalloc(plans)");
        check(cudaMalloc(&h->rep_to_T, (size_t)p->nreplicas_global * 4), "cudaMalloc(rep_to_T)");
        check(cudaMalloc(&h->tracked, nrep * 16), "cudaMalloc(tracked)");
        check(cudaMalloc(&h->reduced, nrep * 16), "cudaMalloc(reduced)");
    }
    if (rc == PYMC_OK) {
        if (h->i8) check(cudaMemsetAsync(h->i8, 0, (size_t)p->nreplicas * 2 * p->lenX * h->nk, h->stream), "memset");
        if (h->bits)
            check(cudaMemsetAsync(h->bits, 0, (size_t)p->nreplicas * 2 * p->lenX * h->nw * 8, h->stream), "memset");
        check(cudaMemsetAsync(h->tracked, 0, (size_t)p->nreplicas * 16, h->stream), "memset");
        // default map: replica r -> row r (or row 0 when there is a single temperature)
        int32_t *map = new int32_t[p->nreplicas_global];
        for (int r = 0; r < p->nreplicas_global; ++r) map[r] = p->ntemps == p->nreplicas_global ? r : 0;
        check(cudaMemcpyAsync(h->rep_to_T, map, (size_t)p->nreplicas_global * 4, cudaMemcpyHostToDevice, h->stream),
              "memcpy(rep_to_T)");
        check(cudaStreamSynchronize(h->stream), "cudaStreamSynchronize");
        delete[] map;
    }
    if (rc != PYMC_OK) {
        char keep[sizeof(g_err)];
        memcpy(keep, g_err, sizeof(keep));
        pymc_destroy(h);
        memcpy(g_err, keep, sizeof(keep));
        return rc;
    }
    *out = h;
    return PYMC_OK;
}

int pymc_set_stream(pymc_handle *h, void *cuda_stream) {
    ENTER(h);
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    h->stream = (cudaStream_t)cuda_stream;  // NULL is CUDA's legacy default stream, a valid choice
    return PYMC_OK;
}

int pymc_reset_stream(pymc_handle *h) {
    ENTER(h);
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    h->stream = h->own_stream;
    return PYMC_OK;
}

int pymc_synchronize(pymc_handle *h) {
    ENTER(h);
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    CHECK_ASYNC(h);
    return PYMC_OK;
}

static int ensure_staging(pymc_handle *h) {
    if (!h->staging) CUDA_TRY(cudaMalloc(&h->staging, (size_t)h->p.lenX * h->p.lenY));
    return PYMC_OK;
}

static int upload_config(pymc_handle *h, int replica, const int8_t *host, size_t nsites, bool wait) {
    ENTER(h);
    REQUIRE(host, PYMC_ERR_BADARG, "pymc_upload_config: null host pointer");
    REQUIRE(replica >= 0 && replica < h->p.nreplicas, PYMC_ERR_BADARG, "pymc_upload_config: replica %d of %d", replica,
            h->p.nreplicas);
    const size_t want_up = (size_t)(h->slab ? h->own_hi - h->own_lo : h->p.lenX) * h->p.lenY;
    REQUIRE(nsites == want_up, PYMC_ERR_BADARG, "pymc_upload_config: %zu sites, lattice (or slab) has %zu", nsites,
            want_up);
    int rc = ensure_staging(h);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(h->staging, host, nsites, cudaMemcpyHostToDevice, h->stream));
    Lat L = make_lat(h);
    dim3 grid((h->p.lenX * h->nw + 127) / 128, 2);
    const bool fast = h->p.lenY % 16 == 0 && (h->p.layout == PYMC_LAYOUT_INT8 || h->p.lenY % 128 == 0);
    if (fast) {
        const int nchunks = (h->slab ? h->own_hi - h->own_lo : h->p.lenX) * (h->p.lenY / 16);
#define CALL_PACKF(LAY, MOD) pack_fast_kernel<LAY, MOD, Lat><<<(nchunks + 255) / 256, 256, 0, h->stream>>>(L, replica, h->staging)
        DISPATCH(h, CALL_PACKF);
#undef CALL_PACKF
    } else if (h->p.layout == PYMC_LAYOUT_BITS)
        pack_kernel<PYMC_LAYOUT_BITS><<<grid, 128, 0, h->stream>>>(L, replica, h->staging);
    else
        pack_kernel<PYMC_LAYOUT_INT8><<<grid, 128, 0, h->stream>>>(L, replica, h->staging);
    CUDA_TRY(cudaGetLastError());
    if (wait) CUDA_TRY(cudaStreamSynchronize(h->stream));  // the host buffer may be reused on return
    h->dirty = true;
    return PYMC_OK;
}

int pymc_upload_config(pymc_handle *h, int replica, const int8_t *host, size_t nsites) {
    return upload_config(h, replica, host, nsites, true);
}

int pymc_upload_config_async(pymc_handle *h, int replica, const int8_t *host, size_t nsites) {
    return upload_config(h, replica, host, nsites, false);
}

static int download_config(pymc_handle *h, int replica, int8_t *host, size_t nsites, bool wait) {
    ENTER(h);
    REQUIRE(host, PYMC_ERR_BADARG, "pymc_download_config: null host pointer");
    REQUIRE(replica >= 0 && replica < h->p.nreplicas, PYMC_ERR_BADARG, "pymc_download_config: replica %d of %d",
            replica, h->p.nreplicas);
    const size_t want_dn = (size_t)(h->slab ? h->own_hi - h->own_lo : h->p.lenX) * h->p.lenY;
    REQUIRE(nsites == want_dn, PYMC_ERR_BADARG, "pymc_download_config: %zu sites, lattice (or slab) has %zu", nsites,
            want_dn);
    int rc = ensure_staging(h);
    if (rc) return rc;
    Lat L = make_lat(h);
    dim3 grid((h->p.lenX * h->nw + 127) / 128, 2);
    const bool fast = h->p.lenY % 16 == 0 && (h->p.layout == PYMC_LAYOUT_INT8 || h->p.lenY % 128 == 0);
    const int nchunks = (h->slab ? h->own_hi - h->own_lo : h->p.lenX) * (h->p.lenY / 16);
#define CALL_UNPACK(LAY, MOD)                                                                                     \
    if (fast) unpack_fast_kernel<LAY, MOD, Lat><<<(nchunks + 255) / 256, 256, 0, h->stream>>>(L, replica, h->staging); \
    else unpack_kernel<LAY, MOD><<<grid, 128, 0, h->stream>>>(L, replica, h->staging)
    DISPATCH(h, CALL_UNPACK);
#undef CALL_UNPACK
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(host, h->staging, nsites, cudaMemcpyDeviceToHost, h->stream));
    if (!wait) return PYMC_OK;  // pymc_synchronize completes the copy and reports a failed barrier
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    CHECK_ASYNC(h);
    return PYMC_OK;
}

int pymc_download_config(pymc_handle *h, int replica, int8_t *host, size_t nsites) {
    return download_config(h, replica, host, nsites, true);
}

int pymc_download_config_async(pymc_handle *h, int replica, int8_t *host, size_t nsites) {
    return download_config(h, replica, host, nsites, false);
}

static int rebase_tracked(pymc_handle *h);

int pymc_init_random(pymc_handle *h) {
    ENTER(h);
    Lat L = make_lat(h);
    if (h->slab) {
        const size_t n = (size_t)h->p.lenX * (h->p.lenY / 4);
        dim3 sgrid((unsigned)((n + 255) / 256), h->p.nreplicas);
        if (h->p.model == PYMC_MODEL_ISING)
            init_slab_kernel<MODEL_ISING><<<sgrid, 256, 0, h->stream>>>(L);
        else
            init_slab_kernel<MODEL_LATGAS><<<sgrid, 256, 0, h->stream>>>(L);
        CUDA_TRY(cudaGetLastError());
        return rebase_tracked(h);
    }
    if (h->bits) CUDA_TRY(cudaMemsetAsync(h->bits, 0, (size_t)h->p.nreplicas * 2 * h->p.lenX * h->nw * 8, h->stream));
    const size_t ncalls = ((size_t)h->p.lenX * h->p.lenY + 3) / 4;
    dim3 grid((unsigned)((ncalls + 255) / 256), h->p.nreplicas);
#define CALL_INIT(LAY, MOD) init_kernel<LAY, MOD><<<grid, 256, 0, h->stream>>>(L)
    DISPATCH(h, CALL_INIT);
#undef CALL_INIT
    CUDA_TRY(cudaGetLastError());
    return rebase_tracked(h);
}

int pymc_set_accept_tables(pymc_handle *h, const uint32_t *thr, int ntemps, int nclass) {
    ENTER(h);
    REQUIRE(thr, PYMC_ERR_BADARG, "pymc_set_accept_tables: null pointer");
    REQUIRE(ntemps == h->p.ntemps && nclass == h->nclass, PYMC_ERR_BADARG,
            "pymc_set_accept_tables: got %d x %d, handle expects %d x %d", ntemps, nclass, h->p.ntemps, h->nclass);
    // dE > 0 ("hard", thr < 2^32-1) comes from dE = -2J(2a-4) (Ising) or dE = -(2c-1)(J nn + Eads - mu),
    // so a consistent table has them only at one end of the Ising class range (a in {3,4} or {0,1})
    // and in at most one occupancy per neighbour count of the lattice gas.  The kernels rely on it.
    for (int t = 0; t < ntemps; ++t) {
        const uint32_t *row = thr + (size_t)t * nclass;
        auto hard = [&](int c) { return row[c] != 0xFFFFFFFFu; };
        if (h->p.model == PYMC_MODEL_ISING) {
            const bool lo = hard(0) || hard(1), hi = hard(3) || hard(4);
            REQUIRE(!hard(2) && !(lo && hi), PYMC_ERR_UNSUPPORTED,
                    "pymc_set_accept_tables: row %d is not an Ising table (dE > 0 at both ends or at dE = 0)", t);
        } else {
            for (int n = 0; n < 5; ++n)
                REQUIRE(!(hard(n) && hard(5 + n)), PYMC_ERR_UNSUPPORTED,
                        "pymc_set_accept_tables: row %d has dE > 0 for both occupancies at %d neighbours", t, n);
        }
    }
    // expand each threshold row into the plan row the kernels read (orientation, enables, bit streams)
    uint32_t *rows = new (std::nothrow) uint32_t[(size_t)ntemps * PLAN_WORDS];
    REQUIRE(rows, PYMC_ERR_BADARG, "pymc_set_accept_tables: out of host memory");
    for (int t = 0; t < ntemps; ++t) {
        if (h->p.model == PYMC_MODEL_ISING)
            Plan<MODEL_ISING>::build_row(thr + (size_t)t * nclass, rows + (size_t)t * PLAN_WORDS);
        else
            Plan<MODEL_LATGAS>::build_row(thr + (size_t)t * nclass, rows + (size_t)t * PLAN_WORDS);
    }
    memcpy(h->plan0, rows, sizeof(h->plan0));
    cudaError_t e = cudaMemcpyAsync(h->plans, rows, (size_t)ntemps * PLAN_WORDS * 4, cudaMemcpyHostToDevice, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    delete[] rows;
    if (e != cudaSuccess) return fail(PYMC_ERR_CUDA, "pymc_set_accept_tables: upload failed: %s", cudaGetErrorString(e));
    h->tables_set = true;
    return PYMC_OK;
}

int pymc_set_replica_to_T(pymc_handle *h, const int32_t *map, int n) {
    ENTER(h);
    REQUIRE(map && n == h->p.nreplicas_global, PYMC_ERR_BADARG, "pymc_set_replica_to_T: need %d entries, got %d",
            h->p.nreplicas_global, n);
    for (int r = 0; r < n; ++r)
        REQUIRE(map[r] >= 0 && map[r] < h->p.ntemps, PYMC_ERR_BADARG, "pymc_set_replica_to_T: map[%d] = %d not in 0..%d",
                r, map[r], h->p.ntemps - 1);
    if (h->rx_ready) {
        bool *seen = new bool[h->p.ntemps]();
        bool perm = n == h->p.ntemps;
        for (int r = 0; r < n && perm; ++r) {
            perm = !seen[map[r]];
            seen[map[r]] = true;
        }
        delete[] seen;
        REQUIRE(perm, PYMC_ERR_BADARG, "pymc_set_replica_to_T: replica exchange needs a permutation of 0..%d", n - 1);
    }
    CUDA_TRY(cudaMemcpyAsync(h->rep_to_T, map, (size_t)n * 4, cudaMemcpyHostToDevice, h->stream));
    if (h->rx_ready) {
        rx_invert_map_kernel<<<(n + 255) / 256, 256, 0, h->stream>>>(h->rep_to_T, h->T_to_rep, n);
        CUDA_TRY(cudaGetLastError());
    }
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    return PYMC_OK;
}

int pymc_get_replica_to_T(pymc_handle *h, int32_t *map, int n) {
    ENTER(h);
    REQUIRE(map && n == h->p.nreplicas_global, PYMC_ERR_BADARG, "pymc_get_replica_to_T: need %d entries, got %d",
            h->p.nreplicas_global, n);
    CUDA_TRY(cudaMemcpyAsync(map, h->rep_to_T, (size_t)n * 4, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    CHECK_ASYNC(h);
    return PYMC_OK;
}

// K11: nsweeps sweeps of every replica of the handle, one cooperative launch per replica and <= 4096 sweeps.
// Returns PYMC_ERR_UNSUPPORTED (and switches the handle back to the per-sweep kernels) if the device refuses the
// cooperative launch, e.g. because the SMs are shared with another context.
static int launch_persistent(pymc_handle *h, int64_t nsweeps) {
    REQUIRE(h->tables_set, PYMC_ERR_STATE, "pymc_sweep: call pymc_set_accept_tables first");
    if (h->dirty) {
        int rc = rebase_tracked(h);
        if (rc) return rc;
    }
    const int m = h->p.model == PYMC_MODEL_ISING ? 0 : 1;
    const void *fn;
    if (h->p.layout == PYMC_LAYOUT_INT8)
        fn = m ? (const void *)persistent_kernel<MODEL_LATGAS, PYMC_LAYOUT_INT8, Lat> : (const void *)persistent_kernel<MODEL_ISING, PYMC_LAYOUT_INT8, Lat>;
    else
        fn = m ? (const void *)persistent_kernel<MODEL_LATGAS, PYMC_LAYOUT_BITS, Lat> : (const void *)persistent_kernel<MODEL_ISING, PYMC_LAYOUT_BITS, Lat>;
    if (!h->persist_attr[m]) {
        if (h->persist_smem > 48 * 1024)
            CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->persist_smem));
        h->persist_attr[m] = true;
    }
    if (!h->persist_counter) {
        CUDA_TRY(cudaMalloc(&h->persist_counter, sizeof(unsigned int)));
        CUDA_TRY(cudaMalloc(&h->persist_halo, (size_t)2 * h->persist_blocks * 2 * h->nw * 8));
    }
    Lat L = make_lat(h);
    while (nsweeps > 0) {
        const int64_t chunk = nsweeps > 4096 ? 4096 : nsweeps;
        for (int rep = 0; rep < h->p.nreplicas; ++rep) {
            PersistArgs a;
            a.nsweeps = chunk;
            a.sweep0 = h->sweep_counter;
            a.rep = rep;
            a.counter = h->persist_counter;
            a.err = h->persist_err;
            a.halo = h->persist_halo;
            CUDA_TRY(cudaMemsetAsync(h->persist_counter, 0, sizeof(unsigned int), h->stream));
            void *args[] = {(void *)&L, (void *)&a};
            cudaError_t e = cudaLaunchCooperativeKernel(fn, dim3(h->persist_blocks), dim3(h->persist_threads), args,
                                                        h->persist_smem, h->stream);
            if (e != cudaSuccess) {
                cudaGetLastError();
                h->persist = false;  // nothing has run for this replica yet: the caller falls back
                return fail(PYMC_ERR_UNSUPPORTED, "cooperative launch refused: %s", cudaGetErrorString(e));
            }
        }
        h->sweep_counter += chunk;
        nsweeps -= chunk;
    }
    return PYMC_OK;
}

// Buffers of the in-kernel exchange; also what pymc_rx_ipc_export hands to the other GPUs.
static int ensure_rx_exchange_buffers(pymc_handle *h) {
    if (!h->pb_esum) {
        CUDA_TRY(cudaMalloc(&h->pb_esum, (size_t)2 * h->p.ntemps * 2 * sizeof(long long)));
        CUDA_TRY(cudaMemset(h->pb_esum, 0, (size_t)2 * h->p.ntemps * 2 * sizeof(long long)));
    }
    if (!h->pb_xflags) {
        CUDA_TRY(cudaMalloc(&h->pb_xflags, PB_MAXRANKS * sizeof(unsigned int)));
        CUDA_TRY(cudaMemset(h->pb_xflags, 0, PB_MAXRANKS * sizeof(unsigned int)));
    }
    if (!h->pb_xpk) {
        const size_t bytes = (size_t)2 * h->p.ntemps * 2 * sizeof(PBPacket);
        CUDA_TRY(cudaMalloc(&h->pb_xpk, bytes));
        CUDA_TRY(cudaMemset(h->pb_xpk, 0, bytes));
    }
    return PYMC_OK;
}

// Pass numbers of K12's packets: every launch starts PB_EPOCH_STEP above the one before (a launch has at most 2 * 4096 + 2
// passes), so the exchange buffer is zeroed once and never again -- which also means that a neighbouring GPU may store its
// first packets here before this GPU has even launched.
static const unsigned int PB_EPOCH_STEP = 16384;

static int slab_refresh_halos(pymc_handle *h);

static int ensure_pbatch_buffers(pymc_handle *h) {
    if (!h->pb_sync) CUDA_TRY(cudaMalloc(&h->pb_sync, 2 * sizeof(unsigned int)));
    if (!h->pb_xbuf) {
        const size_t bytes = (size_t)2 * pb_nslots(h->pb_blocks, h->p.nreplicas) * h->nw * sizeof(PBPacket);
        CUDA_TRY(cudaMalloc(&h->pb_xbuf, bytes));
        CUDA_TRY(cudaMemset(h->pb_xbuf, 0, bytes));
    }
    return PYMC_OK;
}

// K12: nsweeps sweeps of every replica in one cooperative launch per <= 4096 sweeps; with xchg_every > 0 (single-GPU
// replica exchange) an exchange after every xchg_every-th sweep, decided inside the kernel.  Returns
// PYMC_ERR_UNSUPPORTED (and switches the handle back to the other kernels) if the device refuses the launch.
static int launch_pbatch(pymc_handle *h, int64_t nsweeps, int xchg_every, int *scheme_io) {
    REQUIRE(h->tables_set, PYMC_ERR_STATE, "pymc_sweep: call pymc_set_accept_tables first");
    if (nsweeps <= 0) return PYMC_OK;
    const int m = h->p.model == PYMC_MODEL_ISING ? 0 : 1;
    const void *fn;
    const bool keep = h->pb_cache_off > 0;  // room in shared memory for the per-thread item data: the KEEP instantiation
#define PB_FN(MOD, LAY, SLB) \
    (keep ? (const void *)pbatch_kernel<MOD, LAY, true, SLB, Lat> : (const void *)pbatch_kernel<MOD, LAY, false, SLB, Lat>)
    if (h->slab)  // slabs are int8
        fn = m ? PB_FN(MODEL_LATGAS, PYMC_LAYOUT_INT8, true) : PB_FN(MODEL_ISING, PYMC_LAYOUT_INT8, true);
    else if (h->p.layout == PYMC_LAYOUT_INT8)
        fn = m ? PB_FN(MODEL_LATGAS, PYMC_LAYOUT_INT8, false) : PB_FN(MODEL_ISING, PYMC_LAYOUT_INT8, false);
    else
        fn = m ? PB_FN(MODEL_LATGAS, PYMC_LAYOUT_BITS, false) : PB_FN(MODEL_ISING, PYMC_LAYOUT_BITS, false);
#undef PB_FN
    if (!h->pb_attr[m]) {
        if (h->pb_smem > 48 * 1024)
            CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->pb_smem));
        h->pb_attr[m] = true;
    }
    const int B = h->pb_blocks, R = h->p.ntemps;
    {
        int rc = ensure_pbatch_buffers(h);
        if (rc) return rc;
    }
    if (xchg_every > 0 && !h->pb_maps) CUDA_TRY(cudaMalloc(&h->pb_maps, (size_t)4 * R * sizeof(int32_t)));
    if (xchg_every > 0) {
        int rc = ensure_rx_exchange_buffers(h);
        if (rc) return rc;
    }
    Lat L = make_lat(h);
    int scheme = scheme_io ? *scheme_io : 0;
    while (nsweeps > 0) {
        int64_t chunk = nsweeps > 4096 ? 4096 : nsweeps;
        if (xchg_every > 0 && chunk < nsweeps) chunk -= chunk % xchg_every;  // launches end on exchange boundaries
        if (chunk <= 0) chunk = nsweeps > 4096 ? xchg_every : nsweeps;
        PBatchArgs a;
        memset(&a, 0, sizeof(a));
        a.nsweeps = chunk;
        a.sweep0 = h->sweep_counter;
        a.xchg_every = xchg_every;
        a.xchg_phase = 0;
        a.scheme0 = scheme;
        a.round0 = h->rx_round;
        a.bar = h->pb_sync;
        a.err = h->persist_err;
        a.xbuf = h->pb_xbuf;
        a.xbuf_up = h->slab ? h->peer_xbuf_up : h->pb_xbuf;
        a.xbuf_dn = h->slab ? h->peer_xbuf_dn : h->pb_xbuf;
        h->pb_epoch += PB_EPOCH_STEP;
        a.epoch = h->pb_epoch;
        a.row0 = h->slab ? h->slab_index * (h->own_hi - h->own_lo) : 0;
        a.cache_off = h->pb_cache_off;
        a.esum = h->pb_esum;
        a.nranks = h->rx_nranks > 1 ? h->rx_nranks : 1;
        a.rank = h->rx_nranks > 1 ? h->rx_rank : 0;
        a.xpk = h->pb_xpk;
        for (int g = 0; g < a.nranks; ++g) a.peer_xpk[g] = h->peer_xpk[g];
        a.out = h->reduced;
        a.ntemps = R;
        a.rep_to_T[0] = h->rep_to_T;
        const int64_t nex = xchg_every > 0 ? chunk / xchg_every : 0;
        if (xchg_every > 0) {
            // parity 0 = the handle's own maps (current at entry), parity 1 = the spare pair
            a.rep_to_T[1] = h->pb_maps;
            a.T_to_rep[0] = h->T_to_rep;
            a.T_to_rep[1] = h->pb_maps + R;
            a.betas = h->betas;
            a.c_bond = h->c_bond;
            a.c_second = h->c_second;
            a.tried = h->rx_tried;
            a.accepted = h->rx_accepted;
            a.energies = h->energies;
        }
        CUDA_TRY(cudaMemsetAsync(h->pb_sync, 0, 2 * sizeof(unsigned int), h->stream));
        CUDA_TRY(cudaMemsetAsync(h->reduced, 0, (size_t)h->p.nreplicas * 16, h->stream));
        if (xchg_every > 0)  // the blocks add deltas to the sums of MY replicas (other GPUs store theirs whole)
            for (int par = 0; par < 2; ++par)
                CUDA_TRY(cudaMemsetAsync(h->pb_esum + ((size_t)par * R + h->p.replica_offset) * 2, 0,
                                         (size_t)h->p.nreplicas * 16, h->stream));
        void *args[] = {(void *)&L, (void *)&a};
        cudaError_t e = cudaLaunchCooperativeKernel(fn, dim3(B), dim3(h->pb_threads), args, h->pb_smem, h->stream);
        if (e != cudaSuccess) {
            cudaGetLastError();
            h->pbatch = false;  // nothing has run in this chunk: the caller falls back
            return fail(PYMC_ERR_UNSUPPORTED, "cooperative launch refused: %s", cudaGetErrorString(e));
        }
        if (h->slab) {
            // The kernel keeps only the owned rows: bring the halo rows of the per-sweep kernel up to date (mine are written by
            // the neighbours, as after an upload) and let the tracked sums be re-based in the convention of reduce_kernel (a
            // slab's share of the boundary bonds differs between the two reductions; the sum over the slabs does not).
            const int rc = slab_refresh_halos(h);
            if (rc) return rc;
        } else {
            // the exact sums at the end of the launch are the new tracked values
            CUDA_TRY(cudaMemcpyAsync(h->tracked, h->reduced, (size_t)h->p.nreplicas * 16, cudaMemcpyDeviceToDevice, h->stream));
            h->dirty = false;
        }
        if (nex & 1) {  // an odd number of exchanges leaves the current maps in the spare pair
            CUDA_TRY(cudaMemcpyAsync(h->rep_to_T, h->pb_maps, (size_t)R * 4, cudaMemcpyDeviceToDevice, h->stream));
            CUDA_TRY(cudaMemcpyAsync(h->T_to_rep, h->pb_maps + R, (size_t)R * 4, cudaMemcpyDeviceToDevice, h->stream));
        }
        h->sweep_counter += chunk;
        h->rx_round += (uint64_t)nex;
        if (nex & 1) scheme ^= 1;
        nsweeps -= chunk;
    }
    if (scheme_io) *scheme_io = scheme;
    return PYMC_OK;
}

// K10: nsteps steps (flip sweep and / or pair pass each) of every replica in one launch per <= 4096 steps.
static int launch_resident(pymc_handle *h, int64_t nsteps, bool flip, uint32_t mask_even, uint32_t mask_odd) {
    REQUIRE(!flip || h->tables_set, PYMC_ERR_STATE, "sweep: call pymc_set_accept_tables first");
    REQUIRE(!(mask_even | mask_odd) || h->pair_plans, PYMC_ERR_STATE, "pair sweep: call pymc_set_pair_tables first");
    if (h->dirty) {
        int rc = rebase_tracked(h);
        if (rc) return rc;
    }
    const size_t image = (size_t)16 * h->p.lenX * h->nw;
    const int m = h->p.model == PYMC_MODEL_ISING ? 0 : 1;
#define RESIDENT_INST(MOD, LAY) resident_kernel<MOD, LAY, Lat>
    if (!h->resident_attr[m] && image > 48 * 1024) {
        cudaError_t e;
        if (h->p.layout == PYMC_LAYOUT_INT8)
            e = m ? cudaFuncSetAttribute(RESIDENT_INST(MODEL_LATGAS, PYMC_LAYOUT_INT8), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)image)
                  : cudaFuncSetAttribute(RESIDENT_INST(MODEL_ISING, PYMC_LAYOUT_INT8), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)image);
        else
            e = m ? cudaFuncSetAttribute(RESIDENT_INST(MODEL_LATGAS, PYMC_LAYOUT_BITS), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)image)
                  : cudaFuncSetAttribute(RESIDENT_INST(MODEL_ISING, PYMC_LAYOUT_BITS), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)image);
        CUDA_TRY(e);
    }
    h->resident_attr[m] = true;
    Lat L = make_lat(h);
    while (nsteps > 0) {
        ResidentArgs a;
        a.nsteps = nsteps > 4096 ? 4096 : nsteps;
        a.flip = flip ? 1 : 0;
        a.mask_even = mask_even;
        a.mask_odd = mask_odd;
        a.sweep0 = h->sweep_counter;
        a.pair0 = h->pair_counter;
        const int T = h->resident_threads, G = h->p.nreplicas;
        if (h->p.layout == PYMC_LAYOUT_INT8) {
            if (m) RESIDENT_INST(MODEL_LATGAS, PYMC_LAYOUT_INT8)<<<G, T, image, h->stream>>>(L, a);
            else RESIDENT_INST(MODEL_ISING, PYMC_LAYOUT_INT8)<<<G, T, image, h->stream>>>(L, a);
        } else {
            if (m) RESIDENT_INST(MODEL_LATGAS, PYMC_LAYOUT_BITS)<<<G, T, image, h->stream>>>(L, a);
            else RESIDENT_INST(MODEL_ISING, PYMC_LAYOUT_BITS)<<<G, T, image, h->stream>>>(L, a);
        }
        // the counters move exactly as in the kernel
        for (int64_t s = 0; s < a.nsteps; ++s) {
            if (flip) h->sweep_counter++;
            if ((h->pair_counter & 1ull) ? mask_odd : mask_even) h->pair_counter++;
            else if (!flip) break;  // nothing can change any more
        }
        nsteps -= a.nsteps;
    }
#undef RESIDENT_INST
    CUDA_TRY(cudaGetLastError());
    return PYMC_OK;
}

static bool uses_fused(const pymc_handle *h) {
    const int nown = h->slab ? h->own_hi - h->own_lo : h->p.lenX;
    return !h->resident && h->p.layout == PYMC_LAYOUT_INT8 && h->p.kernel == PYMC_KERNEL_AUTO && h->nk % 64 == 0 &&
           h->nw % FUSED_TW == 0 && nown % FUSED_TH == 0;
}

static int slab_sync_async(pymc_handle *h);

// the kernel a call of pymc_sweep(h, nsweeps) runs on (the order of the tests is that of pymc_sweep)
static const char *sweep_kernel_for(const pymc_handle *h, int64_t nsweeps, int *launches_per_sweep) {
    int n = 2;
    const char *name;
    if (h->pbatch && nsweeps >= 2 && (!h->slab || h->slab_pbatch)) {
        name = "pbatch_kernel";  // one cooperative launch per call of >= 2 sweeps (and per exchange cycle on one GPU)
        n = 1;
    } else if (h->resident || (h->resident_big && nsweeps >= 2)) {
        name = "resident_kernel";  // one launch per call (up to 4096 sweeps), not per sweep
        n = 1;
    } else if (h->persist && nsweeps >= 2) {
        name = "persistent_kernel";  // one cooperative launch per call of >= 2 sweeps
        n = 1;
    } else if (h->rowstream && !h->slab) {
        name = "rowstream_kernel";
        n = 1;
    } else if (uses_fused(h)) {
        name = "sweep_fused_i8_kernel";
        n = 1;
    } else if (h->p.layout == PYMC_LAYOUT_BITS && h->p.kernel != PYMC_KERNEL_GENERIC) {
        name = "sweep_bits_rows_kernel";
    } else if (h->p.layout == PYMC_LAYOUT_INT8 && h->p.kernel != PYMC_KERNEL_GENERIC && h->nk % 64 == 0) {
        name = "sweep_tiled_i8_kernel";
    } else {
        name = "sweep_word_kernel";
    }
    if (launches_per_sweep) *launches_per_sweep = n;
    return name;
}

int pymc_sweep_kernel_info(pymc_handle *h, int *launches_per_sweep, const char **kernel_name) {
    REQUIRE(h, PYMC_ERR_BADARG, "pymc_sweep_kernel_info: null handle");
    const char *name = sweep_kernel_for(h, 1 << 20, launches_per_sweep);  // a call of many sweeps
    if (kernel_name) *kernel_name = name;
    return PYMC_OK;
}

int pymc_sweep_kernel_for(pymc_handle *h, int64_t nsweeps, int *launches_per_sweep, const char **kernel_name) {
    REQUIRE(h, PYMC_ERR_BADARG, "pymc_sweep_kernel_for: null handle");
    const char *name = sweep_kernel_for(h, nsweeps, launches_per_sweep);
    if (kernel_name) *kernel_name = name;
    return PYMC_OK;
}

int pymc_sweep(pymc_handle *h, int64_t nsweeps) {
    ENTER(h);
    REQUIRE(nsweeps >= 0, PYMC_ERR_BADARG, "pymc_sweep: nsweeps %lld", (long long)nsweeps);
    REQUIRE(h->tables_set, PYMC_ERR_STATE, "pymc_sweep: call pymc_set_accept_tables first");
    REQUIRE(!h->slab || h->peer_up[0], PYMC_ERR_STATE, "pymc_sweep: connect the slab to its neighbours first");
    if (h->pbatch && nsweeps >= 2 && (!h->slab || h->slab_pbatch)) {
        const int rc = launch_pbatch(h, nsweeps, 0, nullptr);
        if (rc != PYMC_ERR_UNSUPPORTED) return rc;
    }
    if (h->resident || (h->resident_big && nsweeps >= 2)) return launch_resident(h, nsweeps, true, 0, 0);
    if (h->persist && nsweeps >= 2) {
        // Several replicas: one launch each, in turn.  If the first launch is refused nothing has run and the
        // per-sweep kernels take over; a refusal later cannot happen (same launch shape).
        const int rc = launch_persistent(h, nsweeps);
        if (rc != PYMC_ERR_UNSUPPORTED) return rc;
    }
    if (h->dirty) {
        int rc = rebase_tracked(h);
        if (rc) return rc;
    }
    Lat L = make_lat(h);
    if (h->rowstream && !h->slab) return launch_rowstream(h, L, nsweeps);
    dim3 grid((h->p.lenX * h->nw + 127) / 128, h->p.nreplicas);
    // K1 tiled int8 kernel: needs whole words (lenY % 128 == 0); tile = TH rows x TW words, TW | nw
    const bool tiled = h->p.layout == PYMC_LAYOUT_INT8 && h->p.kernel != PYMC_KERNEL_GENERIC && h->nk % 64 == 0;
    // K1f fused two-colour int8 kernel: whole 32 x 8-word tiles, out of place (ping-pong buffers)
    const bool fused = uses_fused(h);
    if (fused && !h->i8_alt && nsweeps > 0)
        CUDA_TRY(cudaMalloc(&h->i8_alt, (size_t)h->p.nreplicas * 2 * h->p.lenX * h->nk));
    if (fused) {
        dim3 fgrid(h->nw / FUSED_TW, (L.own_hi - L.own_lo) / FUSED_TH, h->p.nreplicas);
        for (int64_t s = 0; s < nsweeps; ++s) {
            int8_t *pu = nullptr, *pd = nullptr;
            if (h->slab) {  // the neighbours' destination buffer has the same index as mine (one slab: myself)
                const int k = h->i8_alt == h->buf[0] ? 0 : 1;
                pu = h->peer_up[k];
                pd = h->peer_dn[k];
            }
            SlabSync sync;
            memset(&sync, 0, sizeof(sync));
            if (h->slab && h->nslabs > 1 && !h->local_peers) {  // neighbour barrier carried by the sweep kernel
                sync.mine = h->flags;
                sync.up = h->peer_up_flags;
                sync.dn = h->peer_dn_flags;
                sync.wait_epoch = h->slab_epoch;
                sync.signal_epoch = ++h->slab_epoch;
                sync.done = h->slab_done;
                sync.err = h->slab_error;
            }
#define LAUNCH_FUSED(MOD, SLB) \
    sweep_fused_i8_kernel<MOD, SLB, Lat><<<fgrid, FUSED_THREADS, 0, h->stream>>>(L, h->i8, h->i8_alt, h->sweep_counter, pu, pd, sync)
            if (h->p.model == PYMC_MODEL_ISING) {
                if (h->slab) LAUNCH_FUSED(MODEL_ISING, true);
                else LAUNCH_FUSED(MODEL_ISING, false);
            } else {
                if (h->slab) LAUNCH_FUSED(MODEL_LATGAS, true);
                else LAUNCH_FUSED(MODEL_LATGAS, false);
            }
#undef LAUNCH_FUSED
            int8_t *tmp = h->i8;
            h->i8 = h->i8_alt;
            h->i8_alt = tmp;
            h->sweep_counter++;
            if (h->slab) h->dirty_halo = true;  // my halo rows are complete only once the neighbours' step is
        }
        CUDA_TRY(cudaGetLastError());
        return PYMC_OK;
    }
    REQUIRE(!h->slab, PYMC_ERR_UNSUPPORTED, "pymc_sweep: slab handles use the fused int8 kernel only");
    const bool bitrows = h->p.layout == PYMC_LAYOUT_BITS && h->p.kernel != PYMC_KERNEL_GENERIC;
    const int rpt = h->p.lenX % 4 == 0 ? 4 : 2;
    int tw_log2 = 0;
    while (tw_log2 < 4 && h->nw % (2 << tw_log2) == 0) ++tw_log2;
    const int TH = TILE_THREADS >> tw_log2;
    dim3 tgrid(h->nw >> tw_log2, (h->p.lenX + TH - 1) / TH, h->p.nreplicas);
    for (int64_t s = 0; s < nsweeps; ++s) {
        for (int colour = 0; colour < 2; ++colour) {
            const uint64_t step = 2 * h->sweep_counter + colour;
            if (bitrows) {
                dim3 bgrid(((h->p.lenX / rpt) * h->nw + 127) / 128, h->p.nreplicas);
                if (h->p.model == PYMC_MODEL_ISING) {
                    if (rpt == 4) sweep_bits_rows_kernel<MODEL_ISING, 4><<<bgrid, 128, 0, h->stream>>>(L, colour, step);
                    else sweep_bits_rows_kernel<MODEL_ISING, 2><<<bgrid, 128, 0, h->stream>>>(L, colour, step);
                } else {
                    if (rpt == 4) sweep_bits_rows_kernel<MODEL_LATGAS, 4><<<bgrid, 128, 0, h->stream>>>(L, colour, step);
                    else sweep_bits_rows_kernel<MODEL_LATGAS, 2><<<bgrid, 128, 0, h->stream>>>(L, colour, step);
                }
                continue;
            }
            if (tiled) {
                if (h->p.model == PYMC_MODEL_ISING)
                    sweep_tiled_i8_kernel<MODEL_ISING, Lat><<<tgrid, TILE_THREADS, 0, h->stream>>>(L, colour, step, tw_log2);
                else
                    sweep_tiled_i8_kernel<MODEL_LATGAS, Lat><<<tgrid, TILE_THREADS, 0, h->stream>>>(L, colour, step, tw_log2);
                continue;
            }
#define CALL_SWEEP(LAY, MOD) sweep_word_kernel<LAY, MOD><<<grid, 128, 0, h->stream>>>(L, colour, step)
            DISPATCH(h, CALL_SWEEP);
#undef CALL_SWEEP
        }
        h->sweep_counter++;
    }
    CUDA_TRY(cudaGetLastError());
    return PYMC_OK;
}

int pymc_get_sweep_counter(pymc_handle *h, uint64_t *counter) {
    REQUIRE(h && counter, PYMC_ERR_BADARG, "pymc_get_sweep_counter: null pointer");
    *counter = h->sweep_counter;
    return PYMC_OK;
}

int pymc_set_sweep_counter(pymc_handle *h, uint64_t counter) {
    REQUIRE(h, PYMC_ERR_BADARG, "pymc_set_sweep_counter: null handle");
    REQUIRE(counter < (1ull << 55), PYMC_ERR_BADARG, "pymc_set_sweep_counter: counter too large");
    h->sweep_counter = counter;
    return PYMC_OK;
}

// recompute bond/second into h->reduced and copy them over the tracked values (all async)
static int slab_wait_async(pymc_handle *h);

static int rebase_tracked(pymc_handle *h) {
    Lat L = make_lat(h);
    {
        int rc = slab_wait_async(h);
        if (rc) return rc;
    }
    CUDA_TRY(cudaMemsetAsync(h->reduced, 0, (size_t)h->p.nreplicas * 16, h->stream));
    dim3 grid((2 * h->p.lenX * h->nw + 255) / 256, h->p.nreplicas);
#define CALL_REDUCE(LAY, MOD) reduce_kernel<LAY, MOD><<<grid, 256, 0, h->stream>>>(L, h->reduced)
    DISPATCH(h, CALL_REDUCE);
#undef CALL_REDUCE
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(h->tracked, h->reduced, (size_t)h->p.nreplicas * 16, cudaMemcpyDeviceToDevice, h->stream));
    h->dirty = false;
    return PYMC_OK;
}

static int fetch_pairs(pymc_handle *h, const unsigned long long *dev, int64_t *bond, int64_t *second) {
    const int n = h->p.nreplicas;
    long long *tmp = new (std::nothrow) long long[2 * (size_t)n];
    REQUIRE(tmp, PYMC_ERR_BADARG, "out of host memory");
    cudaError_t e = cudaMemcpyAsync(tmp, dev, (size_t)n * 16, cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    if (e != cudaSuccess) {
        delete[] tmp;
        return fail(PYMC_ERR_CUDA, "observable copy failed: %s", cudaGetErrorString(e));
    }
    for (int r = 0; r < n; ++r) {
        if (bond) bond[r] = tmp[2 * r];
        if (second) second[r] = tmp[2 * r + 1];
    }
    delete[] tmp;
    CHECK_ASYNC(h);
    return PYMC_OK;
}

int pymc_reduce_observables(pymc_handle *h, int64_t *bond, int64_t *second) {
    ENTER(h);
    int rc = rebase_tracked(h);
    if (rc) return rc;
    return fetch_pairs(h, h->reduced, bond, second);
}

int pymc_get_tracked(pymc_handle *h, int64_t *bond, int64_t *second) {
    ENTER(h);
    if (h->dirty) {
        int rc = rebase_tracked(h);
        if (rc) return rc;
    }
    return fetch_pairs(h, h->tracked, bond, second);
}

int pymc_get_tracked_async(pymc_handle *h, int64_t *pairs) {
    ENTER(h);
    REQUIRE(pairs, PYMC_ERR_BADARG, "pymc_get_tracked_async: null pointer");
    if (h->dirty) {
        int rc = rebase_tracked(h);
        if (rc) return rc;
    }
    CUDA_TRY(cudaMemcpyAsync(pairs, h->tracked, (size_t)h->p.nreplicas * 16, cudaMemcpyDeviceToHost, h->stream));
    return PYMC_OK;
}

int pymc_propose(pymc_handle *h, int replica, const int32_t *x, const int32_t *y, int n, int32_t *site_out,
                 int32_t *nnsum_out) {
    ENTER(h);
    REQUIRE(x && y && site_out && nnsum_out, PYMC_ERR_BADARG, "pymc_propose: null pointer");
    REQUIRE(!h->slab, PYMC_ERR_UNSUPPORTED, "pymc_propose: not available on slab handles");
    REQUIRE(replica >= 0 && replica < h->p.nreplicas, PYMC_ERR_BADARG, "pymc_propose: replica %d of %d", replica,
            h->p.nreplicas);
    REQUIRE(n >= 0, PYMC_ERR_BADARG, "pymc_propose: n = %d", n);
    if (n == 0) return PYMC_OK;
    for (int t = 0; t < n; ++t)
        REQUIRE(x[t] >= 0 && x[t] < h->p.lenX && y[t] >= 0 && y[t] < h->p.lenY, PYMC_ERR_BADARG,
                "pymc_propose: site %d = (%d,%d) outside %dx%d", t, x[t], y[t], h->p.lenX, h->p.lenY);
    if (n > h->qcap) {
        cudaFree(h->qbuf);
        h->qbuf = nullptr;
        h->qcap = 0;
        CUDA_TRY(cudaMalloc(&h->qbuf, (size_t)n * 16));
        h->qcap = n;
    }
    int32_t *dx = h->qbuf, *dy = dx + n, *ds = dy + n, *dn = ds + n;
    CUDA_TRY(cudaMemcpyAsync(dx, x, (size_t)n * 4, cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(cudaMemcpyAsync(dy, y, (size_t)n * 4, cudaMemcpyHostToDevice, h->stream));
    Lat L = make_lat(h);
#define CALL_PROPOSE(LAY, MOD) \
    propose_kernel<LAY, MOD><<<(n + 127) / 128, 128, 0, h->stream>>>(L, replica, dx, dy, n, ds, dn)
    DISPATCH(h, CALL_PROPOSE);
#undef CALL_PROPOSE
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(site_out, ds, (size_t)n * 4, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaMemcpyAsync(nnsum_out, dn, (size_t)n * 4, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    return PYMC_OK;
}


// ------------------------------------------------------------------------------------------
// replica exchange
// ------------------------------------------------------------------------------------------
int pymc_rx_setup(pymc_handle *h, const double *betas, int ntemps, double c_bond, double c_second) {
    ENTER(h);
    REQUIRE(betas, PYMC_ERR_BADARG, "pymc_rx_setup: null pointer");
    REQUIRE(ntemps == h->p.ntemps && ntemps == h->p.nreplicas_global, PYMC_ERR_BADARG,
            "pymc_rx_setup: replica exchange needs ntemps == nreplicas_global (got %d temperatures, handle has %d "
            "temperatures and %d replicas)", ntemps, h->p.ntemps, h->p.nreplicas_global);
    if (!h->betas) {
        CUDA_TRY(cudaMalloc(&h->betas, (size_t)ntemps * 8));
        CUDA_TRY(cudaMalloc(&h->energies, (size_t)ntemps * 8));
        CUDA_TRY(cudaMalloc(&h->T_to_rep, (size_t)ntemps * 4));
        CUDA_TRY(cudaMalloc(&h->rx_tried, (size_t)ntemps * 8));
        CUDA_TRY(cudaMalloc(&h->rx_accepted, (size_t)ntemps * 8));
    }
    CUDA_TRY(cudaMemcpyAsync(h->betas, betas, (size_t)ntemps * 8, cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(cudaMemsetAsync(h->energies, 0, (size_t)ntemps * 8, h->stream));
    CUDA_TRY(cudaMemsetAsync(h->rx_tried, 0, (size_t)ntemps * 8, h->stream));
    CUDA_TRY(cudaMemsetAsync(h->rx_accepted, 0, (size_t)ntemps * 8, h->stream));
    rx_invert_map_kernel<<<(ntemps + 255) / 256, 256, 0, h->stream>>>(h->rep_to_T, h->T_to_rep, ntemps);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    h->c_bond = c_bond;
    h->c_second = c_second;
    h->rx_ready = true;
    return PYMC_OK;
}

int pymc_rx_energies(pymc_handle *h) {
    ENTER(h);
    REQUIRE(h->rx_ready, PYMC_ERR_STATE, "pymc_rx_energies: call pymc_rx_setup first");
    if (h->dirty) {
        int rc = rebase_tracked(h);
        if (rc) return rc;
    }
    rx_energy_kernel<<<(h->p.nreplicas + 127) / 128, 128, 0, h->stream>>>(h->tracked, h->p.nreplicas, h->p.replica_offset,
                                                                        h->c_bond, h->c_second, h->energies);
    CUDA_TRY(cudaGetLastError());
    return PYMC_OK;
}

int pymc_rx_buffers(pymc_handle *h, void **energies_all, void **replica_to_T) {
    REQUIRE(h, PYMC_ERR_BADARG, "pymc_rx_buffers: null handle");
    REQUIRE(h->rx_ready, PYMC_ERR_STATE, "pymc_rx_buffers: call pymc_rx_setup first");
    if (energies_all) *energies_all = h->energies;
    if (replica_to_T) *replica_to_T = h->rep_to_T;
    return PYMC_OK;
}

static int rx_exchange_async(pymc_handle *h, int xcscheme) {
    const int npairs = (h->p.ntemps + 1) / 2;
    rx_decide_kernel<<<(npairs + 127) / 128, 128, 0, h->stream>>>(h->energies, h->betas, h->rep_to_T, h->T_to_rep,
                                                                 h->p.ntemps, xcscheme, h->rx_round,
                                                                 make_key(h->p.seed), h->rx_tried, h->rx_accepted);
    h->rx_round++;
    return PYMC_OK;
}

int pymc_rx_exchange(pymc_handle *h, int xcscheme) {
    ENTER(h);
    REQUIRE(h->rx_ready, PYMC_ERR_STATE, "pymc_rx_exchange: call pymc_rx_setup first");
    REQUIRE(xcscheme == 0 || xcscheme == 1, PYMC_ERR_BADARG, "pymc_rx_exchange: scheme %d", xcscheme);
    rx_exchange_async(h, xcscheme);
    CUDA_TRY(cudaGetLastError());
    return PYMC_OK;
}

int pymc_rx_cycle(pymc_handle *h, int64_t ncycles, int64_t sweeps_per_exchange, int first_scheme) {
    ENTER(h);
    REQUIRE(h->rx_ready, PYMC_ERR_STATE, "pymc_rx_cycle: call pymc_rx_setup first");
    const bool partitioned = h->p.nreplicas != h->p.nreplicas_global;
    REQUIRE(!partitioned || (h->rx_nranks > 1 && h->pbatch && sweeps_per_exchange <= 4096), PYMC_ERR_STATE,
            "pymc_rx_cycle: this handle holds %d of %d replicas; connect the GPUs with pymc_rx_ipc_export / pymc_rx_connect "
            "(batch-resident kernel), or all-gather the energies yourself between pymc_rx_energies and pymc_rx_exchange",
            h->p.nreplicas, h->p.nreplicas_global);
    REQUIRE(ncycles >= 0 && sweeps_per_exchange >= 1 && (first_scheme == 0 || first_scheme == 1), PYMC_ERR_BADARG,
            "pymc_rx_cycle: bad arguments");
    int scheme = first_scheme;
    if (h->pbatch && ncycles > 0 && sweeps_per_exchange <= 4096) {
        // everything in one kernel: sweeps, exact energies, grid barrier, pair decisions (pmc_pbatch.cuh)
        const int rc = launch_pbatch(h, ncycles * sweeps_per_exchange, (int)sweeps_per_exchange, &scheme);
        if (rc != PYMC_ERR_UNSUPPORTED || partitioned) return rc;  // (the other GPUs cannot follow a fall-back)
    }
    for (int64_t c = 0; c < ncycles; ++c) {
        int rc = pymc_sweep(h, sweeps_per_exchange);
        if (rc) return rc;
        rc = pymc_rx_energies(h);
        if (rc) return rc;
        rx_exchange_async(h, scheme);
        scheme ^= 1;
    }
    CUDA_TRY(cudaGetLastError());
    return PYMC_OK;
}

// ---- the ladder over several GPUs: exchange of the energies through peer memory, inside the sweep kernel ------
// Each process exports two CUDA IPC handles (its per-replica sums and its flag words); pymc_rx_connect maps those of
// all other GPUs.  After that pymc_rx_cycle on a partitioned handle is one launch per GPU: at an exchange, block 0
// stores the sums of its replicas into every other GPU's array over NVLink and raises a flag there; the decisions
// are then made on every GPU from identical data.  All GPUs must issue the same pymc_rx_cycle calls.
int pymc_rx_ipc_export(pymc_handle *h, void *handles, size_t nbytes) {
    ENTER(h);
    REQUIRE(h->rx_ready, PYMC_ERR_STATE, "pymc_rx_ipc_export: call pymc_rx_setup first");
    REQUIRE(handles && nbytes == 2 * sizeof(cudaIpcMemHandle_t), PYMC_ERR_BADARG, "pymc_rx_ipc_export: need %zu bytes",
            2 * sizeof(cudaIpcMemHandle_t));
    REQUIRE(h->pbatch, PYMC_ERR_UNSUPPORTED,
            "pymc_rx_ipc_export: the in-kernel exchange needs the batch-resident kernel (PYMC_KERNEL_PBATCH or AUTO where it applies)");
    int rc = ensure_rx_exchange_buffers(h);
    if (rc) return rc;
    cudaIpcMemHandle_t *out = (cudaIpcMemHandle_t *)handles;
    CUDA_TRY(cudaIpcGetMemHandle(out + 0, h->pb_xpk));
    CUDA_TRY(cudaIpcGetMemHandle(out + 1, h->pb_xflags));
    return PYMC_OK;
}

int pymc_rx_connect(pymc_handle *h, const void *all_handles, size_t nbytes, int nranks, int rank) {
    ENTER(h);
    REQUIRE(h->rx_ready && h->pb_esum, PYMC_ERR_STATE, "pymc_rx_connect: call pymc_rx_setup and pymc_rx_ipc_export first");
    REQUIRE(nranks >= 2 && nranks <= PB_MAXRANKS && rank >= 0 && rank < nranks, PYMC_ERR_BADARG,
            "pymc_rx_connect: rank %d of %d (at most %d GPUs)", rank, nranks, PB_MAXRANKS);
    REQUIRE(all_handles && nbytes == (size_t)nranks * 2 * sizeof(cudaIpcMemHandle_t), PYMC_ERR_BADARG,
            "pymc_rx_connect: need %zu bytes of handles", (size_t)nranks * 2 * sizeof(cudaIpcMemHandle_t));
    REQUIRE(h->p.nreplicas * nranks == h->p.nreplicas_global && h->p.replica_offset == rank * h->p.nreplicas, PYMC_ERR_BADARG,
            "pymc_rx_connect: the replicas must be block-partitioned evenly, rank g holding [g*R/G, (g+1)*R/G)");
    REQUIRE(h->rx_nranks <= 1, PYMC_ERR_STATE, "pymc_rx_connect: already connected");
    const cudaIpcMemHandle_t *hd = (const cudaIpcMemHandle_t *)all_handles;
    for (int g = 0; g < nranks; ++g) {
        if (g == rank) continue;
        void *p0 = nullptr, *p1 = nullptr;
        CUDA_TRY(cudaIpcOpenMemHandle(&p0, hd[2 * g + 0], cudaIpcMemLazyEnablePeerAccess));
        CUDA_TRY(cudaIpcOpenMemHandle(&p1, hd[2 * g + 1], cudaIpcMemLazyEnablePeerAccess));
        h->peer_xpk[g] = (PBPacket *)p0;
        h->peer_xflags[g] = (unsigned int *)p1;
    }
    h->rx_nranks = nranks;
    h->rx_rank = rank;
    return PYMC_OK;
}

int pymc_rx_get_energies(pymc_handle *h, double *energies, int n) {
    ENTER(h);
    REQUIRE(h->rx_ready, PYMC_ERR_STATE, "pymc_rx_get_energies: call pymc_rx_setup first");
    REQUIRE(energies && n == h->p.nreplicas_global, PYMC_ERR_BADARG, "pymc_rx_get_energies: need %d entries",
            h->p.nreplicas_global);
    CUDA_TRY(cudaMemcpyAsync(energies, h->energies, (size_t)n * 8, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    CHECK_ASYNC(h);
    return PYMC_OK;
}

int pymc_rx_get_stats(pymc_handle *h, uint64_t *tried, uint64_t *accepted, int n) {
    ENTER(h);
    REQUIRE(h->rx_ready, PYMC_ERR_STATE, "pymc_rx_get_stats: call pymc_rx_setup first");
    REQUIRE(tried && accepted && n == h->p.ntemps, PYMC_ERR_BADARG, "pymc_rx_get_stats: need %d entries", h->p.ntemps);
    CUDA_TRY(cudaMemcpyAsync(tried, h->rx_tried, (size_t)n * 8, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaMemcpyAsync(accepted, h->rx_accepted, (size_t)n * 8, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    CHECK_ASYNC(h);
    return PYMC_OK;
}

int pymc_rx_get_round(pymc_handle *h, uint64_t *round) {
    REQUIRE(h && round, PYMC_ERR_BADARG, "pymc_rx_get_round: null pointer");
    *round = h->rx_round;
    return PYMC_OK;
}

int pymc_rx_set_round(pymc_handle *h, uint64_t round) {
    REQUIRE(h, PYMC_ERR_BADARG, "pymc_rx_set_round: null handle");
    if (h->pb_xpk && round < h->rx_round) {  // the packets of other GPUs carry rounds: start them again (do this on every GPU,
        DeviceGuard guard(h->p.device);     // then synchronise the processes, before the next pymc_rx_cycle)
        CUDA_TRY(cudaStreamSynchronize(h->stream));
        CUDA_TRY(cudaMemset(h->pb_xpk, 0, (size_t)2 * h->p.ntemps * 2 * sizeof(PBPacket)));
    }
    h->rx_round = round;
    return PYMC_OK;
}


// ------------------------------------------------------------------------------------------
// particle-hole exchange (lattice gas)
// ------------------------------------------------------------------------------------------
int pymc_set_pair_tables(pymc_handle *h, const uint32_t *thr, int ntemps, int nclass) {
    ENTER(h);
    REQUIRE(thr, PYMC_ERR_BADARG, "pymc_set_pair_tables: null pointer");
    REQUIRE(h->p.model == PYMC_MODEL_LATGAS, PYMC_ERR_UNSUPPORTED, "pymc_set_pair_tables: lattice-gas handles only");
    REQUIRE(ntemps == h->p.ntemps && nclass == PAIR_NCLASS, PYMC_ERR_BADARG,
            "pymc_set_pair_tables: got %d x %d, handle expects %d x %d", ntemps, nclass, h->p.ntemps, PAIR_NCLASS);
    REQUIRE(h->p.lenX % 4 == 0 && h->p.lenY % 4 == 0, PYMC_ERR_UNSUPPORTED,
            "pymc_set_pair_tables: the bond-class schedule needs lenX %% 4 == 0 and lenY %% 4 == 0 (got %dx%d)",
            h->p.lenX, h->p.lenY);
    uint32_t *rows = new (std::nothrow) uint32_t[(size_t)ntemps * PLAN_WORDS];
    REQUIRE(rows, PYMC_ERR_BADARG, "pymc_set_pair_tables: out of host memory");
    for (int t = 0; t < ntemps; ++t) {
        if (!PairPlan::build_row(thr + (size_t)t * nclass, rows + (size_t)t * PLAN_WORDS)) {
            delete[] rows;
            return fail(PYMC_ERR_UNSUPPORTED,
                        "pymc_set_pair_tables: row %d is not a pair-exchange table (dE > 0 on both sides or at d = 0)", t);
        }
    }
    cudaError_t e = cudaSuccess;
    if (!h->pair_plans) e = cudaMalloc(&h->pair_plans, (size_t)ntemps * PLAN_WORDS * 4);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(h->pair_plans, rows, (size_t)ntemps * PLAN_WORDS * 4, cudaMemcpyHostToDevice, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    delete[] rows;
    if (e != cudaSuccess) return fail(PYMC_ERR_CUDA, "pymc_set_pair_tables: %s", cudaGetErrorString(e));
    return PYMC_OK;
}

int pymc_pair_sweep(pymc_handle *h, int64_t nsweeps, uint32_t class_mask) {
    ENTER(h);
    REQUIRE(nsweeps >= 0, PYMC_ERR_BADARG, "pymc_pair_sweep: nsweeps %lld", (long long)nsweeps);
    REQUIRE(h->pair_plans, PYMC_ERR_STATE, "pymc_pair_sweep: call pymc_set_pair_tables first");
    REQUIRE(!h->slab, PYMC_ERR_UNSUPPORTED, "pymc_pair_sweep: not available on slab handles");
    REQUIRE((class_mask & 0xFFFFu) != 0 && (class_mask >> 16) == 0, PYMC_ERR_BADARG, "pymc_pair_sweep: class mask %#x",
            class_mask);
    if (h->resident) return launch_resident(h, nsweeps, false, class_mask, class_mask);
    if (h->dirty) {
        int rc = rebase_tracked(h);
        if (rc) return rc;
    }
    Lat L = make_lat(h);
    for (int64_t s = 0; s < nsweeps; ++s) {
        for (int c = 0; c < 16; ++c) {
            if (!((class_mask >> c) & 1u)) continue;
            const int nrows = h->p.lenX / (c < 8 ? 2 : 4);
            dim3 grid((nrows * h->nw + 127) / 128, h->p.nreplicas);
            const uint64_t step = 16 * h->pair_counter + c;
            if (h->p.layout == PYMC_LAYOUT_BITS)
                pair_class_kernel<PYMC_LAYOUT_BITS><<<grid, 128, 0, h->stream>>>(L, c, step);
            else
                pair_class_kernel<PYMC_LAYOUT_INT8><<<grid, 128, 0, h->stream>>>(L, c, step);
        }
        h->pair_counter++;
    }
    CUDA_TRY(cudaGetLastError());
    return PYMC_OK;
}

int pymc_mixed_sweep(pymc_handle *h, int64_t nsteps, uint32_t mask_even, uint32_t mask_odd) {
    ENTER(h);
    REQUIRE(nsteps >= 0, PYMC_ERR_BADARG, "pymc_mixed_sweep: nsteps %lld", (long long)nsteps);
    REQUIRE((mask_even >> 16) == 0 && (mask_odd >> 16) == 0, PYMC_ERR_BADARG, "pymc_mixed_sweep: class masks %#x %#x",
            mask_even, mask_odd);
    REQUIRE(h->pair_plans || !(mask_even | mask_odd), PYMC_ERR_STATE, "pymc_mixed_sweep: call pymc_set_pair_tables first");
    REQUIRE(!h->slab, PYMC_ERR_UNSUPPORTED, "pymc_mixed_sweep: not available on slab handles");
    if (h->resident) return launch_resident(h, nsteps, true, mask_even, mask_odd);
    for (int64_t s = 0; s < nsteps; ++s) {
        int rc = pymc_sweep(h, 1);
        if (rc) return rc;
        const uint32_t mask = (h->pair_counter & 1ull) ? mask_odd : mask_even;
        if (mask && (rc = pymc_pair_sweep(h, 1, mask))) return rc;
    }
    return PYMC_OK;
}

int pymc_get_pair_counter(pymc_handle *h, uint64_t *counter) {
    REQUIRE(h && counter, PYMC_ERR_BADARG, "pymc_get_pair_counter: null pointer");
    *counter = h->pair_counter;
    return PYMC_OK;
}

int pymc_set_pair_counter(pymc_handle *h, uint64_t counter) {
    REQUIRE(h, PYMC_ERR_BADARG, "pymc_set_pair_counter: null handle");
    REQUIRE(counter < (1ull << 51), PYMC_ERR_BADARG, "pymc_set_pair_counter: counter too large");
    h->pair_counter = counter;
    return PYMC_OK;
}


// ------------------------------------------------------------------------------------------
// hard discs (examples/2D_hardcore)
// ------------------------------------------------------------------------------------------
struct pymc_discs {
    int device;
    cudaStream_t stream;
    DiscParams P;
    int *d_overlap;        // [nrep]
    long long *d_counts;   // [nrep][cap_bins]
    int cap_bins;
    int *d_probe_i;        // id, overlap
    double *d_probe_d;     // new x, y
    uint64_t trial_counter;
};

#define ENTER_D(d)                                                        \
    REQUIRE((d) != nullptr, PYMC_ERR_BADARG, "%s: null handle", __func__); \
    DeviceGuard _guard((d)->device);                                      \
    REQUIRE(_guard.ok, PYMC_ERR_CUDA, "%s: cannot select device %d", __func__, (d)->device)

int pymc_discs_destroy(pymc_discs *d) {
    if (!d) return PYMC_OK;
    DeviceGuard guard(d->device);
    if (guard.ok) {
        if (d->stream) cudaStreamSynchronize(d->stream);
        cudaFree(d->P.coors);
        cudaFree(d->P.accepted);
        cudaFree(d->d_overlap);
        cudaFree(d->d_counts);
        cudaFree(d->d_probe_i);
        cudaFree(d->d_probe_d);
        if (d->stream) cudaStreamDestroy(d->stream);
    }
    delete d;
    return PYMC_OK;
}

int pymc_discs_create(int device, int nreplicas, int replica_offset, int ndisc, double Lcell, double rdisc,
                      double maxstep, uint64_t seed, int isotropic, pymc_discs **out) {
    REQUIRE(out, PYMC_ERR_BADARG, "pymc_discs_create: null pointer");
    *out = nullptr;
    REQUIRE(nreplicas >= 1 && replica_offset >= 0, PYMC_ERR_BADARG, "pymc_discs_create: nreplicas %d offset %d",
            nreplicas, replica_offset);
    REQUIRE(ndisc >= 1 && ndisc <= 32, PYMC_ERR_UNSUPPORTED,
            "pymc_discs_create: 1..32 discs per replica are supported (one disc per warp lane), got %d", ndisc);
    REQUIRE(Lcell > 0.0 && rdisc > 0.0, PYMC_ERR_BADARG, "pymc_discs_create: Lcell and rdisc must be positive");
    int ndev = 0;
    CUDA_TRY(cudaGetDeviceCount(&ndev));
    REQUIRE(device >= 0 && device < ndev, PYMC_ERR_BADARG, "pymc_discs_create: device %d of %d", device, ndev);
    pymc_discs *d = new (std::nothrow) pymc_discs();
    REQUIRE(d, PYMC_ERR_BADARG, "pymc_discs_create: out of host memory");
    memset(d, 0, sizeof(*d));
    d->device = device;
    d->P.nrep = nreplicas;
    d->P.ndisc = ndisc;
    d->P.Lcell = Lcell;
    d->P.two_r = rdisc * 2;
    d->P.maxstep = maxstep;
    d->P.isotropic = isotropic ? 1 : 0;
    d->P.key = make_key(seed);
    d->P.replica_offset = replica_offset;
    DeviceGuard guard(device);
    cudaError_t e = guard.ok ? cudaSuccess : cudaErrorInvalidDevice;
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&d->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc(&d->P.coors, (size_t)nreplicas * ndisc * 16);
    if (e == cudaSuccess) e = cudaMalloc(&d->P.accepted, (size_t)nreplicas * 8);
    if (e == cudaSuccess) e = cudaMalloc(&d->d_overlap, (size_t)nreplicas * 4);
    if (e == cudaSuccess) e = cudaMalloc(&d->d_probe_i, 8);
    if (e == cudaSuccess) e = cudaMalloc(&d->d_probe_d, 16);
    if (e == cudaSuccess) e = cudaMemsetAsync(d->P.coors, 0, (size_t)nreplicas * ndisc * 16, d->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(d->P.accepted, 0, (size_t)nreplicas * 8, d->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(d->stream);
    if (e != cudaSuccess) {
        pymc_discs_destroy(d);
        return fail(PYMC_ERR_CUDA, "pymc_discs_create: %s", cudaGetErrorString(e));
    }
    *out = d;
    return PYMC_OK;
}

int pymc_discs_upload(pymc_discs *d, int replica, const double *coors) {
    ENTER_D(d);
    REQUIRE(coors && replica >= 0 && replica < d->P.nrep, PYMC_ERR_BADARG, "pymc_discs_upload: bad arguments");
    CUDA_TRY(cudaMemcpyAsync(d->P.coors + (size_t)replica * d->P.ndisc * 2, coors, (size_t)d->P.ndisc * 16,
                             cudaMemcpyHostToDevice, d->stream));
    CUDA_TRY(cudaStreamSynchronize(d->stream));
    return PYMC_OK;
}

int pymc_discs_download(pymc_discs *d, int replica, double *coors) {
    ENTER_D(d);
    REQUIRE(coors && replica >= 0 && replica < d->P.nrep, PYMC_ERR_BADARG, "pymc_discs_download: bad arguments");
    CUDA_TRY(cudaMemcpyAsync(coors, d->P.coors + (size_t)replica * d->P.ndisc * 2, (size_t)d->P.ndisc * 16,
                             cudaMemcpyDeviceToHost, d->stream));
    CUDA_TRY(cudaStreamSynchronize(d->stream));
    return PYMC_OK;
}

int pymc_discs_run(pymc_discs *d, int64_t ntrials) {
    ENTER_D(d);
    REQUIRE(ntrials >= 0, PYMC_ERR_BADARG, "pymc_discs_run: ntrials %lld", (long long)ntrials);
    if (ntrials == 0) return PYMC_OK;
    discs_run_kernel<<<(d->P.nrep + 3) / 4, 128, 0, d->stream>>>(d->P, d->trial_counter, (long long)ntrials);
    CUDA_TRY(cudaGetLastError());
    d->trial_counter += (uint64_t)ntrials;
    return PYMC_OK;
}

int pymc_discs_propose(pymc_discs *d, int replica, int *id, double *newcoor, int *overlap) {
    ENTER_D(d);
    REQUIRE(id && newcoor && overlap && replica >= 0 && replica < d->P.nrep, PYMC_ERR_BADARG,
            "pymc_discs_propose: bad arguments");
    discs_propose_kernel<<<1, 32, 0, d->stream>>>(d->P, replica, d->trial_counter, d->d_probe_i, d->d_probe_d,
                                                  d->d_probe_i + 1);
    CUDA_TRY(cudaGetLastError());
    int hi[2];
    CUDA_TRY(cudaMemcpyAsync(hi, d->d_probe_i, 8, cudaMemcpyDeviceToHost, d->stream));
    CUDA_TRY(cudaMemcpyAsync(newcoor, d->d_probe_d, 16, cudaMemcpyDeviceToHost, d->stream));
    CUDA_TRY(cudaStreamSynchronize(d->stream));
    *id = hi[0];
    *overlap = hi[1];
    return PYMC_OK;
}

int pymc_discs_observe(pymc_discs *d, double dr, int nbins, int32_t *overlap, int64_t *counts) {
    ENTER_D(d);
    REQUIRE(nbins >= 0 && (nbins == 0 || (dr > 0.0 && counts)), PYMC_ERR_BADARG, "pymc_discs_observe: bad histogram");
    if (nbins > d->cap_bins) {
        cudaFree(d->d_counts);
        d->d_counts = nullptr;
        d->cap_bins = 0;
        CUDA_TRY(cudaMalloc(&d->d_counts, (size_t)d->P.nrep * nbins * 8));
        d->cap_bins = nbins;
    }
    if (nbins) CUDA_TRY(cudaMemsetAsync(d->d_counts, 0, (size_t)d->P.nrep * nbins * 8, d->stream));
    discs_observe_kernel<<<(d->P.nrep + 3) / 4, 128, 0, d->stream>>>(d->P, dr, nbins, d->d_overlap,
                                                                     nbins ? d->d_counts : nullptr);
    CUDA_TRY(cudaGetLastError());
    if (overlap) CUDA_TRY(cudaMemcpyAsync(overlap, d->d_overlap, (size_t)d->P.nrep * 4, cudaMemcpyDeviceToHost, d->stream));
    if (nbins) CUDA_TRY(cudaMemcpyAsync(counts, d->d_counts, (size_t)d->P.nrep * nbins * 8, cudaMemcpyDeviceToHost, d->stream));
    CUDA_TRY(cudaStreamSynchronize(d->stream));
    return PYMC_OK;
}

int pymc_discs_counters(pymc_discs *d, uint64_t *trial_counter, uint64_t *accepted) {
    ENTER_D(d);
    if (trial_counter) *trial_counter = d->trial_counter;
    if (accepted) {
        CUDA_TRY(cudaMemcpyAsync(accepted, d->P.accepted, (size_t)d->P.nrep * 8, cudaMemcpyDeviceToHost, d->stream));
        CUDA_TRY(cudaStreamSynchronize(d->stream));
    }
    return PYMC_OK;
}

int pymc_discs_set_trial_counter(pymc_discs *d, uint64_t counter) {
    REQUIRE(d, PYMC_ERR_BADARG, "pymc_discs_set_trial_counter: null handle");
    d->trial_counter = counter;
    return PYMC_OK;
}

int pymc_discs_timer(pymc_discs *d, int64_t ntrials, float *elapsed_ms) {
    ENTER_D(d);
    REQUIRE(elapsed_ms, PYMC_ERR_BADARG, "pymc_discs_timer: null pointer");
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0));
    CUDA_TRY(cudaEventCreate(&e1));
    CUDA_TRY(cudaEventRecord(e0, d->stream));
    int rc = pymc_discs_run(d, ntrials);
    cudaEventRecord(e1, d->stream);
    cudaEventSynchronize(e1);
    cudaEventElapsedTime(elapsed_ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return rc;
}


// ------------------------------------------------------------------------------------------
// device-side statistics
// ------------------------------------------------------------------------------------------
int pymc_stats_reset(pymc_handle *h, double c_bond, double c_second) {
    ENTER(h);
    if (!h->stats) {
        CUDA_TRY(cudaMalloc(&h->stats, (size_t)h->p.ntemps * NSTAT * 8));
        CUDA_TRY(cudaMalloc(&h->stats_count, (size_t)h->p.ntemps * 8));
    }
    h->c_bond = c_bond;
    h->c_second = c_second;
    CUDA_TRY(cudaMemsetAsync(h->stats, 0, (size_t)h->p.ntemps * NSTAT * 8, h->stream));
    CUDA_TRY(cudaMemsetAsync(h->stats_count, 0, (size_t)h->p.ntemps * 8, h->stream));
    return PYMC_OK;
}

int pymc_stats_accumulate(pymc_handle *h) {
    ENTER(h);
    REQUIRE(h->stats, PYMC_ERR_STATE, "pymc_stats_accumulate: call pymc_stats_reset first");
    if (h->dirty) {
        int rc = rebase_tracked(h);
        if (rc) return rc;
    }
    stats_accumulate_kernel<<<(h->p.nreplicas + 127) / 128, 128, 0, h->stream>>>(
        h->tracked, h->p.nreplicas, h->p.replica_offset, h->rep_to_T, h->c_bond, h->c_second, h->stats, h->stats_count);
    CUDA_TRY(cudaGetLastError());
    return PYMC_OK;
}

int pymc_stats_get(pymc_handle *h, double *sums, uint64_t *counts, int ntemps, int nstat) {
    ENTER(h);
    REQUIRE(h->stats, PYMC_ERR_STATE, "pymc_stats_get: call pymc_stats_reset first");
    REQUIRE(sums && counts && ntemps == h->p.ntemps && nstat == NSTAT, PYMC_ERR_BADARG,
            "pymc_stats_get: need [%d][%d] sums", h->p.ntemps, NSTAT);
    CUDA_TRY(cudaMemcpyAsync(sums, h->stats, (size_t)ntemps * NSTAT * 8, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaMemcpyAsync(counts, h->stats_count, (size_t)ntemps * 8, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    CHECK_ASYNC(h);
    return PYMC_OK;
}


// ------------------------------------------------------------------------------------------
// time series by temperature + blocking analysis on the device
// ------------------------------------------------------------------------------------------
#define PYMC_SERIES_MAXLEVELS 31

int pymc_series_reset(pymc_handle *h, int capacity, double c_bond, double c_second) {
    ENTER(h);
    REQUIRE(capacity >= 1, PYMC_ERR_BADARG, "pymc_series_reset: capacity %d", capacity);
    if (capacity > h->series_cap) {
        CUDA_TRY(cudaStreamSynchronize(h->stream));
        cudaFree(h->series);
        cudaFree(h->series_work);
        h->series = h->series_work = nullptr;
        h->series_cap = 0;
        CUDA_TRY(cudaMalloc(&h->series, (size_t)capacity * h->p.ntemps * 2 * sizeof(double)));
        CUDA_TRY(cudaMalloc(&h->series_work, (size_t)capacity * h->p.ntemps * 4 * sizeof(double)));
        h->series_cap = capacity;
    }
    if (!h->series_out) CUDA_TRY(cudaMalloc(&h->series_out, (size_t)h->p.ntemps * 2 * (PYMC_SERIES_MAXLEVELS + 2) * sizeof(double)));
    CUDA_TRY(cudaMemsetAsync(h->series, 0, (size_t)h->series_cap * h->p.ntemps * 2 * sizeof(double), h->stream));
    h->series_count = 0;
    h->c_bond = c_bond;
    h->c_second = c_second;
    return PYMC_OK;
}

int pymc_series_record(pymc_handle *h) {
    ENTER(h);
    REQUIRE(h->series, PYMC_ERR_STATE, "pymc_series_record: call pymc_series_reset first");
    REQUIRE(h->series_count < h->series_cap, PYMC_ERR_STATE, "pymc_series_record: the series is full (%d samples)", h->series_cap);
    if (h->dirty) {
        int rc = rebase_tracked(h);
        if (rc) return rc;
    }
    series_record_kernel<<<(h->p.nreplicas + 127) / 128, 128, 0, h->stream>>>(
        h->tracked, h->p.nreplicas, h->p.replica_offset, h->rep_to_T, h->c_bond, h->c_second, h->p.ntemps,
        h->series + (size_t)h->series_count * h->p.ntemps * 2);
    CUDA_TRY(cudaGetLastError());
    h->series_count++;
    return PYMC_OK;
}

int pymc_series_count(pymc_handle *h, int *count) {
    REQUIRE(h && count, PYMC_ERR_BADARG, "pymc_series_count: null pointer");
    *count = h->series_count;
    return PYMC_OK;
}

int pymc_series_get(pymc_handle *h, double *out, size_t n) {
    ENTER(h);
    REQUIRE(h->series && out && n == (size_t)h->series_count * h->p.ntemps * 2, PYMC_ERR_BADARG,
            "pymc_series_get: need %zu values", h->series ? (size_t)h->series_count * h->p.ntemps * 2 : (size_t)0);
    if (n) CUDA_TRY(cudaMemcpyAsync(out, h->series, n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    CHECK_ASYNC(h);
    return PYMC_OK;
}

int pymc_series_put(pymc_handle *h, const double *in, int count) {
    ENTER(h);
    REQUIRE(h->series && in && count >= 0 && count <= h->series_cap, PYMC_ERR_BADARG, "pymc_series_put: %d samples, capacity %d",
            count, h->series_cap);
    if (count) CUDA_TRY(cudaMemcpyAsync(h->series, in, (size_t)count * h->p.ntemps * 2 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    h->series_count = count;
    return PYMC_OK;
}

int pymc_series_binning(pymc_handle *h, int discard, int nlevels, double *mean, double *err) {
    ENTER(h);
    REQUIRE(h->series && mean && err, PYMC_ERR_BADARG, "pymc_series_binning: call pymc_series_reset first; null pointer");
    REQUIRE(nlevels >= 0 && nlevels <= PYMC_SERIES_MAXLEVELS && discard >= 0 && discard < h->series_count &&
                (h->series_count >> nlevels) >= 2,
            PYMC_ERR_BADARG, "pymc_series_binning: %d samples, discard %d, %d levels (the top level needs two blocks)",
            h->series_count, discard, nlevels);
    const int ns = h->p.ntemps * 2;
    double *d_mean = h->series_out, *d_err = h->series_out + ns;
    series_binning_kernel<<<ns, 256, 0, h->stream>>>(h->series, h->series_count, h->p.ntemps, discard, nlevels, h->series_work,
                                                    h->series_cap, d_mean, d_err);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(mean, d_mean, (size_t)ns * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaMemcpyAsync(err, d_err, (size_t)ns * (nlevels + 1) * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    CHECK_ASYNC(h);
    return PYMC_OK;
}

// ------------------------------------------------------------------------------------------
// row-slab mode: one lattice split over GPUs
// ------------------------------------------------------------------------------------------
__global__ void slab_wait_kernel(volatile unsigned long long *my_flags, unsigned long long epoch, int *err) {
    const long long t0 = clock64();
    while (my_flags[0] < epoch || my_flags[1] < epoch) {
        if (clock64() - t0 > 4000000000LL) {
            *err = 1;
            break;
        }
    }
    __threadfence_system();
}

// Before my halo rows are read outside a sweep (observables): the neighbours must have finished the step
// in which they wrote them.
static int slab_wait_async(pymc_handle *h) {
    if (!h->slab || h->nslabs <= 1 || h->local_peers || !h->dirty_halo) return PYMC_OK;
    slab_wait_kernel<<<1, 1, 0, h->stream>>>(h->flags, h->slab_epoch, h->slab_error);
    CUDA_TRY(cudaGetLastError());
    h->dirty_halo = false;
    return PYMC_OK;
}

static int slab_sync_async(pymc_handle *h) {
    if (h->nslabs <= 1 || h->local_peers) return PYMC_OK;
    h->slab_epoch++;
    slab_sync_kernel<<<1, 1, 0, h->stream>>>(h->flags, h->peer_up_flags, h->peer_dn_flags, h->slab_epoch, h->slab_error,
                                            4000000000LL);
    CUDA_TRY(cudaGetLastError());
    return PYMC_OK;
}

int pymc_create_slab(const pymc_params *p, int slab_index, int nslabs, pymc_handle **out) {
    REQUIRE(p && out, PYMC_ERR_BADARG, "pymc_create_slab: null pointer");
    *out = nullptr;
    REQUIRE(nslabs >= 1 && slab_index >= 0 && slab_index < nslabs, PYMC_ERR_BADARG, "pymc_create_slab: slab %d of %d",
            slab_index, nslabs);
    REQUIRE(p->layout == PYMC_LAYOUT_INT8 && p->kernel == PYMC_KERNEL_AUTO, PYMC_ERR_UNSUPPORTED,
            "pymc_create_slab: slabs use the int8 layout and the fused sweep kernel");
    REQUIRE(p->lenX % nslabs == 0 && (p->lenX / nslabs) % FUSED_TH == 0 && p->lenY % (128 * FUSED_TW) == 0,
            PYMC_ERR_UNSUPPORTED,
            "pymc_create_slab: need lenX/nslabs %% %d == 0 and lenY %% %d == 0 (got %dx%d over %d slabs)", FUSED_TH,
            128 * FUSED_TW, p->lenX, p->lenY, nslabs);
    // slab_push_halos_kernel puts 2 * nreplicas colour planes into gridDim.z (limit 65535)
    REQUIRE(p->nreplicas <= 32767, PYMC_ERR_UNSUPPORTED, "pymc_create_slab: at most 32767 replicas per slab handle (got %d)",
            p->nreplicas);
    // the random streams are keyed by the word index in the WHOLE lattice, a 32-bit number
    REQUIRE((uint64_t)p->lenX * (uint64_t)(p->lenY / 128) < (1ull << 31), PYMC_ERR_UNSUPPORTED,
            "pymc_create_slab: lattice too large (%d x %d over all slabs; at most 2^38 sites)", p->lenX, p->lenY);
    const int Xl = p->lenX / nslabs;
    pymc_params q = *p;
    q.lenX = Xl + 4;  // two halo rows above and below the owned rows
    pymc_handle *h = nullptr;
    int rc = pymc_create(&q, &h);
    if (rc) return rc;
    h->persist = h->pbatch = h->resident = h->resident_big = h->rowstream = false;  // single sweeps: the tiled kernel
    h->slab = true;
    h->own_lo = 2;
    h->own_hi = Xl + 2;
    h->row_shift = slab_index * Xl - 2;
    h->lenX_global = p->lenX;
    h->slab_index = slab_index;
    h->nslabs = nslabs;
    DeviceGuard guard(h->p.device);
    const size_t bytes = (size_t)h->p.nreplicas * 2 * h->p.lenX * h->nk;
    cudaError_t e = cudaMalloc(&h->i8_alt, bytes);
    if (e == cudaSuccess) e = cudaMemset(h->i8_alt, 0, bytes);
    if (e == cudaSuccess) e = cudaMalloc(&h->flags, 16);
    if (e == cudaSuccess) e = cudaMemset(h->flags, 0, 16);
    if (e == cudaSuccess) e = cudaMalloc(&h->slab_done, 8);
    if (e == cudaSuccess) e = cudaMemset(h->slab_done, 0, 8);
    if (e != cudaSuccess) {
        pymc_destroy(h);
        return fail(PYMC_ERR_CUDA, "pymc_create_slab: %s", cudaGetErrorString(e));
    }
    h->buf[0] = h->i8;
    h->buf[1] = h->i8_alt;
    // Calls of >= 2 sweeps keep the slab on chip (K12 on the owned rows; the first / last block exchange packets with the
    // GPUs above / below through their exchange buffers).  Needs one replica, at least two blocks (the rows above row 0
    // and below the last row must come through the exchange buffer, not from inside a block) and an even first row.
    {
        int sms = 0, smem = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->p.device);
        cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->p.device);
        const char *off = getenv("PYMC_NO_PERSISTENT");
        h->pbatch = plan_pbatch(h, h->p.nreplicas, Xl, sms, smem, nullptr) && h->p.nreplicas == 1 && h->pb_blocks >= 2 &&
                    (slab_index * Xl) % 2 == 0 && !(off && off[0] == '1');
        if (h->pbatch && ensure_pbatch_buffers(h) != PYMC_OK) h->pbatch = false;
    }
    if (nslabs == 1) {  // periodic onto myself: my own buffers are my neighbours' buffers
        h->peer_up[0] = h->peer_dn[0] = h->buf[0];
        h->peer_up[1] = h->peer_dn[1] = h->buf[1];
        h->peer_xbuf_up = h->peer_xbuf_dn = h->pb_xbuf;
        h->slab_pbatch = h->pbatch;
    }
    *out = h;
    return PYMC_OK;
}

// 4 x 64 bytes: CUDA IPC handles of buffer 0, buffer 1, the flag words and the exchange buffer of the batch-resident kernel
// (zeros if this slab cannot run it), to be handed to both neighbours
int pymc_slab_ipc_export(pymc_handle *h, void *handles, size_t nbytes) {
    ENTER(h);
    REQUIRE(h->slab && handles && nbytes == 4 * sizeof(cudaIpcMemHandle_t), PYMC_ERR_BADARG,
            "pymc_slab_ipc_export: need a slab handle and %zu bytes", 4 * sizeof(cudaIpcMemHandle_t));
    cudaIpcMemHandle_t *out = (cudaIpcMemHandle_t *)handles;
    memset(out, 0, nbytes);
    CUDA_TRY(cudaIpcGetMemHandle(out + 0, h->buf[0]));
    CUDA_TRY(cudaIpcGetMemHandle(out + 1, h->buf[1]));
    CUDA_TRY(cudaIpcGetMemHandle(out + 2, h->flags));
    if (h->pbatch && h->pb_xbuf) CUDA_TRY(cudaIpcGetMemHandle(out + 3, h->pb_xbuf));
    return PYMC_OK;
}

int pymc_slab_connect(pymc_handle *h, const void *up_handles, const void *dn_handles, size_t nbytes) {
    ENTER(h);
    REQUIRE(h->slab && h->nslabs > 1, PYMC_ERR_STATE, "pymc_slab_connect: only for slabs with neighbours");
    REQUIRE(up_handles && dn_handles && nbytes == 4 * sizeof(cudaIpcMemHandle_t), PYMC_ERR_BADARG,
            "pymc_slab_connect: need two sets of %zu bytes", 4 * sizeof(cudaIpcMemHandle_t));
    const cudaIpcMemHandle_t *up = (const cudaIpcMemHandle_t *)up_handles, *dn = (const cudaIpcMemHandle_t *)dn_handles;
    const cudaIpcMemHandle_t none = {};
    // the exchange buffers are mapped only if all three slabs involved can run the batch-resident kernel
    const bool xb = h->pbatch && memcmp(up + 3, &none, sizeof(none)) != 0 && memcmp(dn + 3, &none, sizeof(none)) != 0;
    const int nh = xb ? 4 : 3;
    void *pu[4] = {nullptr, nullptr, nullptr, nullptr}, *pd[4] = {nullptr, nullptr, nullptr, nullptr};
    for (int k = 0; k < nh; ++k) CUDA_TRY(cudaIpcOpenMemHandle(&pu[k], up[k], cudaIpcMemLazyEnablePeerAccess));
    const bool same = memcmp(up, dn, nbytes) == 0;  // two slabs: the neighbour above is the neighbour below
    for (int k = 0; k < nh; ++k) {
        if (same)
            pd[k] = pu[k];
        else
            CUDA_TRY(cudaIpcOpenMemHandle(&pd[k], dn[k], cudaIpcMemLazyEnablePeerAccess));
    }
    h->peer_up[0] = (int8_t *)pu[0];
    h->peer_up[1] = (int8_t *)pu[1];
    h->peer_up_flags = (unsigned long long *)pu[2];
    h->peer_dn[0] = (int8_t *)pd[0];
    h->peer_dn[1] = (int8_t *)pd[1];
    h->peer_dn_flags = (unsigned long long *)pd[2];
    h->peer_xbuf_up = (PBPacket *)pu[3];
    h->peer_xbuf_dn = (PBPacket *)pd[3];
    h->slab_pbatch = xb;  // NOTE: every slab of the lattice must come to the same answer (same shape: they do)
    return PYMC_OK;
}

// Neighbours that live in the same process (several slabs driven by one host thread, possibly on one GPU):
// plain pointers instead of IPC mappings, and no flag barrier -- the caller orders the work itself by
// putting all slabs on one stream and issuing one sweep per slab in turn.
int pymc_slab_connect_local(pymc_handle *h, pymc_handle *up, pymc_handle *down) {
    ENTER(h);
    REQUIRE(h->slab && up && down && up->slab && down->slab, PYMC_ERR_BADARG, "pymc_slab_connect_local: slab handles only");
    REQUIRE(up->p.lenX == h->p.lenX && down->p.lenX == h->p.lenX && up->nk == h->nk && down->nk == h->nk,
            PYMC_ERR_BADARG, "pymc_slab_connect_local: slabs of different shape");
    h->peer_up[0] = up->buf[0];
    h->peer_up[1] = up->buf[1];
    h->peer_dn[0] = down->buf[0];
    h->peer_dn[1] = down->buf[1];
    h->peer_up_flags = h->peer_dn_flags = nullptr;
    h->local_peers = true;
    return PYMC_OK;
}

// My first / last two owned rows into the neighbours' halo rows, then the neighbour barrier (collective over the neighbours).
static int slab_refresh_halos(pymc_handle *h) {
    Lat L = make_lat(h);
    const int k = h->i8 == h->buf[0] ? 0 : 1;
    dim3 grid((h->nk / 16 + 127) / 128, 4, 2 * h->p.nreplicas);
    slab_push_halos_kernel<<<grid, 128, 0, h->stream>>>(L, h->i8, h->peer_up[k], h->peer_dn[k]);
    CUDA_TRY(cudaGetLastError());
    int rc = slab_sync_async(h);
    if (rc) return rc;
    h->dirty = true;
    return PYMC_OK;
}

// After pymc_upload_config on every slab: make the halo rows consistent (collective over the neighbours).
int pymc_slab_push_halos(pymc_handle *h) {
    ENTER(h);
    REQUIRE(h->slab, PYMC_ERR_STATE, "pymc_slab_push_halos: not a slab handle");
    REQUIRE(h->nslabs == 1 || h->peer_up[0], PYMC_ERR_STATE, "pymc_slab_push_halos: call pymc_slab_connect first");
    return slab_refresh_halos(h);
}

// 0 = fine; 1 = a neighbour did not reach a barrier within the time limit (results are invalid)
int pymc_slab_status(pymc_handle *h, int *status) {
    ENTER(h);
    REQUIRE(h->slab && status, PYMC_ERR_BADARG, "pymc_slab_status: bad arguments");
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    *status = h->err_host ? h->err_host[1] : 0;  // reported here without being cleared
    return PYMC_OK;
}

// Test hook for the error path of the in-kernel barriers: raises the same flag a timed-out barrier raises
// (which = 0 persistent kernel, 1 slab neighbour barrier) from a kernel on the handle's stream.
__global__ void raise_flag_kernel(int *flag) { *flag = 1; }

int pymc_debug_raise_async_error(pymc_handle *h, int which) {
    ENTER(h);
    REQUIRE(which == 0 || which == 1, PYMC_ERR_BADARG, "pymc_debug_raise_async_error: which = %d", which);
    REQUIRE(h->persist_err, PYMC_ERR_STATE, "pymc_debug_raise_async_error: no error flags");
    raise_flag_kernel<<<1, 1, 0, h->stream>>>(which ? h->slab_error : h->persist_err);
    CUDA_TRY(cudaGetLastError());
    return PYMC_OK;
}

int pymc_host_alloc(size_t bytes, void **out) {
    REQUIRE(out, PYMC_ERR_BADARG, "pymc_host_alloc: null pointer");
    *out = nullptr;
    CUDA_TRY(cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault));
    return PYMC_OK;
}

int pymc_host_free(void *ptr) {
    if (ptr) CUDA_TRY(cudaFreeHost(ptr));
    return PYMC_OK;
}

int pymc_timer_start(pymc_handle *h) {
    ENTER(h);
    CUDA_TRY(cudaEventRecord(h->ev0, h->stream));
    return PYMC_OK;
}

int pymc_timer_stop(pymc_handle *h, float *elapsed_ms) {
    ENTER(h);
    REQUIRE(elapsed_ms, PYMC_ERR_BADARG, "pymc_timer_stop: null pointer");
    CUDA_TRY(cudaEventRecord(h->ev1, h->stream));
    CUDA_TRY(cudaEventSynchronize(h->ev1));
    CUDA_TRY(cudaEventElapsedTime(elapsed_ms, h->ev0, h->ev1));
    CHECK_ASYNC(h);
    return PYMC_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// Observables without any allocation.
//
// Who this is for: a host program that must not allocate once it is running.  Typical cases are a
// process whose device memory has been handed over in full to another allocator (a framework's
// caching pool, a fixed per-client memory budget), so that a later cudaMalloc may simply fail, and a
// latency-sensitive loop that cannot accept the implicit device-wide synchronisation a cudaMalloc
// or cudaFree brings with it.  pymc_reduce_observables itself keeps its accumulators in memory
// that pymc_create obtained with cudaMalloc and stages the result through a host array from new[];
// this variant needs neither.
//
// The accumulators of reduce_kernel live in the file-scope array below instead: 16 bytes per replica,
// PYMC_STATIC_MAX_REPLICAS replicas, 64 KiB that every GPU which loads this library holds until the
// process exits.  In host code the name g_static_obs is only a stand-in for the device object, not
// a device address, so the address is looked up with cudaGetSymbolAddress -- once per device, with
// that device current -- and cached.
// ------------------------------------------------------------------------------------------
#define PYMC_STATIC_MAX_REPLICAS 4096
#define PYMC_STATIC_MAX_DEVICES 64
static __device__ __align__(256) unsigned long long g_static_obs[2 * PYMC_STATIC_MAX_REPLICAS];
static unsigned long long *g_static_obs_addr[PYMC_STATIC_MAX_DEVICES];  // per device, null = not looked up yet

// Device address of g_static_obs on `device`, which must be the current device.
static int static_obs_address(int device, unsigned long long **out) {
    REQUIRE(device >= 0 && device < PYMC_STATIC_MAX_DEVICES, PYMC_ERR_UNSUPPORTED,
            "static scratch supports device ordinals below %d (got %d)", PYMC_STATIC_MAX_DEVICES, device);
    if (!g_static_obs_addr[device]) {
        unsigned long long *ptr = nullptr;
        CUDA_TRY(cudaGetSymbolAddress((void **)&ptr, g_static_obs));
        REQUIRE(ptr != nullptr, PYMC_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
        g_static_obs_addr[device] = ptr;
    }
    *out = g_static_obs_addr[device];
    return PYMC_OK;
}

extern "C" int pymc_reduce_observables_static(pymc_handle *h, int64_t *bond, int64_t *second) {
    ENTER(h);
    REQUIRE(bond && second, PYMC_ERR_BADARG, "pymc_reduce_observables_static: null output pointer");
    REQUIRE(h->p.nreplicas <= PYMC_STATIC_MAX_REPLICAS, PYMC_ERR_UNSUPPORTED,
            "pymc_reduce_observables_static: %d replicas exceed the static scratch (at most %d replicas per call); "
            "use pymc_reduce_observables",
            h->p.nreplicas, PYMC_STATIC_MAX_REPLICAS);
    unsigned long long *acc = nullptr;
    int rc = static_obs_address(h->p.device, &acc);
    if (rc) return rc;
    const size_t nrep = (size_t)h->p.nreplicas;
    // Row slabs: the reduction reads one halo row, which a neighbouring GPU may still be writing (same wait as
    // in pymc_reduce_observables; a one-thread kernel, no allocation).
    rc = slab_wait_async(h);
    if (rc) return rc;
    // The array still holds whatever the previous call (of any handle on this GPU) left in it, and the
    // kernel adds into it: the part used now has to start from zero.
    CUDA_TRY(cudaMemsetAsync(acc, 0, nrep * 16, h->stream));
    Lat L = make_lat(h);
    dim3 grid((2 * h->p.lenX * h->nw + 255) / 256, h->p.nreplicas);
#define CALL_REDUCE_STATIC(LAY, MOD) reduce_kernel<LAY, MOD><<<grid, 256, 0, h->stream>>>(L, acc)
    DISPATCH(h, CALL_REDUCE_STATIC);
#undef CALL_REDUCE_STATIC
    CUDA_TRY(cudaGetLastError());
    // same side effect as pymc_reduce_observables: the tracked values are re-based on the fresh sums
    CUDA_TRY(cudaMemcpyAsync(h->tracked, acc, nrep * 16, cudaMemcpyDeviceToDevice, h->stream));
    h->dirty = false;
    // pairs (bond, second) are interleaved on the device; copy each column straight into the caller's array
    CUDA_TRY(cudaMemcpy2DAsync(bond, 8, acc, 16, 8, nrep, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaMemcpy2DAsync(second, 8, acc + 1, 16, 8, nrep, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    CHECK_ASYNC(h);
    return PYMC_OK;
}
