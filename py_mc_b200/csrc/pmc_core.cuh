// pmc_core.cuh -- the Metropolis test on one 64-site word, bit-sliced.
//
// What it replaces: model.trialstep + CanonicalMonteCarlo.MCstep for 64 same-colour sites at
// once (examples/ising2D.py:28-52, examples/latgas.py:31-57, py_mc/mc.py:170-185).
//
// A word holds 64 sites of one checkerboard colour as bits (1 = spin +1 / occupied).  The four
// neighbour words U, D (rows i-1, i+1), C (same row, same k) and Sd (same row, k-1 or k+1) belong
// to the other colour and are read-only during the pass.  Each lane falls in one *class* that
// fixes its dE (Ising: number of aligned neighbours 0..4; lattice gas: 5*c + occupied
// neighbours), the caller supplies one 32-bit threshold per class
// (thr = floor(exp(-dE/kT) 2^32), 2^32-1 = always accept), and a lane is accepted iff its
// uniform x32 <= thr.  x32 is defined bit-serially by the Philox planes of pmc_stream.cuh, so the
// comparison runs MSB-first on all 64 lanes in parallel for the first eight bits; the few lanes that
// are still undecided then are settled one per thread from their own Philox block.
#pragma once
#include "pmc_stream.cuh"

namespace pmc {

constexpr uint32_t THR_ALWAYS = 0xFFFFFFFFu;

enum : int { MODEL_ISING = 0, MODEL_LATGAS = 1 };

template <int MODEL>
struct ModelTraits;
template <>
struct ModelTraits<MODEL_ISING> {
    static constexpr int NCLASS = 5;
};
template <>
struct ModelTraits<MODEL_LATGAS> {
    static constexpr int NCLASS = 10;
};

// bit-sliced sum of four one-bit operands: count = a0 + 2 a1 + 4 a2
__device__ __forceinline__ void add4(uint64_t x1, uint64_t x2, uint64_t x3, uint64_t x4, uint64_t &a0, uint64_t &a1,
                                     uint64_t &a2) {
    uint64_t s01 = x1 ^ x2, c01 = x1 & x2;
    uint64_t s23 = x3 ^ x4, c23 = x3 & x4;
    a0 = s01 ^ s23;
    uint64_t c0 = s01 & s23;
    a1 = c01 ^ c23 ^ c0;
    a2 = c01 & c23;
}

// One-hot class masks from a bit-sliced count 0..4.
__device__ __forceinline__ void count_masks(uint64_t a0, uint64_t a1, uint64_t a2, uint64_t (&m)[5]) {
    m[0] = ~(a0 | a1 | a2);
    m[1] = a0 & ~a1;
    m[2] = a1 & ~a0;
    m[3] = a1 & a0;
    m[4] = a2;
}

// x32 <= thr for all lanes.  Every thread of the warp must call this together.
//
// Only the lanes of *hard* classes (thr < 2^32-1, i.e. dE > 0) need random bits; the caller
// passes them compacted as hm[k] (one-hot lane masks) / ht[k] (thresholds, warp-uniform).
// Head: the 8 most significant bits are compared bit-serially for all 64 lanes at once -- per plane
// the per-lane threshold bit is T = OR_k (hm[k] & -bit_j(ht[k])), 2 logic operations per hard class plus 3
// for the lt/und update on each 32-bit half -- which leaves about 1 lane in 256 undecided.
// Tail: each undecided lane is settled by one Philox block of its own (its low bits), by the thread
// that owns the word.  (Queueing the warp's undecided lanes in shared memory and sharing them out one
// per thread was tried and measured slower: the bookkeeping cost more than the blocks it saved.)
template <int MAXH>
__device__ __forceinline__ uint64_t decide_hard(const uint64_t (&hm)[MAXH], const uint32_t (&ht)[MAXH],
                                                uint64_t always, uint64_t valid, uint32_t word_idx,
                                                uint32_t replica, uint64_t step, uint32_t domain, const PhiloxKey &key) {
    uint64_t needed = 0;
#pragma unroll
    for (int k = 0; k < MAXH; ++k) needed |= hm[k];
    uint64_t und = needed & valid;  // lanes whose x32 still equals thr on all bits seen so far
    uint64_t lt = 0;                // lanes already known to have x32 < thr
#ifdef PMC_HEAD_ROLLED
#pragma unroll 1
#else
#pragma unroll
#endif
    for (uint32_t call = 0; call < HEAD_PLANES / 2; ++call) {
#ifdef PMC_HEAD_CHECK_EVERY_CALL
        if (!__any_sync(0xFFFFFFFFu, und != 0)) break;
#else
        // one check before the first block (a warp whose lanes all accept unconditionally draws nothing);
        // after that some lane of 2048 is practically always still open, so the test would be wasted work
        if (call == 0 && !__any_sync(0xFFFFFFFFu, und != 0)) break;
#endif
        uint64_t pe, po;
        word_planes(word_idx, replica, step, domain, call, key, pe, po);
        const int be = 31 - 2 * (int)call, bo = be - 1;
        // per-lane threshold bit of this plane: T = OR_k (hm[k] if bit_j(ht[k])).  The class masks are
        // disjoint, so the OR is a sum and the selection a multiplication by 0 or 1: integer
        // multiply-adds run on the FMA pipe, which is otherwise idle here, instead of the logic pipe.
        uint32_t te_lo = 0, te_hi = 0, to_lo = 0, to_hi = 0;
#pragma unroll
        for (int k = 0; k < MAXH; ++k) {
            const uint32_t b_e = (ht[k] >> be) & 1u, b_o = (ht[k] >> bo) & 1u;
            te_lo += (uint32_t)hm[k] * b_e;
            te_hi += (uint32_t)(hm[k] >> 32) * b_e;
            to_lo += (uint32_t)hm[k] * b_o;
            to_hi += (uint32_t)(hm[k] >> 32) * b_o;
        }
        const uint64_t te = ((uint64_t)te_hi << 32) | te_lo, to = ((uint64_t)to_hi << 32) | to_lo;
        lt |= und & te & ~pe;
        und &= ~(te ^ pe);
        lt |= und & to & ~po;
        und &= ~(to ^ po);
    }
    // Tail: every thread settles its own undecided lanes, one Philox block per lane; the warp loops
    // until the thread with the most of them is done (usually one or two rounds).
#pragma unroll 1
    while (__any_sync(0xFFFFFFFFu, und != 0)) {
        if (und) {
            const int l = __ffsll((long long)und) - 1;
            und &= und - 1;
            uint32_t thr = ht[0];
#pragma unroll
            for (int k = 1; k < MAXH; ++k)
                if ((hm[k] >> l) & 1ull) thr = ht[k];
            const uint32_t x = lane_tail(word_idx, replica, step, domain, (uint32_t)l, key);
            if (x <= (thr & ((1u << TAIL_BITS) - 1u))) lt |= 1ull << l;  // head bits equal: the low bits decide
        }
    }
    return (always & valid) | lt;
}

// ------------------------------------------------------------------------------------------
// Per-temperature plans: everything that depends only on the (warp-uniform) threshold row is
// worked out once per thread, outside the per-word path.
// ------------------------------------------------------------------------------------------
constexpr int PLAN_WORDS = 16;  // 32-bit words per temperature in the device plan table

// Where a thread reads its plan row from: the global table (one row per temperature, indexed through
// replica_to_T), or -- when the handle has a single temperature -- a copy of the row carried in the kernel
// parameters, i.e. in constant memory, which costs no load instruction and is known to be warp-uniform.
struct GlobalRow {
    const uint32_t *p;
    __device__ __forceinline__ uint32_t operator[](int k) const { return __ldg(p + k); }
};
struct ParamRow {
    const uint32_t (&a)[PLAN_WORDS];
    __device__ __forceinline__ uint32_t operator[](int k) const { return a[k]; }
};

template <int MODEL>
struct Plan;

// Ising.  dE = -2J(2a-4) with a = aligned neighbours, so dE > 0 for a in {3,4} when J < 0 and for
// a in {0,1} when J > 0 (pymc_set_accept_tables rejects anything else).  Counting *anti*-aligned
// neighbours in the second case maps it onto the first: the hard classes are always count 3 and 4.
// Row layout (built on the host by build_row): 0 count-aligned flag, 1-2 thresholds of count 3 / 4,
// 3-4 enable flags, 5 orientation (1 = flipped).
template <>
struct Plan<MODEL_ISING> {
    uint64_t notmask;  // ~0: count aligned neighbours; 0: count anti-aligned neighbours
    uint64_t en3, en4;  // ~0 where the class needs a draw
    uint32_t ht[2];
    int sign;           // +1 / -1: orientation of the bond-sum change
    template <class Row>
    __device__ __forceinline__ void init(const Row row) {
        notmask = row[0] ? ~0ull : 0ull;
        ht[0] = row[1];
        ht[1] = row[2];
        en3 = row[3] ? ~0ull : 0ull;
        en4 = row[4] ? ~0ull : 0ull;
        sign = row[5] ? -1 : 1;
    }
    static __host__ void build_row(const uint32_t *thr, uint32_t *row) {
        const bool low = thr[0] != THR_ALWAYS || thr[1] != THR_ALWAYS;
        const uint32_t tA = low ? thr[1] : thr[3], tB = low ? thr[0] : thr[4];
        row[0] = low ? 0u : 1u;
        row[1] = tA;
        row[2] = tB;
        row[3] = tA != THR_ALWAYS;
        row[4] = tB != THR_ALWAYS;
        row[5] = low ? 1u : 0u;
        for (int k = 6; k < PLAN_WORDS; ++k) row[k] = 0;
    }
};

// Lattice gas.  Class = 5c + nn; dE(c=0,nn) = -dE(c=1,nn), so for each nn at most one of the two
// occupancies needs a draw.  Row layout: 0-4 "hard class is the empty site" flags, 5-9 thresholds,
// 10-14 enable flags.
template <>
struct Plan<MODEL_LATGAS> {
    uint64_t sel[5];  // XOR-ed with S: result 1 = lane is in the hard occupancy for this nn
    uint64_t en[5];   // ~0 if that nn has a hard class at all
    uint32_t ht[5];
    template <class Row>
    __device__ __forceinline__ void init(const Row row) {
#pragma unroll
        for (int n = 0; n < 5; ++n) {
            sel[n] = row[n] ? ~0ull : 0ull;
            ht[n] = row[5 + n];
            en[n] = row[10 + n] ? ~0ull : 0ull;
        }
    }
    static __host__ void build_row(const uint32_t *thr, uint32_t *row) {
        for (int n = 0; n < 5; ++n) {
            const uint32_t te = thr[n], to = thr[5 + n];  // empty / occupied site
            const bool hard_occ = to != THR_ALWAYS;
            row[n] = hard_occ ? 0u : 1u;
            row[5 + n] = hard_occ ? to : te;
            row[10 + n] = row[5 + n] != THR_ALWAYS;
        }
        row[15] = 0;
    }
};

// Full update of one word.  Returns the new word; dB / dS are the exact integer changes of the
// bond sum and of the second observable (magnetisation or particle number) caused by this word.
template <int MODEL>
__device__ __forceinline__ uint64_t update_word(uint64_t S, uint64_t U, uint64_t D, uint64_t C, uint64_t Sd,
                                                uint64_t valid, const Plan<MODEL> &plan, uint32_t word_idx,
                                                uint32_t replica, uint64_t step, const PhiloxKey &key, int &dB, int &dS) {
    if constexpr (MODEL == MODEL_ISING) {
        uint64_t c0, c1, c2;
        const uint64_t nm = plan.notmask;
        add4(S ^ U ^ nm, S ^ D ^ nm, S ^ C ^ nm, S ^ Sd ^ nm, c0, c1, c2);
        uint64_t hm[2] = {c1 & c0 & plan.en3, c2 & plan.en4};
        const uint64_t acc = decide_hard<2>(hm, plan.ht, ~(hm[0] | hm[1]), valid, word_idx, replica, step,
                                            DOMAIN_SWEEP, key);
        // flipping a spin with a aligned neighbours changes the bond sum by 8 - 4a, M by -2s
        const int na = __popcll(acc);
        dB = plan.sign * (8 * na - 4 * (__popcll(acc & c0) + 2 * __popcll(acc & c1) + 4 * __popcll(acc & c2)));
        dS = -2 * (2 * __popcll(acc & S) - na);
        return S ^ acc;
    } else {
        uint64_t n0, n1, n2, cnt[5], hm[5], needed = 0;
        add4(U, D, C, Sd, n0, n1, n2);  // occupied neighbours
        count_masks(n0, n1, n2, cnt);
#pragma unroll
        for (int n = 0; n < 5; ++n) {
            hm[n] = cnt[n] & (S ^ plan.sel[n]) & plan.en[n];
            needed |= hm[n];
        }
        const uint64_t acc = decide_hard<5>(hm, plan.ht, ~needed, valid, word_idx, replica, step, DOMAIN_SWEEP, key);
        // toggling c -> 1-c with nn occupied neighbours changes the bond count by (1-2c) nn, N by (1-2c)
        const uint64_t add = acc & ~S, rem = acc & S;
        dB = (__popcll(add & n0) + 2 * __popcll(add & n1) + 4 * __popcll(add & n2)) -
             (__popcll(rem & n0) + 2 * __popcll(rem & n1) + 4 * __popcll(rem & n2));
        dS = __popcll(add) - __popcll(rem);
        return S ^ acc;
    }
}

}  // namespace pmc
