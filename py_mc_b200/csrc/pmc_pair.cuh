// pmc_pair.cuh -- K6: nearest-neighbour particle-hole exchange (Kawasaki) on conflict-free bond classes.
//
// The reference's particle move (examples/latgas_move.py:56-83) picks one occupied and one empty site
// anywhere on the lattice by materialising both index lists per trial (np.where, O(N)) and uses a
// defective dE.  Here particles hop between *neighbouring* sites with the
// correct energy difference of latticegas.energy (latgas_move.py:17-25),
//     dE = J * (occupied neighbours of the destination other than the source
//               - occupied neighbours of the source other than the destination),
// accepted by the Metropolis rule of py_mc/mc.py:175-184.  The 2N bonds are split into 16 classes
// (oracle/pairmove.py): bonds of one class share no site and no neighbour, so a class is one launch.
//   class 0..7   horizontal (i, j)-(i, j+1), i % 2 == c >> 2, j % 4 == c & 3
//   class 8..15  vertical   (i, j)-(i+1, j), i % 4 == (c-8) >> 1, j % 2 == (c-8) & 1
// One thread handles the 64-lane word that holds the first sites of its bonds; the neighbour counts
// are bit-sliced, the comparison is the shared bit-serial decide_hard (stream domain 3,
// step = 16*sweep + class, word and lane of the first site).
#pragma once
#include "pmc_core.cuh"

namespace pmc {

constexpr int PAIR_NCLASS = 7;  // d = -3..3

struct PairPlan {
    uint64_t notmask;  // 0: hard moves lose neighbours (J < 0); ~0: hard moves gain neighbours (J > 0)
    uint64_t en[3];
    uint32_t ht[3];
    template <class Row>
    __device__ __forceinline__ void init(const Row row) {
        notmask = row[0] ? 0ull : ~0ull;
#pragma unroll
        for (int e = 0; e < 3; ++e) {
            ht[e] = row[1 + e];
            en[e] = row[4 + e] ? ~0ull : 0ull;
        }
    }
    // thr[d+3], d = -3..3.  Returns false if the table has dE > 0 on both sides or at d = 0.
    static __host__ bool build_row(const uint32_t *thr, uint32_t *row) {
        const bool neg = thr[0] != THR_ALWAYS || thr[1] != THR_ALWAYS || thr[2] != THR_ALWAYS;
        const bool pos = thr[4] != THR_ALWAYS || thr[5] != THR_ALWAYS || thr[6] != THR_ALWAYS;
        if ((neg && pos) || thr[3] != THR_ALWAYS) return false;
        row[0] = neg ? 1u : 0u;
        for (int e = 1; e <= 3; ++e) {
            row[e] = neg ? thr[3 - e] : thr[3 + e];
            row[3 + e] = row[e] != THR_ALWAYS;
        }
        for (int k = 7; k < PLAN_WORDS; ++k) row[k] = 0;
        return true;
    }
};

// bit-sliced count of three one-bit operands
__device__ __forceinline__ void add3(uint64_t x, uint64_t y, uint64_t z, uint64_t &s0, uint64_t &s1) {
    const uint64_t t = x ^ y;
    s0 = t ^ z;
    s1 = (x & y) | (z & t);
}

// Decide all bonds whose first site is in this word.  S: first sites, Bw: second sites aligned lane by
// lane, (A0,A1)/(B0,B1): occupied neighbours of the first/second site other than the partner.
// Returns the mask of accepted exchanges; dB = change of the bond count.
__device__ __forceinline__ uint64_t pair_update(uint64_t S, uint64_t Bw, uint64_t A0, uint64_t A1, uint64_t B0,
                                                uint64_t B1, uint64_t active, const PairPlan &plan, uint32_t word_idx,
                                                uint32_t replica, uint64_t step, const PhiloxKey &key, int &dB) {
    const uint64_t movable = (S ^ Bw) & active;
    // source = where the particle sits.  P - Q is the quantity whose positive values are the hard moves.
    const uint64_t sel = S ^ plan.notmask;  // neg orientation: P = source count, Q = destination count
    const uint64_t P0 = (sel & A0) | (~sel & B0), P1 = (sel & A1) | (~sel & B1);
    const uint64_t Q0 = (sel & B0) | (~sel & A0), Q1 = (sel & B1) | (~sel & A1);
    const uint64_t p1 = ~P1 & P0, p2 = P1 & ~P0, p3 = P1 & P0;
    const uint64_t q0 = ~Q1 & ~Q0, q1 = ~Q1 & Q0, q2 = Q1 & ~Q0;
    uint64_t hm[3];
    hm[0] = ((p1 & q0) | (p2 & q1) | (p3 & q2)) & movable & plan.en[0];
    hm[1] = ((p2 & q0) | (p3 & q1)) & movable & plan.en[1];
    hm[2] = (p3 & q0) & movable & plan.en[2];
    const uint64_t acc = decide_hard<3>(hm, plan.ht, ~(hm[0] | hm[1] | hm[2]), movable, word_idx, replica, step,
                                        DOMAIN_PAIR, key);
    // destination minus source, summed over accepted lanes
    const uint64_t D0 = (S & B0) | (~S & A0), D1 = (S & B1) | (~S & A1);
    const uint64_t R0 = (S & A0) | (~S & B0), R1 = (S & A1) | (~S & B1);
    dB = __popcll(acc & D0) + 2 * __popcll(acc & D1) - __popcll(acc & R0) - 2 * __popcll(acc & R1);
    return acc;
}

}  // namespace pmc
