// pmc_stream.cuh -- Philox4x32-10 in registers and the "pmc stream v1" counter layout.
//
// The reference draws from CPython's MT19937 (py_mc/mc.py:2,182; examples/ising2D.py:32-33);
// a sweep kernel needs a counter-based generator instead.  The layout below is the single
// definition shared by every kernel; oracle/philox.py restates it for the CPU checker.
//
//   key     = (seed_lo, seed_hi)
//   counter = (idx, replica, step_lo, domain<<28 | call<<24 | lane<<18 | step_hi[17:0])
//
// Domain 0 (sweeps): idx = word index (row*nwords_per_row + wcol) of a 64-site word of one
// colour, step = 2*sweep + colour.  The 32-bit uniform x32 of lane l of that word has two parts:
//   head  its HEAD_PLANES = 8 most significant bits are bit l of the bit planes 0..7; call m
//         (lane field 0) yields planes 2m (o0 | o1<<32) and 2m+1 (o2 | o3<<32), plane 0 = MSB;
//   tail  its 24 low bits are o0 >> 8 of the lane's own call (call = TAIL_CALL = 15, lane field l).
// The head serves all 64 lanes with one Philox block per two bits and decides 255/256 of the
// comparisons; the few lanes still undecided take their remaining bits from one block each.
#pragma once
#include <stdint.h>

namespace pmc {

constexpr uint32_t PHILOX_M0 = 0xD2511F53u;
constexpr uint32_t PHILOX_M1 = 0xCD9E8D57u;
constexpr uint32_t PHILOX_W0 = 0x9E3779B9u;
constexpr uint32_t PHILOX_W1 = 0xBB67AE85u;

enum : uint32_t { DOMAIN_SWEEP = 0, DOMAIN_INIT = 1, DOMAIN_RX = 2, DOMAIN_PAIR = 3, DOMAIN_DISC = 4 };

// The ten round keys of Philox4x32-10, (k0 + r W0, k1 + r W1), worked out once on the host.  A kernel receives them
// in its parameters, i.e. in constant memory, so a round's key is an operand of the XOR itself and the key schedule
// costs no instruction at all (20 uniform additions per block before: 13 % of all issued instructions of a sweep).
struct PhiloxKey {
    uint32_t rk[10][2];
};

__host__ __device__ __forceinline__ PhiloxKey make_key(uint64_t seed) {
    PhiloxKey k;
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    for (int r = 0; r < 10; ++r) {
        k.rk[r][0] = k0;
        k.rk[r][1] = k1;
        k0 += PHILOX_W0;
        k1 += PHILOX_W1;
    }
    return k;
}

#ifndef PMC_HEAD_PLANES
#define PMC_HEAD_PLANES 8
#endif
constexpr int HEAD_PLANES = PMC_HEAD_PLANES;  // planes compared bit-serially for all lanes of a word (even)
constexpr uint32_t TAIL_CALL = 15;   // call number of a lane's own block
constexpr int TAIL_BITS = 32 - HEAD_PLANES;

__host__ __device__ __forceinline__ uint32_t ctr3(uint32_t domain, uint32_t call, uint64_t step, uint32_t lane = 0) {
    return (domain << 28) | (call << 24) | (lane << 18) | ((uint32_t)(step >> 32) & 0x3FFFFu);
}

// One Philox4x32-10 block with the round keys of make_key.
__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, const PhiloxKey &key) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(PHILOX_M0, c0), lo0 = PHILOX_M0 * c0;
        uint32_t hi1 = __umulhi(PHILOX_M1, c2), lo1 = PHILOX_M1 * c2;
        uint32_t n0 = hi1 ^ c1 ^ key.rk[r][0];
        uint32_t n2 = hi0 ^ c3 ^ key.rk[r][1];
        c0 = n0;
        c1 = lo1;
        c2 = n2;
        c3 = lo0;
    }
    return make_uint4(c0, c1, c2, c3);
}

// Two bit planes (2*call, 2*call+1) of one word.
__device__ __forceinline__ void word_planes(uint32_t word_idx, uint32_t replica, uint64_t step, uint32_t domain,
                                            uint32_t call, const PhiloxKey &key, uint64_t &even, uint64_t &odd) {
    uint4 o = philox4x32_10(word_idx, replica, (uint32_t)step, ctr3(domain, call, step), key);
    even = (uint64_t)o.x | ((uint64_t)o.y << 32);
    odd = (uint64_t)o.z | ((uint64_t)o.w << 32);
}

// The 26 low bits of lane `lane` of a word.
__device__ __forceinline__ uint32_t lane_tail(uint32_t word_idx, uint32_t replica, uint64_t step, uint32_t domain,
                                              uint32_t lane, const PhiloxKey &key) {
    return philox4x32_10(word_idx, replica, (uint32_t)step, ctr3(domain, TAIL_CALL, step, lane), key).x >> HEAD_PLANES;
}

}  // namespace pmc
