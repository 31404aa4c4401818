// pmc_pack.cuh -- 16 sites of the int8 layout <-> 16 bits of a word (shared by every int8 kernel; plain integer
// code, so the host emulation of the test-suite compiles it too).
#pragma once
#include "pmc_core.cuh"

namespace pmc {

// 16 bytes (+1/-1 or 0/1) -> 16 bits (1 = +1 / occupied).  Three operations per four sites:
// isolate one bit per byte, gather the four bits at the top of a product, shift into place.
template <int MODEL>
__device__ __forceinline__ uint32_t pack16(uint4 v) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t r = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        uint32_t nib;
        if (MODEL == MODEL_ISING) {
            // sign bit of each byte (set for -1): bits 7,15,23,31 -> product bits 28..31
            nib = ((w[j] & 0x80808080u) * 0x00204081u) >> 28;
        } else {
            // bit 0 of each byte: bits 0,8,16,24 -> product bits 28..31
            nib = ((w[j] & 0x01010101u) * 0x10204080u) >> 28;
        }
        r += nib * (1u << (4 * j));  // disjoint nibbles: a multiply-add (FMA pipe) instead of shift + or
    }
    return (MODEL == MODEL_ISING) ? (~r & 0xFFFFu) : r;
}

// 16 bits -> 16 bytes
template <int MODEL>
__device__ __forceinline__ uint4 unpack16(uint32_t bits) {
    uint32_t w[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint32_t nib = (bits >> (4 * j)) & 0xFu;
        const uint32_t spread = (nib * 0x00204081u) & 0x01010101u;  // bit b -> byte b (0 or 1)
        // 1 -> 0x01, 0 -> 0xFF:  0xFF - 0xFE*bit per byte, one multiply-add
        w[j] = (MODEL == MODEL_ISING) ? 0xFFFFFFFFu - spread * 0xFEu : spread;
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}

}  // namespace pmc
