// pmc_layout.cuh -- natural row-major int8 image  <->  colour-split device storage, coalesced.
//
// Host format (what the reference keeps in config.config, examples/ising2D.py:56-70, as integers):
// int8 [lenX][lenY].  In row i the even columns belong to colour (i & 1) and the odd columns to the
// other colour, so a 16-byte chunk of a natural row de-interleaves into 8 bytes of each colour.
// One thread handles one 16-byte chunk; all global accesses are 8- or 16-byte wide and contiguous
// across a warp.  Needs lenY % 16 == 0 (int8 layout) or lenY % 128 == 0 (bit layout); other widths
// use the simple per-word kernels in pymc_b200.cu.
#pragma once
#include "pmc_tiled_i8.cuh"

namespace pmc {

__device__ __forceinline__ void deinterleave16(uint4 v, uint2 &even, uint2 &odd) {
    even.x = __byte_perm(v.x, v.y, 0x6420);
    even.y = __byte_perm(v.z, v.w, 0x6420);
    odd.x = __byte_perm(v.x, v.y, 0x7531);
    odd.y = __byte_perm(v.z, v.w, 0x7531);
}

__device__ __forceinline__ uint4 interleave16(uint2 even, uint2 odd) {
    uint4 v;
    v.x = __byte_perm(even.x, odd.x, 0x5140);
    v.y = __byte_perm(even.x, odd.x, 0x7362);
    v.z = __byte_perm(even.y, odd.y, 0x5140);
    v.w = __byte_perm(even.y, odd.y, 0x7362);
    return v;
}

// 8 bytes -> 8 bits and back (same multiply-gather as pack16 / unpack16)
template <int MODEL>
__device__ __forceinline__ uint32_t pack8(uint2 v) {
    uint32_t r;
    if (MODEL == MODEL_ISING) {
        r = (((v.x & 0x80808080u) * 0x00204081u) >> 28) | ((((v.y & 0x80808080u) * 0x00204081u) >> 28) << 4);
        r = ~r & 0xFFu;
    } else {
        r = (((v.x & 0x01010101u) * 0x10204080u) >> 28) | ((((v.y & 0x01010101u) * 0x10204080u) >> 28) << 4);
    }
    return r;
}

template <int MODEL>
__device__ __forceinline__ uint2 unpack8(uint32_t bits) {
    uint2 v;
    const uint32_t lo = ((bits & 0xFu) * 0x00204081u) & 0x01010101u;
    const uint32_t hi = (((bits >> 4) & 0xFu) * 0x00204081u) & 0x01010101u;
    v.x = (MODEL == MODEL_ISING) ? ~(lo * 0xFEu) : lo;
    v.y = (MODEL == MODEL_ISING) ? ~(hi * 0xFEu) : hi;
    return v;
}

// grid = ceil(lenX * lenY/16 / block)
template <int LAYOUT, int MODEL, class LatT>
__global__ void __launch_bounds__(256) pack_fast_kernel(LatT L, int rep, const int8_t *__restrict__ natural) {
    const int cpr = L.lenY >> 4;  // chunks per natural row
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = idx < (L.own_hi - L.own_lo) * cpr;
    const int irow = active ? idx / cpr : 0, c = active ? idx - irow * cpr : 0;
    const int i = L.own_lo + irow;  // device row; the host image holds the owned rows only
    uint2 even = make_uint2(0, 0), odd = make_uint2(0, 0);
    if (active) deinterleave16(__ldg(reinterpret_cast<const uint4 *>(natural + (size_t)irow * L.lenY) + c), even, odd);
    const int ce = i & 1, co = ce ^ 1;  // colour of the even / odd columns of row i
    if (LAYOUT == PYMC_LAYOUT_INT8) {
        if (active) {
            reinterpret_cast<uint2 *>(L.i8 + ((size_t)(rep * 2 + ce) * L.lenX + i) * L.nk)[c] = even;
            reinterpret_cast<uint2 *>(L.i8 + ((size_t)(rep * 2 + co) * L.lenX + i) * L.nk)[c] = odd;
        }
    } else {
        // eight consecutive lanes hold the eight bytes of one 64-site word (lenY % 128 == 0)
        const int sh = 8 * (threadIdx.x & 7);
        uint64_t we = (uint64_t)pack8<MODEL>(even) << sh, wo = (uint64_t)pack8<MODEL>(odd) << sh;
#pragma unroll
        for (int d = 1; d < 8; d <<= 1) {
            we |= __shfl_xor_sync(0xFFFFFFFFu, we, d);
            wo |= __shfl_xor_sync(0xFFFFFFFFu, wo, d);
        }
        if (active && (threadIdx.x & 7) == 0) {
            L.bits[((size_t)(rep * 2 + ce) * L.lenX + i) * L.nw + (c >> 3)] = we;
            L.bits[((size_t)(rep * 2 + co) * L.lenX + i) * L.nw + (c >> 3)] = wo;
        }
    }
}

template <int LAYOUT, int MODEL, class LatT>
__global__ void __launch_bounds__(256) unpack_fast_kernel(LatT L, int rep, int8_t *__restrict__ natural) {
    const int cpr = L.lenY >> 4;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (L.own_hi - L.own_lo) * cpr) return;
    const int irow = idx / cpr, c = idx - irow * cpr;
    const int i = L.own_lo + irow;
    const int ce = i & 1, co = ce ^ 1;
    uint2 even, odd;
    if (LAYOUT == PYMC_LAYOUT_INT8) {
        even = reinterpret_cast<const uint2 *>(L.i8 + ((size_t)(rep * 2 + ce) * L.lenX + i) * L.nk)[c];
        odd = reinterpret_cast<const uint2 *>(L.i8 + ((size_t)(rep * 2 + co) * L.lenX + i) * L.nk)[c];
    } else {
        const int sh = 8 * (c & 7);
        const uint64_t we = L.bits[((size_t)(rep * 2 + ce) * L.lenX + i) * L.nw + (c >> 3)];
        const uint64_t wo = L.bits[((size_t)(rep * 2 + co) * L.lenX + i) * L.nw + (c >> 3)];
        even = unpack8<MODEL>((uint32_t)(we >> sh) & 0xFFu);
        odd = unpack8<MODEL>((uint32_t)(wo >> sh) & 0xFFu);
    }
    reinterpret_cast<uint4 *>(natural + (size_t)irow * L.lenY)[c] = interleave16(even, odd);
}

}  // namespace pmc
