// cuda_host_shim.h -- just enough of the CUDA device vocabulary to compile the kernels' *core* headers
// (py_mc_b200/csrc/pmc_stream.cuh, pmc_core.cuh, pmc_pair.cuh) with g++ for the CPU test-suite.
//
// TEST INFRASTRUCTURE ONLY.  The product never includes this file and has no CPU path: it lets the
// `-m "not gpu"` tests run the very source the GPU executes -- the bit-sliced Metropolis test, the Philox
// planes, the plan rows -- against the reference-pinned fixtures, so that a change to the core is checked
// before any GPU time is spent.  A "warp" here is one thread: the warp votes reduce to their argument.
#pragma once
#include <stddef.h>
#include <stdint.h>

#define __device__
#define __host__
#define __global__
#define __forceinline__ inline
#define __restrict__

struct uint4 {
    uint32_t x, y, z, w;
};
static inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
    uint4 v = {x, y, z, w};
    return v;
}
static inline uint32_t __umulhi(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * (uint64_t)b) >> 32); }
static inline int __popcll(unsigned long long x) { return __builtin_popcountll(x); }
static inline int __ffsll(long long x) { return __builtin_ffsll(x); }
static inline int __any_sync(unsigned, int pred) { return pred; }
template <class T>
static inline T __ldg(const T *p) {
    return *p;
}
