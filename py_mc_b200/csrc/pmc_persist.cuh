// pmc_persist.cuh -- K11: ONE lattice resident in the shared memory of the WHOLE GPU for all sweeps of a call.
//
// The reference keeps its one lattice in host memory and touches one site per MCstep
// (py_mc/mc.py:170-185).  The 16384 x 16384 headline lattice is 32 MiB as bits; a B200 has 148 SMs x 227 KiB
// = 32.8 MiB of shared memory.  So the lattice is cut into one row slab per SM (110 or 111 rows of 128 words
// of both colours = 222 KiB), every block of a cooperative launch loads its slab once (int8 -> bits, or
// bits as stored), runs ALL sweeps of the call on it, and stores it once.  Per colour pass a block needs one
// row of the colour just updated from each neighbour slab: the blocks publish their first and last row in
// a small global buffer (2 x 1 KiB per block and colour), meet at a grid barrier, and read the neighbours'
// rows.  HBM sees one read and one write of the lattice per call; the byte <-> bit packing of the int8
// layout, a quarter of the fused kernel's instructions, is paid once per call instead of once per sweep.
//
// Same random-stream indexing (word = global row * nw + w), so the trajectory is that of every other
// sweep kernel (tests/test_zy_gpu_persist.py; the per-block logic also runs on the host in
// tests/test_hostemu_core.py, block after block, which checks the slab geometry and the halo plumbing).
//
// The per-thread pieces are plain functions of (thread index t, thread count T) so that they can be driven
// from the host shim as well; the __global__ wrapper adds the barriers.
#pragma once
#include "pmc_core.cuh"
#ifdef __CUDACC__
#include "pmc_tiled_i8.cuh"  // pack16 / unpack16
#endif

namespace pmc {

struct PersistArgs {
    long long nsweeps;
    unsigned long long sweep0;   // the handle's sweep counter at entry
    int rep;                     // replica (of this handle) the launch works on
    unsigned int *counter;       // grid barrier: zeroed by the host before the launch, counts arrivals
    int *err;                    // set if a block waited too long at the barrier
    uint64_t *halo;              // [2 colours][nblocks][2 sides][nw]
};

// first row of slab b when lenX rows are cut into nblocks slabs (sizes differ by at most one)
__host__ __device__ __forceinline__ int persist_row0(int b, int lenX, int nblocks) {
    return (int)(((long long)b * lenX) / nblocks);
}

// Shared-memory image of a slab: [2 colours][nrows + 2][nw]; local row -1 / nrows are the neighbours' rows.
__device__ __forceinline__ uint64_t &persist_word(uint64_t *s, int nrows, int nw, int colour, int lr, int w) {
    return s[((size_t)colour * (nrows + 2) + (lr + 1)) * nw + w];
}

// One colour pass over the slab.  Every thread of a warp must call this together (update_word votes).
template <int MODEL>
__device__ __forceinline__ void persist_pass(uint64_t *s, int nrows, int nw, int r0, int colour, const Plan<MODEL> &plan,
                                             uint32_t greplica, uint64_t step, const PhiloxKey &key, int t, int T, long long &dB,
                                             long long &dS) {
    const int oc = colour ^ 1, items = nrows * nw;
#pragma unroll 1
    for (int base = 0; base < items; base += T) {
        const bool on = base + t < items;
        const int item = on ? base + t : 0;
        const int lr = item / nw, w = item - lr * nw;
        const int i = r0 + lr;  // row of the whole lattice
        const uint64_t S = persist_word(s, nrows, nw, colour, lr, w);
        const uint64_t C = persist_word(s, nrows, nw, oc, lr, w);
        uint64_t Sd;
        if ((i + colour) & 1)
            Sd = (C >> 1) | (persist_word(s, nrows, nw, oc, lr, w == nw - 1 ? 0 : w + 1) << 63);
        else
            Sd = (C << 1) | (persist_word(s, nrows, nw, oc, lr, w == 0 ? nw - 1 : w - 1) >> 63);
        int b, s2;
        const uint64_t Snew =
            update_word<MODEL>(S, persist_word(s, nrows, nw, oc, lr - 1, w), persist_word(s, nrows, nw, oc, lr + 1, w), C, Sd,
                               on ? ~0ull : 0ull, plan, (uint32_t)(i * nw + w), greplica, step, key, b, s2);
        if (on) {
            persist_word(s, nrows, nw, colour, lr, w) = Snew;
            dB += b;
            dS += s2;
        }
    }
}

// After a pass of `colour`: my first and last row of that colour go to the exchange buffer ...
__device__ __forceinline__ void persist_publish(const uint64_t *s, int nrows, int nw, int colour, int b, int nblocks,
                                                uint64_t *halo, int t, int T) {
    uint64_t *mine = halo + ((size_t)colour * nblocks + b) * 2 * nw;
    for (int k = t; k < 2 * nw; k += T) {
        const int side = k / nw, w = k - side * nw;
        mine[k] = s[((size_t)colour * (nrows + 2) + (side ? nrows : 1)) * nw + w];
    }
}

// ... and, once every block has published, the neighbours' rows come into my halo rows (L1 is bypassed: the same
// addresses held older rows two passes ago).
__device__ __forceinline__ uint64_t persist_ld(const uint64_t *p) {
#ifdef __CUDA_ARCH__
    return __ldcg(p);
#else
    return *p;
#endif
}
__device__ __forceinline__ void persist_collect(uint64_t *s, int nrows, int nw, int colour, int b, int nblocks,
                                                const uint64_t *halo, int t, int T) {
    const int up = b == 0 ? nblocks - 1 : b - 1, dn = b == nblocks - 1 ? 0 : b + 1;
    const uint64_t *above = halo + (((size_t)colour * nblocks + up) * 2 + 1) * nw;  // its last row
    const uint64_t *below = halo + (((size_t)colour * nblocks + dn) * 2 + 0) * nw;  // its first row
    for (int w = t; w < nw; w += T) {
        persist_word(s, nrows, nw, colour, -1, w) = persist_ld(above + w);
        persist_word(s, nrows, nw, colour, nrows, w) = persist_ld(below + w);
    }
}

#ifdef __CUDACC__

constexpr int PERSIST_MAX_THREADS = 1024;

// All blocks of the (cooperative, hence co-resident) launch meet here.  `target` = arrivals expected so far.
// The wait is bounded: a time-out raises *err instead of hanging the GPU.
__device__ __forceinline__ void persist_grid_barrier(unsigned int *counter, unsigned int target, int *err) {
    __threadfence();  // my halo rows are visible before my arrival is
    __syncthreads();
    if (threadIdx.x == 0) {
        atomicAdd(counter, 1u);
        const long long t0 = clock64();
        while (*(volatile unsigned int *)counter < target) {
            if (clock64() - t0 > 4000000000LL) {
                *err = 1;
                break;
            }
        }
        __threadfence();
    }
    __syncthreads();
}

// grid = nblocks (<= number of SMs, cooperative launch), block = multiple of 32 threads,
// dynamic shared memory = 2 * (max slab rows + 2) * nw * 8 bytes.
template <int MODEL, int LAYOUT, class LatT>
__global__ void __launch_bounds__(PERSIST_MAX_THREADS, 1) persistent_kernel(LatT L, PersistArgs a) {
    extern __shared__ __align__(16) uint64_t s_slab[];
    __shared__ long long s_red[2][PERSIST_MAX_THREADS / 32];
    const int b = blockIdx.x, B = gridDim.x, t = threadIdx.x, T = blockDim.x;
    const int lenX = L.lenX, nw = L.nw, nk = L.nk, rep = a.rep;
    const int r0 = persist_row0(b, lenX, B), nrows = persist_row0(b + 1, lenX, B) - r0;
    const uint32_t greplica = (uint32_t)(L.rep_offset + rep);

    // ---- load my rows and the two neighbouring rows of both colours ---------------------------------
    for (int colour = 0; colour < 2; ++colour) {
        if (LAYOUT == PYMC_LAYOUT_BITS) {
            const uint64_t *src = L.bits + (size_t)(rep * 2 + colour) * lenX * nw;
            for (int idx = t; idx < (nrows + 2) * nw; idx += T) {
                const int lr = idx / nw - 1, w = idx - (lr + 1) * nw;
                int i = r0 + lr;
                i = i < 0 ? i + lenX : (i >= lenX ? i - lenX : i);
                persist_word(s_slab, nrows, nw, colour, lr, w) = src[(size_t)i * nw + w];
            }
        } else {
            const int8_t *src = L.i8 + (size_t)(rep * 2 + colour) * lenX * nk;
            uint16_t *s16 = reinterpret_cast<uint16_t *>(s_slab + (size_t)colour * (nrows + 2) * nw);
            for (int c = t; c < (nrows + 2) * nw * 4; c += T) {  // 16-byte chunks, 4 per word
                const int lr = c / (nw * 4) - 1, cc = c - (lr + 1) * nw * 4;
                int i = r0 + lr;
                i = i < 0 ? i + lenX : (i >= lenX ? i - lenX : i);
                s16[c] = (uint16_t)pack16<MODEL>(__ldg(reinterpret_cast<const uint4 *>(src + (size_t)i * nk) + cc));
            }
        }
    }
    __syncthreads();

    Plan<MODEL> plan;
    fetch_plan(L, rep, plan);
    long long dB = 0, dS = 0;
    unsigned int arrivals = 0;
    for (long long sw = 0; sw < a.nsweeps; ++sw) {
#pragma unroll 1
        for (int colour = 0; colour < 2; ++colour) {
            persist_pass<MODEL>(s_slab, nrows, nw, r0, colour, plan, greplica, 2 * (a.sweep0 + sw) + colour, L.key, t, T, dB, dS);
            if (sw == a.nsweeps - 1 && colour == 1) break;  // nobody reads the halo rows after the last pass
            __syncthreads();
            persist_publish(s_slab, nrows, nw, colour, b, B, a.halo, t, T);
            arrivals += (unsigned int)B;
            persist_grid_barrier(a.counter, arrivals, a.err);
            persist_collect(s_slab, nrows, nw, colour, b, B, a.halo, t, T);
            __syncthreads();
        }
    }
    __syncthreads();

    // ---- store my rows ----------------------------------------------------------------------------
    for (int colour = 0; colour < 2; ++colour) {
        if (LAYOUT == PYMC_LAYOUT_BITS) {
            uint64_t *dst = L.bits + ((size_t)(rep * 2 + colour) * lenX + r0) * nw;
            for (int idx = t; idx < nrows * nw; idx += T) dst[idx] = s_slab[((size_t)colour * (nrows + 2) + 1) * nw + idx];
        } else {
            uint4 *dst = reinterpret_cast<uint4 *>(L.i8 + ((size_t)(rep * 2 + colour) * lenX + r0) * nk);
            const uint16_t *s16 = reinterpret_cast<const uint16_t *>(s_slab + ((size_t)colour * (nrows + 2) + 1) * nw);
            for (int c = t; c < nrows * nw * 4; c += T) dst[c] = unpack16<MODEL>(s16[c]);
        }
    }

#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        dB += __shfl_down_sync(0xFFFFFFFFu, dB, d);
        dS += __shfl_down_sync(0xFFFFFFFFu, dS, d);
    }
    if ((t & 31) == 0) {
        s_red[0][t >> 5] = dB;
        s_red[1][t >> 5] = dS;
    }
    __syncthreads();
    if (t < 2) {
        long long v = 0;
        for (int k = 0; k < (T >> 5); ++k) v += s_red[t][k];
        if (v) atomicAdd(L.tracked + 2 * (size_t)rep + t, (unsigned long long)v);
    }
}

#endif  // __CUDACC__

}  // namespace pmc
