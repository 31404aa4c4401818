// pmc_resident.cuh -- K10: whole-lattice-resident kernel for batches of small lattices.
//
// The reference runs one replica per MPI rank and touches its lattice one site at a time
// (py_mc/mc.py:170-185, py_mc/mc_mpi.py:47-56).  When a handle holds many replicas whose bit image fits
// into shared memory (BASELINE config 4: 512 lattice-gas replicas of 256 x 256 = 8 KiB each; config 3:
// 1024 x 1024 = 128 KiB), the per-pass kernels are launch-bound and under-fill the GPU: a colour pass of
// such a batch is tens of microseconds, a bond class of the pair-exchange sweep (pmc_pair.cuh) a few.
// Here ONE block owns ONE replica for a whole call: it loads the lattice once (int8 -> bits, or bits as
// they are), runs every requested step -- flip sweeps (both colours), pair-exchange passes (any subset
// of the 16 bond classes), or both in turn -- on the copy in shared memory with block barriers between
// the passes, and stores it once.  HBM sees one read and one write of the lattice per call, whatever
// the number of sweeps; the kernel is bound by instruction issue like the bit-packed kernel.
//
// The random numbers are indexed by (word, replica, step) exactly as in the per-pass kernels, so the
// trajectories are identical to theirs (tests/test_gpu_resident.py compares all of them).
//
// Needs lenY % 128 == 0 (whole 64-site words).  grid = nreplicas, block = any multiple of 32 threads,
// dynamic shared memory = 2 * lenX * nw * 8 bytes.
#pragma once
#include "pmc_pair.cuh"
#include "pmc_tiled_i8.cuh"

namespace pmc {

struct ResidentArgs {
    long long nsteps;               // outer steps
    int flip;                       // 1: every step starts with a flip sweep (colour 0, then colour 1)
    unsigned int mask_even, mask_odd;  // bond classes of the step's pair pass, chosen by the parity of the
                                       // pair counter; 0 = no pair pass
    unsigned long long sweep0, pair0;  // the handle's counters at entry
};

constexpr int RESIDENT_MAX_THREADS = 1024;

template <int MODEL, int LAYOUT, class LatT>
__global__ void __launch_bounds__(RESIDENT_MAX_THREADS, 1) resident_kernel(LatT L, ResidentArgs a) {
    extern __shared__ __align__(16) uint64_t s_lat[];  // [2][lenX][nw]
    __shared__ long long s_red[2][RESIDENT_MAX_THREADS / 32];
    const int rep = blockIdx.x, t = threadIdx.x, T = blockDim.x;
    const int lenX = L.lenX, nw = L.nw, plane = lenX * nw;
    const uint32_t greplica = (uint32_t)(L.rep_offset + rep);

    // ---- load ------------------------------------------------------------------------------------
    if (LAYOUT == PYMC_LAYOUT_BITS) {
        const uint64_t *src = L.bits + (size_t)rep * 2 * plane;
        for (int idx = t; idx < 2 * plane; idx += T) s_lat[idx] = src[idx];
    } else {
        const uint4 *src = reinterpret_cast<const uint4 *>(L.i8 + (size_t)rep * 2 * lenX * L.nk);
        uint16_t *s16 = reinterpret_cast<uint16_t *>(s_lat);
        for (int c = t; c < 8 * plane; c += T) s16[c] = (uint16_t)pack16<MODEL>(__ldg(src + c));
    }
    __syncthreads();

    auto W = [&](int colour, int i, int w) -> uint64_t & { return s_lat[(colour * lenX + i) * nw + w]; };
    // word whose lane l holds site k-1 / k+1 of word w (periodic in the row)
    auto prev_site = [&](int colour, int i, int w) {
        return (W(colour, i, w) << 1) | (W(colour, i, w == 0 ? nw - 1 : w - 1) >> 63);
    };
    auto next_site = [&](int colour, int i, int w) {
        return (W(colour, i, w) >> 1) | (W(colour, i, w == nw - 1 ? 0 : w + 1) << 63);
    };
    // xor into a word that another thread of the same pass may also xor into (different bits)
    auto xor_shared = [&](int colour, int i, int w, uint64_t m) {
        unsigned int *p = reinterpret_cast<unsigned int *>(&W(colour, i, w));
        if ((unsigned int)m) atomicXor(p, (unsigned int)m);
        if ((unsigned int)(m >> 32)) atomicXor(p + 1, (unsigned int)(m >> 32));
    };

    long long dB = 0, dS = 0;  // per-thread partial sums cancel only across threads: keep them wide
    unsigned long long sweep = a.sweep0, pair = a.pair0;
    for (long long s = 0; s < a.nsteps; ++s) {
        if (a.flip) {
            Plan<MODEL> plan;
            fetch_plan(L, rep, plan);
#pragma unroll 1
            for (int colour = 0; colour < 2; ++colour) {
                const int oc = colour ^ 1;
#pragma unroll 1
                for (int base = 0; base < plane; base += T) {  // whole warps stay in the loop: the update votes
                    const bool on = base + t < plane;
                    const int item = on ? base + t : 0;
                    const int i = item / nw, w = item - i * nw;
                    const int im = i == 0 ? lenX - 1 : i - 1, ip = i == lenX - 1 ? 0 : i + 1;
                    const uint64_t S = W(colour, i, w);
                    const uint64_t Sd = ((i + colour) & 1) ? next_site(oc, i, w) : prev_site(oc, i, w);
                    int b, s2;
                    const uint64_t Snew = update_word<MODEL>(S, W(oc, im, w), W(oc, ip, w), W(oc, i, w), Sd, on ? ~0ull : 0ull,
                                                             plan, (uint32_t)item, greplica, 2 * sweep + colour, L.key, b, s2);
                    if (on) {
                        W(colour, i, w) = Snew;
                        dB += b;
                        dS += s2;
                    }
                }
                __syncthreads();
            }
            ++sweep;
        }
        const unsigned int mask = (pair & 1ull) ? a.mask_odd : a.mask_even;
        if (mask) {
            PairPlan pplan;
            pplan.init(GlobalRow{L.pair_plans + (size_t)L.rep_to_T[L.rep_offset + rep] * PLAN_WORDS});
#pragma unroll 1
            for (int cls = 0; cls < 16; ++cls) {
                if (!((mask >> cls) & 1u)) continue;
                // bond classes as in pmc_pair.cuh: horizontal (i, j)-(i, j+1) with i % 2, j % 4 fixed;
                // vertical (i, j)-(i+1, j) with i % 4, j % 2 fixed
                const bool horiz = cls < 8;
                const int ri = horiz ? cls >> 2 : (cls - 8) >> 1, cj = horiz ? cls & 3 : (cls - 8) & 1;
                const int rstride = horiz ? 2 : 4;
                const int nitems = (lenX / rstride) * nw;
                const int q = cj & 1;  // column parity of the first site: where its colour starts in the row
#pragma unroll 1
                for (int base = 0; base < nitems; base += T) {  // whole warps stay in the loop (see above)
                    const bool on = base + t < nitems;
                    const int item = on ? base + t : 0;
                    const int ridx = item / nw, w = item - ridx * nw;
                    const int i = ri + rstride * ridx;
                    const int ca = (i + cj) & 1, oc = ca ^ 1;
                    const int im = i == 0 ? lenX - 1 : i - 1, ip = i == lenX - 1 ? 0 : i + 1;
                    const uint64_t S = W(ca, i, w);
                    uint64_t Bw, A0, A1, B0, B1, active;
                    int i1 = ip;
                    if (horiz) {
                        active = ((cj >> 1) & 1) ? 0xAAAAAAAAAAAAAAAAull : 0x5555555555555555ull;
                        uint64_t La, Ub, Db;
                        if (!q) {  // partner at the same k of the other colour; left neighbour of the first at k-1
                            Bw = W(oc, i, w);
                            La = prev_site(oc, i, w);
                            Ub = W(ca, im, w);
                            Db = W(ca, ip, w);
                        } else {  // partner at k+1; left neighbour of the first at k
                            Bw = next_site(oc, i, w);
                            La = W(oc, i, w);
                            Ub = next_site(ca, im, w);
                            Db = next_site(ca, ip, w);
                        }
                        add3(W(oc, im, w), W(oc, ip, w), La, A0, A1);
                        add3(Ub, Db, next_site(ca, i, w), B0, B1);  // (i, j+2) is the next site of the first colour
                    } else {
                        active = ~0ull;
                        const int i2 = i1 == lenX - 1 ? 0 : i1 + 1;
                        Bw = W(oc, i1, w);
                        add3(W(oc, im, w), W(oc, i, w), q ? next_site(oc, i, w) : prev_site(oc, i, w), A0, A1);
                        add3(W(ca, i2, w), W(ca, i1, w), q ? next_site(ca, i1, w) : prev_site(ca, i1, w), B0, B1);
                    }
                    int b;
                    const uint64_t acc = pair_update(S, Bw, A0, A1, B0, B1, on ? active : 0ull, pplan, (uint32_t)(i * nw + w),
                                                     greplica, 16 * pair + cls, L.key, b);
                    if (on && acc) {
                        dB += b;
                        W(ca, i, w) = S ^ acc;  // first sites: this thread is the only writer of the word
                        if (!horiz) {
                            W(oc, i1, w) = Bw ^ acc;
                        } else if (!q) {
                            W(oc, i, w) = Bw ^ acc;
                        } else {  // partners sit one lane up; lane 63 carries into the next word, which the
                                  // thread of that word also updates
                            xor_shared(oc, i, w, acc << 1);
                            if (acc >> 63) xor_shared(oc, i, w == nw - 1 ? 0 : w + 1, 1ull);
                        }
                    }
                }
                __syncthreads();
            }
            ++pair;
        }
    }

    // ---- store -----------------------------------------------------------------------------------
    if (LAYOUT == PYMC_LAYOUT_BITS) {
        uint64_t *dst = L.bits + (size_t)rep * 2 * plane;
        for (int idx = t; idx < 2 * plane; idx += T) dst[idx] = s_lat[idx];
    } else {
        uint4 *dst = reinterpret_cast<uint4 *>(L.i8 + (size_t)rep * 2 * lenX * L.nk);
        const uint16_t *s16 = reinterpret_cast<const uint16_t *>(s_lat);
        for (int c = t; c < 8 * plane; c += T) dst[c] = unpack16<MODEL>(s16[c]);
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

}  // namespace pmc
