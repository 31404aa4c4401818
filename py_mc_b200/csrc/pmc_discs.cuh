// pmc_discs.cuh -- K8: continuum hard discs in a periodic box, one warp per replica.
//
// Replaces HC_2D.energy / trialstep / newconfig and g_r of examples/2D_hardcore/model_setup.py
// (lines 24-37, 40-67, 106-125) driven by CanonicalMonteCarlo.MCstep (py_mc/mc.py:170-185).
//
// A trial (model_setup.py:45-60): pick disc id = randint(Ndisc); move it by maxstep*u with
// u = uniform(-1, 1) -- the SAME scalar on both coordinates in the reference (a diagonal move; kept as
// the "compat" mode, an isotropic move with two independent draws is the "physical" mode); dE = 0 if
// the moved disc overlaps no other disc under the minimum-image convention, else +inf.  MCstep then
// accepts dE = 0 always (exp(0) = 1 >= random()) and dE = inf only if random() == 0.
// All float64 operations are written with explicit round-to-nearest intrinsics in the reference's
// order (no FMA contraction), so coordinates stay bit-identical to numpy's.
//
// Random stream, domain 4: counter (trial_lo, replica, trial_hi, domain<<28 | call<<24); call 0 gives
//   o0 -> id = (o0 * Ndisc) >> 32;  (o1, o2) -> 53-bit uniform for the displacement;  o3 -> x32 of the
//   Metropolis test.  Call 1 (isotropic mode only) gives the second displacement from (o1, o2).
// The state is 16 bytes per disc: no HBM roofline applies, the kernel is latency/issue bound.
#pragma once
#include "pmc_stream.cuh"

namespace pmc {

struct DiscParams {
    double *coors;  // [nrep][ndisc][2]
    unsigned long long *accepted;  // [nrep]
    int nrep, ndisc;
    double Lcell, two_r, maxstep;
    int isotropic;
    PhiloxKey key;
    int replica_offset;
};

__device__ __forceinline__ double uniform53(uint32_t a, uint32_t b) {
    // (a >> 5) * 2^26 + (b >> 6), scaled by 2^-53: the construction CPython's random() uses
    return __dmul_rn(__dadd_rn(__dmul_rn((double)(a >> 5), 67108864.0), (double)(b >> 6)), 1.0 / 9007199254740992.0);
}

// minimum-image distance test, operation order of model_setup.py:53-55
__device__ __forceinline__ bool apart(double ox, double oy, double nx, double ny, double L, double two_r) {
    double sx = __dsub_rn(ox, nx), sy = __dsub_rn(oy, ny);
    sx = __dsub_rn(sx, __dmul_rn(rint(__ddiv_rn(sx, L)), L));
    sy = __dsub_rn(sy, __dmul_rn(rint(__ddiv_rn(sy, L)), L));
    const double dist = __dsub_rn(__dsqrt_rn(__dadd_rn(__dmul_rn(sx, sx), __dmul_rn(sy, sy))), two_r);
    return dist > 0.0;
}

struct DiscTrial {
    int id;
    double nx, ny;
    uint32_t x32;
};

__device__ __forceinline__ DiscTrial disc_trial(const DiscParams &P, int rep, uint64_t trial, double cx, double cy,
                                                int lane) {
    // every lane evaluates the same counter; lane `id` supplies the old coordinates
    const uint4 o = philox4x32_10((uint32_t)trial, (uint32_t)(P.replica_offset + rep), (uint32_t)(trial >> 32),
                                  ctr3(DOMAIN_DISC, 0, 0), P.key);
    DiscTrial t;
    t.id = (int)(((uint64_t)o.x * (uint64_t)P.ndisc) >> 32);
    t.x32 = o.w;
    const double u = __dadd_rn(-1.0, __dmul_rn(2.0, uniform53(o.y, o.z)));
    double ux = u, uy = u;
    if (P.isotropic) {
        const uint4 o2 = philox4x32_10((uint32_t)trial, (uint32_t)(P.replica_offset + rep), (uint32_t)(trial >> 32),
                                       ctr3(DOMAIN_DISC, 1, 0), P.key);
        uy = __dadd_rn(-1.0, __dmul_rn(2.0, uniform53(o2.y, o2.z)));
    }
    const double ox = __shfl_sync(0xFFFFFFFFu, cx, t.id & 31), oy = __shfl_sync(0xFFFFFFFFu, cy, t.id & 31);
    t.nx = __dadd_rn(ox, __dmul_rn(P.maxstep, ux));
    t.ny = __dadd_rn(oy, __dmul_rn(P.maxstep, uy));
    (void)lane;
    return t;
}

// One warp per replica, up to 32 discs (one per lane).  grid = ceil(nrep / warps_per_block).
__global__ void __launch_bounds__(128) discs_run_kernel(DiscParams P, uint64_t first_trial, long long ntrials) {
    const int lane = threadIdx.x & 31;
    const int rep = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (rep >= P.nrep) return;
    double *c = P.coors + (size_t)rep * P.ndisc * 2;
    const bool have = lane < P.ndisc;
    double cx = have ? c[2 * lane] : 0.0, cy = have ? c[2 * lane + 1] : 0.0;
    unsigned long long nacc = 0;
    for (long long s = 0; s < ntrials; ++s) {
        const DiscTrial t = disc_trial(P, rep, first_trial + (uint64_t)s, cx, cy, lane);
        const bool ok = !have || lane == t.id || apart(cx, cy, t.nx, t.ny, P.Lcell, P.two_r);
        const bool free_move = __all_sync(0xFFFFFFFFu, ok);
        // dE = 0: exp(0) = 1 >= u always;  dE = inf: exp(-inf) = 0, accepted only for u == 0
        const bool accept = free_move || t.x32 == 0u;
        if (accept) {
            ++nacc;
            if (lane == t.id) {
                cx = t.nx;
                cy = t.ny;
            }
        }
    }
    if (have) {
        c[2 * lane] = cx;
        c[2 * lane + 1] = cy;
    }
    if (lane == 0) P.accepted[rep] += nacc;
}

// trialstep probe: what trial `trial` of replica `rep` would propose, without applying it
__global__ void discs_propose_kernel(DiscParams P, int rep, uint64_t trial, int *id_out, double *new_out, int *overlap_out) {
    const int lane = threadIdx.x & 31;
    const double *c = P.coors + (size_t)rep * P.ndisc * 2;
    const bool have = lane < P.ndisc;
    const double cx = have ? c[2 * lane] : 0.0, cy = have ? c[2 * lane + 1] : 0.0;
    const DiscTrial t = disc_trial(P, rep, trial, cx, cy, lane);
    const bool ok = !have || lane == t.id || apart(cx, cy, t.nx, t.ny, P.Lcell, P.two_r);
    const bool free_move = __all_sync(0xFFFFFFFFu, ok);
    if (lane == 0) {
        *id_out = t.id;
        new_out[0] = t.nx;
        new_out[1] = t.ny;
        *overlap_out = free_move ? 0 : 1;
    }
}

// energy(): 1 if any pair overlaps (model_setup.py:24-37); g(r) pair counts (model_setup.py:116-122).
// One warp per replica; pair (a, b), a < b, handled by lane a looping over b.
__global__ void __launch_bounds__(128) discs_observe_kernel(DiscParams P, double dr, int nbins, int *overlap,
                                                            long long *counts) {
    const int lane = threadIdx.x & 31;
    const int rep = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (rep >= P.nrep) return;
    const double *c = P.coors + (size_t)rep * P.ndisc * 2;
    bool any = false;
    if (lane < P.ndisc) {
        const double ax = c[2 * lane], ay = c[2 * lane + 1];
        for (int b = lane + 1; b < P.ndisc; ++b) {
            // sij = coors[a] - coors[b] (model_setup.py:30,116), minimum image, norm
            double sx = __dsub_rn(ax, c[2 * b]), sy = __dsub_rn(ay, c[2 * b + 1]);
            sx = __dsub_rn(sx, __dmul_rn(rint(__ddiv_rn(sx, P.Lcell)), P.Lcell));
            sy = __dsub_rn(sy, __dmul_rn(rint(__ddiv_rn(sy, P.Lcell)), P.Lcell));
            const double dist = __dsqrt_rn(__dadd_rn(__dmul_rn(sx, sx), __dmul_rn(sy, sy)));
            if (!(__dsub_rn(dist, P.two_r) > 0.0)) any = true;
            if (counts) {
                const double bin = rint(__ddiv_rn(dist, dr));  // np.around(dist/dr)
                if (bin >= 1.0 && bin <= (double)nbins)
                    atomicAdd((unsigned long long *)(counts + (size_t)rep * nbins + ((int)bin - 1)), 2ull);
            }
        }
    }
    any = __any_sync(0xFFFFFFFFu, any);
    if (lane == 0 && overlap) overlap[rep] = any ? 1 : 0;
}

}  // namespace pmc
