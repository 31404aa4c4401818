// pmc_rx.cuh -- K4: temperature replica exchange decided on the device.
//
// Replaces TemperatureRX_MPI.Xtrial (py_mc/mc_mpi.py:102-140).  The reference pairs neighbouring
// temperatures (t, t+1) with (t + XCscheme) even; the rank holding the lower temperature receives
// the partner's energy, forms  delta = (beta[t+1] - beta[t]) * (E_t - E_{t+1})  and accepts iff
// delta < 0 or random() <= exp(-delta); on acceptance the two ranks swap temperature labels
// (configurations stay where they are) and the rank->temperature map is re-synchronised by an
// Allgather.  Here every GPU holds the all-gathered energies E[R] and runs this kernel on them, so
// all GPUs derive the same new map with no second message: the uniform of pair t in exchange
// round n is Philox(domain 2, idx = t, step = n), a pure function of (seed, n, t).
#pragma once
#include "pmc_stream.cuh"

namespace pmc {

// E[offset + r] = c_bond * bond_r + c_second * second_r, evaluated without FMA contraction so that
// it equals the host expression  J*B + (Eads-mu)*N  bit for bit.
__global__ void rx_energy_kernel(const unsigned long long *__restrict__ tracked, int nrep, int offset, double c_bond,
                                 double c_second, double *__restrict__ energies) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nrep) return;
    const double b = (double)(long long)tracked[2 * r], s = (double)(long long)tracked[2 * r + 1];
    double e = __dmul_rn(c_bond, b);
    if (c_second != 0.0) e = __dadd_rn(e, __dmul_rn(s, c_second));
    energies[offset + r] = e;
}

__global__ void rx_decide_kernel(const double *__restrict__ energies, const double *__restrict__ betas,
                                 int32_t *rep_to_T, int32_t *T_to_rep, int ntemps, int xcscheme, uint64_t round,
                                 PhiloxKey key, unsigned long long *tried, unsigned long long *accepted) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    const int t = 2 * p + (xcscheme & 1);  // lower temperature index of the pair
    if (t + 1 >= ntemps) return;
    const int r_lo = T_to_rep[t], r_hi = T_to_rep[t + 1];
    const double delta = __dmul_rn(__dsub_rn(betas[t + 1], betas[t]), __dsub_rn(energies[r_lo], energies[r_hi]));
    bool accept = delta < 0.0;
    if (!accept) {
        const uint4 o = philox4x32_10((uint32_t)t, 0u, (uint32_t)round, ctr3(DOMAIN_RX, 0, round), key);
        const double u = (double)o.x * (1.0 / 4294967296.0);
        accept = u <= exp(-delta);
    }
    tried[t] += 1;
    if (accept) {
        accepted[t] += 1;
        rep_to_T[r_lo] = t + 1;
        rep_to_T[r_hi] = t;
        T_to_rep[t] = r_hi;
        T_to_rep[t + 1] = r_lo;
    }
}

// K9: per-temperature accumulators, the device-side form of  obs[rank_to_T[rank]] += obs_step
// (py_mc/mc_mpi.py:172) for the standard observables: E, E^2, |M|, M^2, M^4, M (M = second tracked
// observable: magnetisation or particle number).  One thread per local replica; several replicas may
// share a temperature row (independent replicas at equal T), hence atomics.
constexpr int NSTAT = 6;
__global__ void stats_accumulate_kernel(const unsigned long long *__restrict__ tracked, int nrep, int offset,
                                        const int32_t *__restrict__ rep_to_T, double c_bond, double c_second,
                                        double *stats, unsigned long long *counts) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nrep) return;
    const double b = (double)(long long)tracked[2 * r], m = (double)(long long)tracked[2 * r + 1];
    double e = __dmul_rn(c_bond, b);
    if (c_second != 0.0) e = __dadd_rn(e, __dmul_rn(m, c_second));
    const int t = rep_to_T[offset + r];
    double *row = stats + (size_t)t * NSTAT;
    const double m2 = m * m;
    atomicAdd(row + 0, e);
    atomicAdd(row + 1, e * e);
    atomicAdd(row + 2, fabs(m));
    atomicAdd(row + 3, m2);
    atomicAdd(row + 4, m2 * m2);
    atomicAdd(row + 5, m);
    atomicAdd(counts + t, 1ull);
}

__global__ void rx_invert_map_kernel(const int32_t *__restrict__ rep_to_T, int32_t *T_to_rep, int n) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < n) T_to_rep[rep_to_T[r]] = r;
}

}  // namespace pmc
