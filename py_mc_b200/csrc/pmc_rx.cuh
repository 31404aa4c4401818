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

// ---- time series by temperature and their blocking analysis on the device --------------------------
// series[sample][T][2] = (E, |M|) of whichever replica sits at temperature index T at that sample: the regrouping
// examples/analyze_ising2D_mpi.py:28-34 does afterwards from obs_save.npy / Trank_hist.npy (np.where + Reduce).
// One thread per local replica; entries of replicas on other GPUs stay 0 (the caller adds the GPUs up).
__global__ void series_record_kernel(const unsigned long long *__restrict__ tracked, int nrep, int offset,
                                     const int32_t *__restrict__ rep_to_T, double c_bond, double c_second, int ntemps,
                                     double *__restrict__ row) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nrep) return;
    const double b = (double)(long long)tracked[2 * r], m = (double)(long long)tracked[2 * r + 1];
    double e = __dmul_rn(c_bond, b);
    if (c_second != 0.0) e = __dadd_rn(e, __dmul_rn(m, c_second));
    const int t = rep_to_T[offset + r];
    row[2 * t] = e;
    row[2 * t + 1] = fabs(m);
}

// Sum of the n values v(i) over a block of 256 threads (every thread gets the result).
template <class F>
__device__ __forceinline__ double block_sum_256(int n, F v) {
    __shared__ double s_part[8];
    __shared__ double s_total;
    double a = 0.0;
    for (int i = threadIdx.x; i < n; i += 256) a += v(i);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) a += __shfl_down_sync(0xFFFFFFFFu, a, d);
    __syncthreads();  // s_part / s_total of a previous call have been read
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int k = 0; k < 8; ++k) t += s_part[k];
        s_total = t;
    }
    __syncthreads();
    return s_total;
}

// binning() of py_mc/mc.py:47-61 for every (temperature, observable) series at once: block (T, j) drops the leading
// count % 2^nlevels samples, then per level takes sqrt(var(x, ddof=1) / len(x)) and averages neighbouring pairs
// ((a + b) / 2.0, the reference's expression).  Also the mean over samples [discard, count) that
// analyze_ising2D_mpi.py:36 prints.  work: [ntemps * 2][2][capacity] ping-pong buffers.  block = 256 threads.
__global__ void __launch_bounds__(256) series_binning_kernel(const double *__restrict__ series, int count, int ntemps,
                                                             int discard, int nlevels, double *work, int capacity,
                                                             double *__restrict__ mean_out, double *__restrict__ err_out) {
    const int sidx = blockIdx.x;  // T * 2 + j
    double *buf0 = work + (size_t)sidx * 2 * capacity, *buf1 = buf0 + capacity;
    const int drop = count % (1 << nlevels);
    int n = count - drop;
    for (int i = threadIdx.x; i < count; i += 256) {
        const double v = series[(size_t)i * ntemps * 2 + sidx];
        if (i >= drop) buf0[i - drop] = v;
    }
    __syncthreads();
    const int nm = count - discard;
    const double msum = block_sum_256(nm, [&](int i) { return series[(size_t)(i + discard) * ntemps * 2 + sidx]; });
    if (threadIdx.x == 0) mean_out[sidx] = nm > 0 ? msum / nm : 0.0;
    double *x = buf0, *y = buf1;
    for (int lvl = 0; lvl <= nlevels; ++lvl) {
        const double mu = block_sum_256(n, [&](int i) { return x[i]; }) / n;
        const double ss = block_sum_256(n, [&](int i) { const double d = x[i] - mu; return d * d; });
        if (threadIdx.x == 0) err_out[(size_t)sidx * (nlevels + 1) + lvl] = sqrt(ss / (n - 1) / n);
        if (lvl == nlevels) break;
        for (int i = threadIdx.x; i < n / 2; i += 256) y[i] = (x[2 * i] + x[2 * i + 1]) / 2.0;
        __syncthreads();
        n /= 2;
        double *tmp = x;
        x = y;
        y = tmp;
    }
}

__global__ void rx_invert_map_kernel(const int32_t *__restrict__ rep_to_T, int32_t *T_to_rep, int n) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < n) T_to_rep[rep_to_T[r]] = r;
}

}  // namespace pmc
