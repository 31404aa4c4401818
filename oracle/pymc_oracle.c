/* pymc_oracle.c -- C restatement of the Ising checkerboard Metropolis sweep.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): built by oracle/cbuild.py into
 * oracle/_build/libpymc_oracle.so and called from tests (and nothing else) to check the CUDA kernels on
 * lattices too large for the numpy oracle.  It is a scalar, site-by-site program on purpose -- no bit
 * slicing, no tiles -- so that it shares no structure with the kernels it checks.
 *
 * What it follows (reference tree): the trial of examples/ising2D.py:28-47
 * (dE = -2 J s (up+down+left+right), periodic), the update of :49-52 (s -> -s) and the acceptance rule of
 * py_mc/mc.py:175-184 in its integer form x32 <= thr[a] (a = number of aligned neighbours; thresholds
 * computed by the caller with the host libm, oracle/ising.py:threshold_table), visiting the sites in
 * checkerboard order with the counter-based stream of oracle/philox.py ("pmc stream": 8 head bit planes
 * per 64-site word + a 24-bit tail block per lane).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define M0 0xD2511F53u
#define M1 0xCD9E8D57u
#define W0 0x9E3779B9u
#define W1 0xBB67AE85u
#define HEAD_PLANES 8
#define TAIL_CALL 15u

static void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)M0 * c[0], p1 = (uint64_t)M1 * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        c[0] = n0;
        c[1] = (uint32_t)p1;
        c[2] = n2;
        c[3] = (uint32_t)p0;
        k0 += W0;
        k1 += W1;
    }
}

static uint32_t ctr3(uint32_t domain, uint32_t call, uint64_t step, uint32_t lane) {
    return (domain << 28) | (call << 24) | (lane << 18) | ((uint32_t)(step >> 32) & 0x3FFFFu);
}

/* x32 of all lanes of one word */
static void word_x32(uint64_t seed, uint32_t replica, uint64_t step, uint32_t domain, uint32_t word, int nlanes,
                     uint32_t *x) {
    const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    memset(x, 0, 64 * sizeof(uint32_t));
    for (uint32_t call = 0; call < HEAD_PLANES / 2; ++call) {
        uint32_t c[4] = {word, replica, (uint32_t)step, ctr3(domain, call, step, 0)};
        philox4x32_10(c, k0, k1);
        const uint64_t even = (uint64_t)c[0] | ((uint64_t)c[1] << 32), odd = (uint64_t)c[2] | ((uint64_t)c[3] << 32);
        for (int l = 0; l < nlanes; ++l) {
            x[l] |= (uint32_t)((even >> l) & 1u) << (31 - 2 * call);
            x[l] |= (uint32_t)((odd >> l) & 1u) << (30 - 2 * call);
        }
    }
    for (int l = 0; l < nlanes; ++l) {
        uint32_t c[4] = {word, replica, (uint32_t)step, ctr3(domain, TAIL_CALL, step, (uint32_t)l)};
        philox4x32_10(c, k0, k1);
        x[l] |= c[0] >> HEAD_PLANES;
    }
}

/* One full sweep (colour 0, then colour 1) of an int8 +1/-1 lattice [lenX][lenY], in place.
 * thr[a], a = aligned neighbours 0..4.  Returns the number of accepted flips; *dB / *dM get the changes
 * of the bond sum and the magnetisation. */
long long pymc_oracle_ising_sweep(int8_t *cfg, int lenX, int lenY, const uint32_t *thr, uint64_t seed,
                                  uint32_t replica, uint64_t sweep, long long *dB, long long *dM) {
    const int nk = lenY / 2, nw = (nk + 63) / 64;
    long long nacc = 0, db = 0, dm = 0;
    uint32_t x[64];
    for (int colour = 0; colour < 2; ++colour) {
        const uint64_t step = 2 * sweep + (uint64_t)colour;
        for (int i = 0; i < lenX; ++i) {
            const int q = (i + colour) & 1;
            const int im = i == 0 ? lenX - 1 : i - 1, ip = i == lenX - 1 ? 0 : i + 1;
            for (int w = 0; w < nw; ++w) {
                const int nl = nk - 64 * w < 64 ? nk - 64 * w : 64;
                word_x32(seed, replica, step, 0u, (uint32_t)(i * nw + w), nl, x);
                for (int l = 0; l < nl; ++l) {
                    const int j = 2 * (64 * w + l) + q;
                    const int jm = j == 0 ? lenY - 1 : j - 1, jp = j == lenY - 1 ? 0 : j + 1;
                    const int s = cfg[(size_t)i * lenY + j];
                    const int nn = cfg[(size_t)im * lenY + j] + cfg[(size_t)ip * lenY + j] + cfg[(size_t)i * lenY + jm] +
                                   cfg[(size_t)i * lenY + jp];
                    const int a = (s * nn + 4) / 2; /* aligned neighbours */
                    if (x[l] <= thr[a]) {
                        cfg[(size_t)i * lenY + j] = (int8_t)-s;
                        ++nacc;
                        db += -2 * s * nn;
                        dm += -2 * s;
                    }
                }
            }
        }
    }
    if (dB) *dB = db;
    if (dM) *dM = dm;
    return nacc;
}

/* B = sum s(i,j) (s(i-1,j) + s(i,j-1)), M = sum s  (examples/ising2D.py:15-26) */
void pymc_oracle_ising_observables(const int8_t *cfg, int lenX, int lenY, long long *B, long long *M) {
    long long b = 0, m = 0;
    for (int i = 0; i < lenX; ++i) {
        const int im = i == 0 ? lenX - 1 : i - 1;
        for (int j = 0; j < lenY; ++j) {
            const int jm = j == 0 ? lenY - 1 : j - 1;
            const int s = cfg[(size_t)i * lenY + j];
            b += s * (cfg[(size_t)im * lenY + j] + cfg[(size_t)i * lenY + jm]);
            m += s;
        }
    }
    *B = b;
    *M = m;
}

/* hot start, stream domain 1: +1 iff x32 >= 2^31 (examples/ising2D.py:64-70) */
void pymc_oracle_ising_init(int8_t *cfg, int lenX, int lenY, uint64_t seed, uint32_t replica) {
    const size_t n = (size_t)lenX * lenY;
    for (size_t call = 0; call * 4 < n; ++call) {
        uint32_t c[4] = {(uint32_t)call, replica, 0u, ctr3(1u, 0u, 0u, 0u)};
        philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
        for (int e = 0; e < 4 && call * 4 + e < n; ++e) cfg[call * 4 + e] = c[e] >= 0x80000000u ? 1 : -1;
    }
}
