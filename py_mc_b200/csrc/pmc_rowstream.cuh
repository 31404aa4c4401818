// pmc_rowstream.cuh -- K13: one full checkerboard sweep (both colours) of the int8 layout per launch, with the rows
// of the lattice STREAMED through shared memory.
//
// What it replaces: lenX*lenY calls of CanonicalMonteCarlo.MCstep (py_mc/mc.py:170-185) on ising2D / latticegas
// trialstep (examples/ising2D.py:28-52, examples/latgas.py:31-57), driven in checkerboard order -- for calls of a
// single sweep (sample_frequency = 1) and for lattices too large to stay on chip between sweeps.
//
// Why another int8 sweep kernel.  The tiled kernel K1f (pmc_fused_i8.cuh) cuts the lattice into 32 x 512-site
// tiles: every tile recomputes a ring of halo sites, spends a fifth of its instructions on tile addressing and
// per-thread global loads, and parks 6 of 8 warps at a barrier behind the ring (profiles/README.md: 960
// instructions per 64-site word against 547 for the bit-packed kernel, 47 % of the HBM roofline).  Here a block owns
// a PIECE of whole rows of one replica and walks down it:
//   * rows arrive in shared memory by bulk asynchronous copies (cp.async.bulk + mbarrier: one instruction per row
//     from one lane, no per-thread address arithmetic, no registers held across the latency), two staging slots;
//   * every thread packs 64 bytes to bits, updates one 64-site word and unpacks one per iteration; full-width rows
//     wrap onto themselves, so there is no halo in y at all, and in x a piece recomputes ONE extra row of colour 0
//     at either end (2 row updates per piece instead of a ring per tile);
//   * colour 1 of row j follows colour 0 of row j+1 at a fixed distance in one task sequence, so both colours are
//     done in one pass over the data: every site is read once and written once per sweep (2 B per attempt), out
//     of place (the handle ping-pongs two buffers); the warps of a block synchronise through mbarriers, two
//     iterations back (below), not at a block barrier.
// Random-stream indexing is that of every other sweep kernel (word = row * nw + w, replica, step = 2*sweep +
// colour), so the trajectory is identical.
//
// Task sequence of a piece with rows [a, a+n).  c0[m] = colour 0 of row a-1+m (m = 0 .. n+1; the first and last
// are the recomputed halo rows, not stored), c1[j] = colour 1 of row a+j (j = 0 .. n-1).  c1[j] needs the NEW c0 of
// rows a+j-1 .. a+j+1, i.e. c0[j], c0[j+1], c0[j+2]; c0[m] needs the OLD c1 of rows a+m-2 .. a+m, the last of which
// c1[m-2] overwrites.  With Q row tasks per iteration (Q = threads / words per row) the order
//       c0[0] .. c0[2Q]   c0[2Q+1] c1[0]   c0[2Q+2] c1[1]   ...   c0[n+1] c1[n-2Q]   c1[n-2Q+1] .. c1[n-1]
// keeps every such pair at least 2Q positions = TWO iterations apart.  That is what lets the block run without a
// barrier in the usual sense: a warp starting iteration `it` waits only until every warp has finished iteration
// it - 2 (an mbarrier per iteration parity, one arrival per warp), so the warps of a block may be a whole iteration
// apart and the spread of the per-lane tail rounds of the Metropolis test -- which parked warps at a __syncthreads
// for 12 % of their time -- is absorbed.  Everything obeys the same rule: nothing is read that was written less
// than two iterations earlier (packing runs that far ahead of the sequence, the stores of finished rows trail it
// by two), nothing is overwritten that was read less than two iterations earlier (ring size), a staging slot is
// refilled two iterations after it was packed from, and the tables of an iteration are built two iterations
// before it (by a different warp each time).  Bit words live in two rings of rs_ring_rows(Q) rows (one per colour);
// the emulation in tests/hostemu checks every read and write against that rule, for what the ring slot holds at
// that time.
#pragma once
#include "pmc_core.cuh"
#include "pmc_pack.cuh"

#ifndef PMC_RS_THREADS
#define PMC_RS_THREADS 512
#endif

namespace pmc {

constexpr int RS_THREADS = PMC_RS_THREADS;
constexpr int RS_STAGES = 2;   // staging slots for the bulk copies (one iteration's rows each)
constexpr int RS_MAXQ = 64;    // row tasks per iteration: words per colour row >= RS_THREADS / 64
constexpr int RS_TABLES = 4;   // iterations whose tables are kept: built two iterations ahead, read by warps up to one apart

// launch-uniform schedule constants (kernel parameter: constant memory)
struct RSParams {
    int lenX, nk, nw, nw_log2;
    int Q, q_log2;     // row tasks per iteration (a power of two)
    int lead;          // c0 tasks before the first c1 task
    int prologue;      // iterations that only pack
    int rw;            // ring rows per colour
    int pieces;        // pieces per replica
};

// The schedule constants below were found by searching the smallest values for which no hazard occurs when the warps
// of a block may be up to one whole iteration apart (see the kernel), and are checked for every geometry by the emulation
// in tests/hostemu (which refuses a read of anything written less than two iterations earlier, and a write over anything
// read less than two iterations earlier).
// c0 tasks before the first c1 task: colour 1 of row j reads colour 0 of row j+1 from two iterations back
__host__ __device__ constexpr int rs_lead(int Q) { return 2 * Q + 1; }
// ring rows per colour
__host__ __device__ constexpr int rs_ring_rows(int Q) { return (11 * Q + 1) / 2 + (Q < 4 ? 1 : 0); }
// iterations that only pack: a task reads rows packed at least two iterations before it runs
__host__ __device__ constexpr int rs_prologue(int Q) { return (2 * Q + 4 + Q - 1) / Q + 2; }
__host__ __device__ constexpr int rs_min_rows(int Q) { return 4 * Q + 4; }  // smallest piece the task order is valid for
constexpr int RS_LAG = 2;  // iterations between a task and the storing of its row (and between any write and a read of it)

// false if the shape cannot be streamed with RS_THREADS threads
__host__ inline bool rs_make_params(int lenX, int nk, int pieces, RSParams &p) {
    if (nk % 64) return false;
    const int nw = nk / 64;
    if (nw < 1 || (nw & (nw - 1)) || nw > RS_THREADS || RS_THREADS / nw > RS_MAXQ) return false;  // 1 <= Q <= RS_MAXQ
    p.lenX = lenX;
    p.nk = nk;
    p.nw = nw;
    p.nw_log2 = 0;
    while ((1 << p.nw_log2) < nw) ++p.nw_log2;
    p.Q = RS_THREADS / nw;
    p.q_log2 = 0;
    while ((1 << p.q_log2) < p.Q) ++p.q_log2;
    p.lead = rs_lead(p.Q);
    p.prologue = rs_prologue(p.Q);
    p.rw = rs_ring_rows(p.Q);
    p.pieces = pieces;
    return pieces >= 1 && lenX / pieces >= rs_min_rows(p.Q);
}

__host__ __device__ inline size_t rs_smem_bytes(const RSParams &p) {
    return (size_t)RS_STAGES * RS_THREADS * 64 + (size_t)(2 * p.rw + 1) * p.nw * 8;  // staging, two rings, one spare row
}

__host__ __device__ __forceinline__ void rs_piece(const RSParams &p, int piece, int &a, int &n) {
    a = (int)(((long long)piece * p.lenX) / p.pieces);
    n = (int)(((long long)(piece + 1) * p.lenX) / p.pieces) - a;
}

// A row to copy in.  info: bits 0-15 ring row, bit 16 colour, bit 17 valid.  row: row in the lattice.
struct RSTask {
    int info, row;
};
constexpr int RS_COLOUR = 1 << 16, RS_VALID = 1 << 17;

__host__ __device__ __forceinline__ int rs_wrap(int i, int lenX) { return i < 0 ? i + lenX : (i >= lenX ? i - lenX : i); }

// position in the task sequence -> (colour, index): c0[idx] or c1[idx]; false beyond the end
__host__ __device__ inline bool rs_task_index(const RSParams &p, int n, int pos, int &colour, int &idx) {
    if (pos < 0 || pos >= 2 * n + 2) return false;
    const int L = p.lead;  // n >= rs_min_rows(Q) > L
    if (pos < L) {
        colour = 0;
        idx = pos;
        return true;
    }
    const int q = pos - L, ninter = n + 2 - L;  // (c0, c1) pairs after the lead
    if (q < 2 * ninter) {
        colour = q & 1;
        idx = colour ? q >> 1 : L + (q >> 1);
    } else {
        colour = 1;
        idx = ninter + (q - 2 * ninter);
    }
    return true;
}

// position in the packing sequence -> row to pack:  c1h[0] c1h[1]  c0[0] c1h[2]  c0[1] c1h[3] ...  (c1h[u] = colour 1
// of row a-2+u, u = 0 .. n+3)
// (colour, index) of packing position pp: colour 0 -> c0[idx], colour 1 -> c1h[idx]
__host__ __device__ inline bool rs_pack_index(int n, int pp, int &colour, int &idx) {
    if (pp < 0 || pp >= 2 * n + 6) return false;
    if (pp < 2 || (pp & 1)) {
        colour = 1;
        idx = pp < 2 ? pp : ((pp - 3) >> 1) + 2;
    } else {
        colour = 0;
        idx = (pp - 2) >> 1;
    }
    return true;
}

__host__ __device__ inline RSTask rs_pack_task(const RSParams &p, int a, int n, int pp) {
    RSTask t = {0, 0};
    int colour, idx;
    if (!rs_pack_index(n, pp, colour, idx)) return t;
    t.row = rs_wrap(a - (colour ? 2 : 1) + idx, p.lenX);
    t.info = (idx % p.rw) | (colour ? RS_COLOUR : 0) | RS_VALID;
    return t;
}

// ---- per-iteration tables ------------------------------------------------------------------------------------
// Everything a thread needs about the row task it works on is worked out once per task and iteration (by one lane
// per task) and read from shared memory, instead of being re-derived by every thread.  The two rings are one array W
// of 2*RW + 1 rows (colour 1 follows colour 0); a "word base" indexes it as 64-bit words.
// The last row is a spare: a position beyond the end of a sequence is given that row to work on, so that the phases
// need no branch around their loads and stores (only the store to global memory is predicated).
struct RSCompute {
    int own, ru, rc, rd;  // word bases of the task's row and of the rows above / same / below in the other colour
    int widx;             // row * nw: random-stream index of word 0 of the row
    int flags;            // bit 0 valid, 1 owned, 2 colour, 3 the same-row neighbour is site k+1 (else k-1)
    int pad0, pad1;
};
struct RSMove {
    int base;  // word base of the row in W (the spare row if there is nothing to do)
    int off;   // unpack: offset of the row in the destination, in 16-byte units, or -1: do not store
};

__host__ __device__ inline RSCompute rs_build_compute(const RSParams &p, int a, int n, int pos) {
    const int RW = p.rw, spare = (2 * RW) << p.nw_log2;
    RSCompute c = {spare, spare, spare, spare, 0, 0, 0, 0};
    int colour, idx;
    if (!rs_task_index(p, n, pos, colour, idx)) return c;
    // c0[m]: ring row m, reads c1h[m .. m+2];  c1[j] = c1h[j+2]: ring row j+2, reads c0[j .. j+2]
    const int r = colour ? idx + 2 : idx, row = rs_wrap(a - 1 + idx + colour, p.lenX);
    const int oth = colour ? 0 : RW;
    c.own = ((colour ? RW : 0) + r % RW) << p.nw_log2;
    c.ru = (oth + idx % RW) << p.nw_log2;
    c.rc = (oth + (idx + 1) % RW) << p.nw_log2;
    c.rd = (oth + (idx + 2) % RW) << p.nw_log2;
    c.widx = row * p.nw;
    const bool owned = colour || (idx >= 1 && idx <= n);
    c.flags = 1 | (owned ? 2 : 0) | (colour ? 4 : 0) | (((row + colour) & 1) ? 8 : 0);
    return c;
}

__host__ __device__ inline RSMove rs_build_pack(const RSParams &p, int n, int pp) {
    RSMove m = {(2 * p.rw) << p.nw_log2, -1};
    int colour, idx;
    if (!rs_pack_index(n, pp, colour, idx)) return m;
    m.base = ((colour ? p.rw : 0) + idx % p.rw) << p.nw_log2;
    return m;
}

// the task at position pos was computed RS_LAG iterations ago: where its row goes
__host__ __device__ inline RSMove rs_build_unpack(const RSParams &p, int a, int n, int pos) {
    RSMove m = {(2 * p.rw) << p.nw_log2, -1};
    int colour, idx;
    if (!rs_task_index(p, n, pos, colour, idx)) return m;
    if (!colour && (idx < 1 || idx > n)) return m;  // recomputed halo row: another piece stores it
    const int r = colour ? idx + 2 : idx, row = rs_wrap(a - 1 + idx + colour, p.lenX);
    m.base = ((colour ? p.rw : 0) + r % p.rw) << p.nw_log2;
    m.off = (int)((((long long)colour * p.lenX + row) * p.nk) >> 4);
    return m;
}

// 16 sites -> 16 bits with the integer dot product: for bytes s = +1/-1 the weights (1,2,4,8 | 16,32,64,-128) give
// v = 2 n + 1 - 256 b7 for the byte n of occupation bits (b = (s+1)/2), so n = ((v - 1) >> 1) mod 256; for bytes 0/1
// v = n - 256 b7.  Two chained DP4A per 8 sites instead of two (mask, multiply, shift, merge) groups.
template <int MODEL>
__device__ __forceinline__ uint32_t rs_pack16(uint4 v) {
#if defined(__CUDA_ARCH__) && !defined(PMC_RS_NO_DP4A)
    const int K0 = 0x08040201, K1 = (int)0x80402010;
    if (MODEL == MODEL_ISING) {
        const int lo = __dp4a((int)v.x, K0, __dp4a((int)v.y, K1, -1));
        const int hi = __dp4a((int)v.z, K0, __dp4a((int)v.w, K1, -1));
        return (((uint32_t)lo >> 1) & 0xFFu) | (((uint32_t)hi << 7) & 0xFF00u);
    } else {
        const int lo = __dp4a((int)v.x, K0, __dp4a((int)v.y, K1, 0));
        const int hi = __dp4a((int)v.z, K0, __dp4a((int)v.w, K1, 0));
        return ((uint32_t)lo & 0xFFu) | (((uint32_t)hi << 8) & 0xFF00u);
    }
#else
    return pack16<MODEL>(v);
#endif
}

// ---- the three per-thread phases of an iteration (also driven on the host by tests/hostemu) --------------------
// chunk = 16 bytes = 16 sites of one colour row; a row has 4*nw chunks.  In the packing and unpacking phases thread t
// handles chunks t + e*RS_THREADS, e = 0..3 (four different rows): consecutive lanes read / store consecutive 16 bytes.
// (Measured slower: one whole word = 64 contiguous bytes per thread, with the four 128-bit shared-memory loads rotated
// so that they do not conflict -- +4 % instructions for the variable shifts when packing, and 128-bit global stores
// 64 bytes apart when unpacking: 0.164 and 0.185 ms against 0.160.)  NWL2 = log2(words per colour row) is a template
// parameter: the index arithmetic folds into constants.

// bytes of the rows of this iteration's packing tasks (staged contiguously, chunk c at stage[c]) -> bits in the rings
template <int MODEL, int NWL2>
__device__ __forceinline__ void rs_pack(const RSMove *tab, const uint4 *stage, uint64_t *W, int t) {
    constexpr int CSH = NWL2 + 2;
    uint16_t *s16 = reinterpret_cast<uint16_t *>(W);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        const int c = t + e * RS_THREADS;
        s16[4 * tab[c >> CSH].base + (c & ((1 << CSH) - 1))] = (uint16_t)rs_pack16<MODEL>(stage[c]);
    }
}

// one word of one row task
template <int MODEL, int NWL2>
__device__ __forceinline__ void rs_compute(const RSCompute *tab, uint64_t *W, const Plan<MODEL> &plan, uint32_t greplica,
                                           uint64_t sweep, const PhiloxKey &key, int t, int &dB, int &dS) {
    constexpr int NW = 1 << NWL2;
    const int w = t & (NW - 1);
    const RSCompute c = tab[t >> NWL2];
    const uint64_t S = W[c.own + w], U = W[c.ru + w], C = W[c.rc + w], D = W[c.rd + w];
    // the other same-row neighbour: site k+1 (bit 3 of flags) or k-1 -- one load of the word beside C, one funnel shift
    const bool right = c.flags & 8;
    const uint64_t nb = W[c.rc + ((w + (right ? 1 : NW - 1)) & (NW - 1))];
    const uint64_t Sd = right ? ((C >> 1) | (nb << 63)) : ((C << 1) | (nb >> 63));
    int b, s2;
    const uint64_t Snew = update_word<MODEL>(S, U, D, C, Sd, (c.flags & 1) ? ~0ull : 0ull, plan, (uint32_t)(c.widx + w), greplica,
                                             2 * sweep + (uint64_t)((c.flags >> 2) & 1), key, b, s2);
    W[c.own + w] = Snew;  // a position beyond the end works on the spare row, with no valid lane
    if (c.flags & 2) {
        dB += b;
        dS += s2;
    }
}

// bits of the rows the previous iteration computed -> bytes in the destination lattice of this replica
template <int MODEL, int NWL2>
__device__ __forceinline__ void rs_unpack(const RSMove *tab, const uint64_t *W, uint4 *dst16, int t) {
    constexpr int CSH = NWL2 + 2;
    const uint16_t *s16 = reinterpret_cast<const uint16_t *>(W);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        const int c = t + e * RS_THREADS, cm = c & ((1 << CSH) - 1);
        const RSMove m = tab[c >> CSH];
        const uint4 v = unpack16<MODEL>(s16[4 * m.base + cm]);
        if (m.off >= 0) dst16[(uint32_t)(m.off + cm)] = v;
    }
}

#ifdef __CUDACC__

// ---- mbarrier / bulk-copy wrappers (PTX; sm_90+) -------------------------------------------------------------
__device__ __forceinline__ uint32_t rs_smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void rs_mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(rs_smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void rs_mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(rs_smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void rs_mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(rs_smem_addr(bar)) : "memory");
}
// One attempt to see phase `parity` of the barrier complete.  The hardware suspends the thread until the phase completes or
// about `hint_ns` have passed, so a waiting warp issues next to nothing (a loop of plain polls took 12 % of all issue slots).
__device__ __forceinline__ bool rs_mbar_try_wait(uint64_t *bar, uint32_t parity, uint32_t hint_ns = 1000000u) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(rs_smem_addr(bar)), "r"(parity), "r"(hint_ns)
        : "memory");
    return ok != 0;
}
// global -> shared, `bytes` a multiple of 16, both addresses 16-byte aligned; completion is counted on `bar`
__device__ __forceinline__ void rs_bulk_load(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     rs_smem_addr(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(rs_smem_addr(bar))
                 : "memory");
}

// what a block works out once and reads back where it is needed (registers are too scarce to keep it there)
struct RSBlock {
    const int8_t *src_rep;
    uint4 *dst16;
    int a, n, ntasks, npacks, niter;
};

// grid = (pieces, nrep), block = RS_THREADS, dynamic shared memory = rs_smem_bytes(p)
template <int MODEL, int NWL2, class LatT>
__global__ void __launch_bounds__(RS_THREADS, 1024 / RS_THREADS)
    rowstream_kernel(const __grid_constant__ LatT L, const __grid_constant__ RSParams p_in, const int8_t *__restrict__ src,
                     int8_t *__restrict__ dst, uint64_t sweep, int *err) {
    constexpr int Q = RS_THREADS >> NWL2, QL2 = 9 - NWL2 + (RS_THREADS == 256 ? -1 : (RS_THREADS == 1024 ? 1 : 0));
    static_assert((1 << QL2) == Q && Q >= 1 && Q <= RS_MAXQ, "row tasks per iteration");
    constexpr int NWARPS = RS_THREADS / 32;
    extern __shared__ __align__(128) unsigned char rs_smem[];
    __shared__ __align__(8) uint64_t s_full[RS_STAGES];  // the bulk copies of a staging slot have landed
    __shared__ __align__(8) uint64_t s_done[2];          // every warp has finished an iteration (even / odd ones)
    __shared__ __align__(16) RSCompute s_ct[RS_TABLES * Q];
    __shared__ __align__(8) RSMove s_pt[RS_TABLES * Q], s_ut[RS_TABLES * Q];
    __shared__ RSBlock s_blk;
    __shared__ int s_red[2][NWARPS];
    uint4 *const stage = reinterpret_cast<uint4 *>(rs_smem);
    uint64_t *const W = reinterpret_cast<uint64_t *>(rs_smem + (size_t)RS_STAGES * RS_THREADS * 64);

    // the schedule constants that follow from the template parameter, as compile-time values (the table builders
    // below divide by the ring size: a constant divisor costs a multiplication, a run-time one a division routine)
    RSParams p = p_in;
    p.nw = 1 << NWL2;
    p.nw_log2 = NWL2;
    p.Q = Q;
    p.q_log2 = QL2;
    p.lead = rs_lead(Q);
    p.prologue = rs_prologue(Q);
    p.rw = rs_ring_rows(Q);

    const int t = threadIdx.x, rep = blockIdx.y, warp = t >> 5;
    if (t == 0) {
        int a, n;
        rs_piece(p, blockIdx.x, a, n);
        s_blk.a = a;
        s_blk.n = n;
        s_blk.ntasks = 2 * n + 2;
        s_blk.npacks = 2 * n + 6;
        s_blk.niter = p.prologue + ((2 * n + 2 + Q - 1) >> QL2) + RS_LAG;
        s_blk.src_rep = src + (size_t)rep * 2 * p.lenX * p.nk;
        s_blk.dst16 = reinterpret_cast<uint4 *>(dst + (size_t)rep * 2 * p.lenX * p.nk);
        for (int s = 0; s < RS_STAGES; ++s) rs_mbar_init(&s_full[s], 1);
        rs_mbar_init(&s_done[0], NWARPS);
        rs_mbar_init(&s_done[1], NWARPS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    Plan<MODEL> plan;
    fetch_plan(L, rep, plan);

    // rows of packing group g -> staging slot g % RS_STAGES (one warp; lane = threadIdx.x & 31)
    auto issue_group = [&](int g) {
        const int first = g << QL2, npacks = s_blk.npacks, lane = t & 31;
        if (first >= npacks) return;
        const int nvalid = npacks - first < Q ? npacks - first : Q;
        uint64_t *bar = &s_full[g % RS_STAGES];
        if (lane == 0) rs_mbar_expect_tx(bar, (uint32_t)nvalid * (uint32_t)p.nk);
        __syncwarp();
        unsigned char *slot = rs_smem + (size_t)(g % RS_STAGES) * RS_THREADS * 64;
        for (int k = lane; k < nvalid; k += 32) {
            const RSTask task = rs_pack_task(p, s_blk.a, s_blk.n, first + k);
            const int colour = (task.info >> 16) & 1;
            rs_bulk_load(slot + (size_t)k * p.nk, s_blk.src_rep + ((size_t)colour * p.lenX + task.row) * p.nk, (uint32_t)p.nk, bar);
        }
    };
    // tables of iteration `it`, entry k (one lane per row task)
    auto build_entry = [&](int it, int k) {
        const int kc = it - p.prologue, a = s_blk.a, n = s_blk.n, at = ((it % RS_TABLES) << QL2) + k;
        s_ct[at] = rs_build_compute(p, a, n, (kc << QL2) + k);
        s_pt[at] = rs_build_pack(p, n, (it << QL2) + k);
        s_ut[at] = rs_build_unpack(p, a, n, ((kc - RS_LAG) << QL2) + k);
    };
    if (t < 2 * Q) build_entry(t >> QL2, t & (Q - 1));  // iterations 0 and 1
    __syncthreads();
    if (warp == 0)
        for (int g = 0; g < RS_STAGES; ++g) issue_group(g);

    int dB = 0, dS = 0;
    const uint32_t greplica = (uint32_t)(L.rep_offset + rep);
    auto wait_on = [&](uint64_t *bar, uint32_t parity) {
        if (rs_mbar_try_wait(bar, parity)) return;
        const long long t0 = clock64();
        while (!rs_mbar_try_wait(bar, parity)) {
            if (clock64() - t0 > 4000000000LL) {  // about two seconds: a copy never arrived / a warp never did
                *err = 1;
                break;
            }
        }
    };
    const int niter = s_blk.niter, prologue = p.prologue;
#pragma unroll 1
    for (int it = 0; it < niter; ++it) {
        const int tb = (it % RS_TABLES) << QL2, kc = it - prologue;
        if (it >= 2) {
            // every warp has finished iteration it - 2: what it computed and packed may be read, what it read may be
            // overwritten, the staging slot it packed from may be refilled, the tables built during it are complete
            wait_on(&s_done[it & 1], (uint32_t)((it - 2) >> 1) & 1u);
            if (warp == (it + NWARPS / 2) % NWARPS) issue_group(it);
        }
        if (kc >= 0) {
            rs_unpack<MODEL, NWL2>(s_ut + tb, W, s_blk.dst16, t);
            if ((kc << QL2) < s_blk.ntasks) rs_compute<MODEL, NWL2>(s_ct + tb, W, plan, greplica, sweep, L.key, t, dB, dS);
        }
        if ((it << QL2) < s_blk.npacks) {
            wait_on(&s_full[it % RS_STAGES], (uint32_t)(it / RS_STAGES) & 1u);
            rs_pack<MODEL, NWL2>(s_pt + tb, stage + (size_t)(it % RS_STAGES) * RS_THREADS * 4, W, t);
        }
        {  // the tables of iteration it + 2: a different warp (two for Q = 64) every time, so that no warp is always behind
            const int k = (((warp + NWARPS - it % NWARPS) % NWARPS) << 5) + (t & 31);
            if (k < Q) build_entry(it + 2, k);
        }
        __syncwarp();
        if ((t & 31) == 0) rs_mbar_arrive(&s_done[it & 1]);
    }
    __syncthreads();

    // tracked observables: block sum, one atomic pair
    dB = __reduce_add_sync(0xFFFFFFFFu, dB);
    dS = __reduce_add_sync(0xFFFFFFFFu, dS);
    if ((t & 31) == 0) {
        s_red[0][t >> 5] = dB;
        s_red[1][t >> 5] = dS;
    }
    __syncthreads();
    if (t < 2) {
        int v = 0;
#pragma unroll
        for (int k = 0; k < NWARPS; ++k) v += s_red[t][k];
        if (v) atomicAdd(L.tracked + 2 * (size_t)rep + t, (unsigned long long)(long long)v);
    }
}

#endif  // __CUDACC__

}  // namespace pmc
