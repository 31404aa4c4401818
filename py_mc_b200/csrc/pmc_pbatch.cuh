// pmc_pbatch.cuh -- K12: ALL replicas of a handle resident in the shared memory of the whole GPU for all sweeps
// (and all replica exchanges) of a call.
//
// What it replaces: the run loop of the reference -- MCstep after MCstep with an exchange every
// RXtrial_frequency steps (py_mc/mc_mpi.py:165-169, Xtrial :102-140) -- for a whole batch of replicas, or
// CanonicalMonteCarlo.run's loop (py_mc/mc.py:198-199) for one large lattice.
//
// The rows of all replicas of the handle form one list of nrep * lenX rows; it is cut into B contiguous pieces
// of (almost) equal length, one per block of a cooperative launch (B = number of SMs), whatever the replica
// boundaries are.  A block loads its rows once (int8 -> bits, or bits as stored), keeps them in shared memory,
// runs every sweep of the call on them and stores them once.  256 replicas of 1024 x 1024 are 32 MiB of bits,
// exactly as much as one 16384 x 16384 lattice: both fill the 148 x 227 KiB of a B200 to 97 %, and 32 or 8
// replicas (replica exchange spread over 8 GPUs) use all SMs just the same, which a one-block-per-replica
// kernel cannot.
//
// Rows a block needs from other blocks (at most four per colour: the row above its first row, the row below its
// last row, and -- because lattices are periodic -- row 0 / row lenX-1 of a replica that starts / ends in
// another block) travel through a small exchange buffer in global memory.  There is NO grid barrier in a sweep and
// no flag either: every 64-bit word of an exported row travels as two 8-byte packets, each holding 32 bits of the
// word and the number of the pass that produced it (the low-latency protocol of collective libraries: the flag
// rides in the same atomic store as the data, so the writer needs no fence and the reader a single load):
//   * the rows other blocks need ("export rows") are updated FIRST in every colour pass and stored to the exchange
//     buffer by the threads that computed them -- while the rest of the block is already working on the other rows;
//   * the only readers of remote rows are those same first items of the NEXT pass; each loads the packet pair of the
//     word it needs from L2 and repeats the load until both packets carry that pass number (they do already,
//     unless the other block lags a whole pass behind).
// So the neighbour exchange is fully overlapped with the update and costs no block-wide synchronisation beyond
// the one __syncthreads that ends a colour pass.  (The single-lattice kernel K11, pmc_persist.cuh, meets at a
// grid barrier twice per sweep: measured 2.9 us per pass, 5.5 % of a 16384^2 sweep and half of a 1024^2 one.  The
// first form of this kernel published a per-block flag after a fence and a count of the exporting warps: three
// dependent trips to L2 per hand-shake, about 3 us per pass once a block holds only a few rows.)
// A word is never overwritten before it has been read: the thread that stores word w of an export row in pass p+2
// has, in that same item, loaded word w of the neighbour's row from pass p+1, which the neighbour's thread stored
// after loading my word w of pass p.
//
// Replica exchange inside the kernel (one GPU holds all replicas): after the sweeps of an exchange interval the
// blocks recompute the exact bond sum and magnetisation of their rows per replica, publish them, meet at ONE grid
// barrier, and every block evaluates the same pair decisions as rx_decide_kernel (pmc_rx.cuh) from the same
// inputs; block 0 also writes the new replica <-> temperature maps (double-buffered by exchange parity).  The
// lattices never leave the chip between exchanges.
//
// Observables: the update loop keeps no running sums at all (the per-word popcounts of the other kernels are
// gone); the exact sums are recomputed from the resident lattice where they are needed -- at an exchange and at
// the end of the launch.
//
// Random-stream indexing is that of every other sweep kernel (word = row-in-replica * nw + w, replica, step), so
// the trajectories are identical (tests/test_gpu_pbatch.py; the per-thread pieces below also run on the host,
// block after block, in tests/test_hostemu_core.py, which checks the geometry and the exchange-buffer plumbing).
#pragma once
#include "pmc_core.cuh"
#ifdef __CUDACC__
#include "pmc_tiled_i8.cuh"  // pack16 / unpack16
#endif

#ifndef PMC_WARP
#define PMC_WARP 32  // the host shim compiles with 1: a "warp" of one thread
#endif

namespace pmc {

constexpr int PB_MAXSEG = 8;  // replicas a block may hold rows of
constexpr int PB_MAXX = 4;    // rows a block takes from / gives to other blocks, per colour
constexpr int PB_MAX_THREADS = 1024;
constexpr int PB_MAXRANKS = 16;  // GPUs one ladder of replicas may be spread over

// ---- geometry of block b (pure function of b, B, nrep, lenX) -------------------------------------------------
struct PBGeom {
    int R0, nrows;             // first row (in the list of all rows of all replicas) and number of rows
    int rep_first, nseg;       // first replica with rows here; number of replicas with rows here
    int i0;                    // row-in-replica of local row 0
    int nimp;                  // rows taken from other blocks
    int imp_slot[PB_MAXX];     // ... their slot in the exchange buffer
    int imp_blk[PB_MAXX];      // ... the block that writes it
    int up0, dn1, h2, h3;      // import index of: row above local row 0, row below the last local row, row 0 of the
                               // first replica (it began in an earlier block), last row of the last replica; or -1
    int nexp;                  // (local row -> slot) pairs given to other blocks
    int exp_row[PB_MAXX], exp_slot[PB_MAXX];
    int nxrow;                 // distinct export rows, ascending
    int xrow[PB_MAXX];
};

__host__ __device__ __forceinline__ int pb_row0(int b, long long NR, int B) { return (int)(((long long)b * NR) / B); }
__host__ __device__ __forceinline__ int pb_block_of_row(long long g, long long NR, int B) {
    return (int)(((g + 1) * B - 1) / NR);
}
// exchange-buffer slots: first / last row of block b, row 0 / row lenX-1 of replica r
__host__ __device__ __forceinline__ int pb_bslot(int b, int last) { return 2 * b + last; }
__host__ __device__ __forceinline__ int pb_rslot(int B, int r, int last) { return 2 * B + 2 * r + last; }
__host__ __device__ __forceinline__ int pb_nslots(int B, int nrep) { return 2 * B + 2 * nrep; }

__host__ __device__ inline void pb_geometry(int b, int B, int nrep, int lenX, PBGeom &g) {
    const long long NR = (long long)nrep * lenX;
    g.R0 = pb_row0(b, NR, B);
    const int R1 = pb_row0(b + 1, NR, B);
    g.nrows = R1 - g.R0;
    g.rep_first = g.R0 / lenX;
    const int rep_last = (R1 - 1) / lenX;
    g.nseg = rep_last - g.rep_first + 1;
    g.i0 = g.R0 - g.rep_first * lenX;
    const int i1 = (R1 - 1) - rep_last * lenX;
    g.nimp = g.nexp = g.nxrow = 0;
    g.up0 = g.dn1 = g.h2 = g.h3 = -1;
    for (int k = 0; k < PB_MAXX; ++k) g.imp_slot[k] = g.imp_blk[k] = g.exp_row[k] = g.exp_slot[k] = g.xrow[k] = 0;
    const long long first_last = (long long)g.rep_first * lenX + lenX - 1;  // last row of the first replica
    const long long last_first = (long long)rep_last * lenX;                // row 0 of the last replica
#define PB_IMP(slot, blk) (g.imp_slot[g.nimp] = (slot), g.imp_blk[g.nimp] = (blk), g.nimp++)
#define PB_EXP(row, slot) (g.exp_row[g.nexp] = (row), g.exp_slot[g.nexp] = (slot), g.nexp++)
    // local row 0: its upper neighbour
    if (g.i0 > 0) {  // ... is the last row of the block before
        g.up0 = PB_IMP(pb_bslot(b - 1, 1), b - 1);
        PB_EXP(0, pb_bslot(b, 0));
    } else if (first_last >= R1) {  // ... is the last row of this replica, which lives in a later block
        g.up0 = PB_IMP(pb_rslot(B, g.rep_first, 1), pb_block_of_row(first_last, NR, B));
        PB_EXP(0, pb_rslot(B, g.rep_first, 0));
    }
    // last local row: its lower neighbour
    if (i1 < lenX - 1) {
        g.dn1 = PB_IMP(pb_bslot(b + 1, 0), b + 1);
        PB_EXP(g.nrows - 1, pb_bslot(b, 1));
    } else if (last_first < g.R0) {  // row 0 of this replica, in an earlier block
        g.dn1 = PB_IMP(pb_rslot(B, rep_last, 0), pb_block_of_row(last_first, NR, B));
        PB_EXP(g.nrows - 1, pb_rslot(B, rep_last, 1));
    }
    // the first replica began in an earlier block and ends inside this one (not on my last row: that is the case above)
    if (g.i0 > 0 && first_last < R1 - 1) {
        g.h2 = PB_IMP(pb_rslot(B, g.rep_first, 0), pb_block_of_row((long long)g.rep_first * lenX, NR, B));
        PB_EXP((int)(first_last - g.R0), pb_rslot(B, g.rep_first, 1));
    }
    // the last replica begins inside this block (not on my first row) and ends in a later one
    if (i1 < lenX - 1 && last_first > g.R0) {
        g.h3 = PB_IMP(pb_rslot(B, rep_last, 1), pb_block_of_row(last_first + lenX - 1, NR, B));
        PB_EXP((int)(last_first - g.R0), pb_rslot(B, rep_last, 0));
    }
#undef PB_IMP
#undef PB_EXP
    // distinct export rows, ascending (insertion into a list of at most four)
    for (int e = 0; e < g.nexp; ++e) {
        const int r = g.exp_row[e];
        int pos = 0;
        bool dup = false;
        for (int k = 0; k < g.nxrow; ++k) {
            if (g.xrow[k] == r) dup = true;
            if (g.xrow[k] < r) pos = k + 1;
        }
        if (dup) continue;
        for (int k = g.nxrow; k > pos; --k) g.xrow[k] = g.xrow[k - 1];
        g.xrow[pos] = r;
        ++g.nxrow;
    }
}

// One 64-bit word in the exchange buffer: two 8-byte packets {32 bits of the word, pass number}.
struct alignas(16) PBPacket {
    uint64_t lo, hi;
};

// What a thread needs besides the geometry: everything that sits in shared or global memory.
struct PBCtx {
    uint64_t *s;                 // [2 colours][nrows][nw] resident rows of this block
    const int *segoff;           // [nseg] row-in-replica minus local row, per replica of this block
    const int *segstart;         // [nseg] first local row of each replica of this block (segstart[0] = 0)
    const uint32_t *splan;       // [nseg][PLAN_WORDS] plan row of each replica (its current temperature)
    PBPacket *xbuf;              // [2 colours][nslots][nw] exchange buffer (global); pass numbers only ever grow
    PBPacket *xbuf_up, *xbuf_dn; // where row 0 / the last row of the lattice is exported to: this buffer, or -- one lattice
                                 // in row slabs over several GPUs -- the exchange buffer of the GPU above / below (peer memory)
    unsigned int epoch;          // added to every pass number: launches never see each other's packets
    int row0;                    // row of the whole lattice of local row 0 of replica 0 (random-stream index; slabs)
    long long wait_cycles;       // how long a reader waits for a packet before it gives up (longer when other GPUs feed me)
    unsigned int *abort_word;    // global: set when a wait has timed out
    int *err;
    int lenX, nw, nw_shift;      // nw_shift >= 0: nw is that power of two
    int nslots, b, nblocks;
    uint32_t rep_offset;
    const PhiloxKey *key;        // round keys (in the kernel parameters: constant memory)
};

struct SmemRow {
    const uint32_t *p;
    __device__ __forceinline__ uint32_t operator[](int k) const { return p[k]; }
};

// Store word v of an export row as the packet pair of pass `ev` (each 8-byte half is one atomic store).
__device__ __forceinline__ void pb_put(PBPacket *p, uint64_t v, unsigned int ev) {
    const uint64_t a = (v & 0xFFFFFFFFull) | ((uint64_t)ev << 32), b = (v >> 32) | ((uint64_t)ev << 32);
#ifdef __CUDA_ARCH__
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(a), "l"(b) : "memory");
#else
    p->lo = a;
    p->hi = b;
#endif
}

// Load the word whose packets carry pass number `need` (or a later one: see the header).  The poll backs off and is
// bounded: after about two seconds (ten when the rows of another GPU are awaited) *err is raised, and so is the abort word
// in device memory that lets every other waiter of the launch give up at once.
__device__ __forceinline__ uint64_t pb_get(const PBPacket *p, unsigned int need, int *err, unsigned int *abort_word,
                                           long long wait_cycles) {
#ifdef __CUDA_ARCH__
    uint64_t a, b;
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
    if ((int)((unsigned int)(a >> 32) - need) < 0 || (int)((unsigned int)(b >> 32) - need) < 0) {
        const long long t0 = clock64();
        unsigned int spins = 0;
        for (;;) {
            __nanosleep(20);
            asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
            if ((int)((unsigned int)(a >> 32) - need) >= 0 && (int)((unsigned int)(b >> 32) - need) >= 0) break;
            if ((++spins & 1023u) == 0) {
                if (*(volatile unsigned int *)abort_word) break;
                if (clock64() - t0 > wait_cycles) {
                    *(volatile unsigned int *)abort_word = 1u;
                    *err = 1;
                    break;
                }
            }
        }
    }
    return (a & 0xFFFFFFFFull) | (b << 32);
#else
    // the host drives the blocks in lock step: the packets must be there already
    if ((int)((unsigned int)(p->lo >> 32) - need) < 0 || (int)((unsigned int)(p->hi >> 32) - need) < 0) *err = 1;
    (void)abort_word;
    (void)wait_cycles;
    return (p->lo & 0xFFFFFFFFull) | (p->hi << 32);
#endif
}

// Acquire / release at GPU scope without the full (sequentially consistent) __threadfence (grid barrier of the
// exchange step).
__device__ __forceinline__ unsigned int pb_ld_acquire(const unsigned int *p) {
#ifdef __CUDA_ARCH__
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
#else
    return *p;
#endif
}
__device__ __forceinline__ void pb_fence() {
#ifdef __CUDA_ARCH__
    asm volatile("fence.acq_rel.gpu;" ::: "memory");
#endif
}

// local row, word column, replica segment and row-in-replica of ordinal `o` when the export rows come first
__device__ __forceinline__ void pb_locate(const PBGeom &g, const PBCtx &c, int o, bool exports_first, int &lr, int &w, int &seg,
                                          int &i) {
    int ro;
    if (c.nw_shift >= 0) {
        ro = o >> c.nw_shift;
        w = o & (c.nw - 1);
    } else {
        ro = o / c.nw;
        w = o - ro * c.nw;
    }
    if (!exports_first) {
        lr = ro;
    } else if (ro < g.nxrow) {
        lr = ro == 0 ? g.xrow[0] : (ro == 1 ? g.xrow[1] : (ro == 2 ? g.xrow[2] : g.xrow[3]));
    } else {
        lr = ro - g.nxrow;
#pragma unroll
        for (int k = 0; k < PB_MAXX; ++k)
            if (k < g.nxrow && lr >= g.xrow[k]) ++lr;
    }
    seg = 0;
    for (int k = 1; k < g.nseg; ++k) seg += lr >= c.segstart[k];
    i = lr + c.segoff[seg];
}

// rows above / below local row lr (row-in-replica i): a local row (>= 0) or -1 - import index
__device__ __forceinline__ void pb_neighbours(const PBGeom &g, int lenX, int lr, int i, int &up, int &dn) {
    up = lr - 1;
    dn = lr + 1;
    if (i == 0) {
        const int G = lr + lenX - 1;  // the replica's last row
        up = G < g.nrows ? G : -1 - (lr == 0 ? g.up0 : g.h3);
    } else if (lr == 0) {
        up = -1 - g.up0;
    }
    if (i == lenX - 1) {
        const int G = lr - (lenX - 1);  // the replica's row 0
        dn = G >= 0 ? G : -1 - (lr == g.nrows - 1 ? g.dn1 : g.h2);
    } else if (lr == g.nrows - 1) {
        dn = -1 - g.dn1;
    }
}

// word w of the other colour in the row above / below: shared memory, or the exchange buffer once its writer is there
__device__ __forceinline__ uint64_t pb_row_word(const PBGeom &g, const PBCtx &c, int oc, int row, int w, unsigned int need) {
    if (row >= 0) return c.s[((size_t)oc * g.nrows + row) * c.nw + w];
    const int k = -1 - row;
    return pb_get(c.xbuf + ((size_t)oc * c.nslots + g.imp_slot[k]) * c.nw + w, c.epoch + need, c.err, c.abort_word, c.wait_cycles);
}

// What a thread works out about its items that does not change from pass to pass: the first-round item (general path)
// and where its later rounds start.  Kept in shared memory when there is room (32 bytes per thread: every launch shape except
// the ones that fill the shared memory with rows, where a pass has so many rounds that the set-up does not matter); with a
// few rows per block a pass is two or three rounds long and this set-up was a tenth of its instructions.
struct PBI4 {
    int x, y, z, w;
};
struct PBItems {
    int idx, i, up, dn;     // first round: word index in the block (lr * nw + w), row in the replica, rows above / below
    int lr, w, seg, on;     // ... its local row, word column, replica segment; on = 0: no item (thread beyond the items)
    int lr2, w2, xk2;       // later rounds: local row and word column of the second item, export rows already skipped
};

__device__ __forceinline__ void pb_items_first(const PBGeom &g, const PBCtx &c, int t, PBItems &it) {
    const int nw = c.nw, items = g.nrows * nw;
    it.on = t < items;
    const int o = it.on ? t : items - 1;
    pb_locate(g, c, o, true, it.lr, it.w, it.seg, it.i);
    pb_neighbours(g, c.lenX, it.lr, it.i, it.up, it.dn);
    it.idx = it.lr * nw + it.w;
}
__device__ __forceinline__ void pb_items_later(const PBGeom &g, const PBCtx &c, int t, int T, PBItems &it) {
    const int nw = c.nw, ro = (T + t) / nw;
    it.w2 = (T + t) - ro * nw;
    it.lr2 = ro - g.nxrow;  // rank among the rows that are not export rows -> row: skip over the export rows
    it.xk2 = 0;
    while (it.xk2 < g.nxrow && it.lr2 >= g.xrow[it.xk2]) {
        ++it.lr2;
        ++it.xk2;
    }
}
__device__ __forceinline__ void pb_items(const PBGeom &g, const PBCtx &c, int t, int T, PBItems &it) {
    pb_items_first(g, c, t, it);
    pb_items_later(g, c, t, T, it);
}
__device__ __forceinline__ void pb_items_store(const PBItems &it, PBI4 *ca, PBI4 *cb, int t) {
    ca[t] = PBI4{it.idx, it.i, it.up, it.dn};
    cb[t] = PBI4{it.lr | (it.on << 30), it.w | (it.seg << 16), it.lr2, it.w2 | (it.xk2 << 16)};
}
__device__ __forceinline__ void pb_items_load(PBItems &it, const PBI4 *ca, const PBI4 *cb, int t) {
    const PBI4 a = ca[t], b = cb[t];
    it.idx = a.x;
    it.i = a.y;
    it.up = a.z;
    it.dn = a.w;
    it.lr = b.x & 0x3FFFFFFF;
    it.on = (b.x >> 30) & 1;
    it.w = b.y & 0xFFFF;
    it.seg = b.y >> 16;
    it.lr2 = b.z;
    it.w2 = b.w & 0xFFFF;
    it.xk2 = b.w >> 16;
}

// Where the word of export slot `slot` goes: the slots of a replica's row 0 / last row (what a block at the other end of
// the replica reads) are those a neighbouring GPU reads when one lattice is cut into row slabs.
__device__ __forceinline__ PBPacket *pb_export_buffer(const PBCtx &c, int slot, int B) {
    return slot < 2 * B ? c.xbuf : (((slot - 2 * B) & 1) ? c.xbuf_dn : c.xbuf_up);
}

// One colour pass of one thread.  `events` = 1 + number of passes since the load: the pass number the packets of
// the rows this pass reads must carry; what it exports carries events + 1.  Every thread of a warp must call this
// together (update_word votes).
//
// Items are taken export rows first, then the other rows in their natural order: ordinal o = base + t, ordinal row
// o / nw, column o % nw.  The first round (ordinals 0 .. T-1) holds all items of the export rows (the host makes T at
// least that large): it takes the general path -- rows of other blocks from the exchange buffer, new words to the
// exchange buffer.  Every later round touches only rows whose neighbours are all in this block, and a thread
// walks down the rows with a fixed stride: its row, replica segment and plan are carried from one item to the next
// instead of being recomputed.  (Measured alternatives, slower: recomputing the row from the ordinal per item; the
// natural order with the export items pulled forward by their owners -- with few words per row those owners' warps do
// a whole extra item.)
template <int MODEL, bool KEEP, bool SLAB>
__device__ __forceinline__ void pb_pass(const PBGeom &g, const PBCtx &c, int colour, uint64_t step, unsigned int events, int t,
                                        int T, Plan<MODEL> &kept_plan, int &kept_cur, const PBI4 *ca, const PBI4 *cb) {
    // block-uniform values out of shared memory once: the stores of the loop could alias them for all the compiler knows
    const int nrows = g.nrows, nxrow = g.nxrow, nseg = g.nseg, nw = c.nw, lenX = c.lenX;
    const int oc = colour ^ 1, items = nrows * nw, nexport = nxrow * nw;
    uint64_t *const sc = c.s + (size_t)colour * items;
    const uint64_t *const so = c.s + (size_t)oc * items;
    const uint32_t rep0 = c.rep_offset + (uint32_t)g.rep_first;
    // KEEP (launch shapes with room in shared memory: few rows per block, two or three rounds per pass): the per-thread
    // item data comes from shared memory, and the Ising plan (9 registers) with the replica segment it belongs to (-1:
    // none) lives across passes in the caller (the lattice-gas plan, 25 registers, would spill there).  Otherwise everything
    // is worked out per pass, the later rounds' starting point only after the first round (registers).
    Plan<MODEL> pass_plan;
    int pass_cur = -1;
    Plan<MODEL> &plan = (KEEP && MODEL == MODEL_ISING) ? kept_plan : pass_plan;
    int &cur = (KEEP && MODEL == MODEL_ISING) ? kept_cur : pass_cur;
    PBItems f;
    if (KEEP)
        pb_items_load(f, ca, cb, t);
    else
        pb_items_first(g, c, t, f);

    // ---- first round: general path -----------------------------------------------------------------------
    {
        const bool on = f.on;
        const int lr = f.lr, w = f.w, seg = f.seg, i = f.i, up = f.up, dn = f.dn, idx = f.idx;
        if (seg != cur) {
            plan.init(SmemRow{c.splan + seg * PLAN_WORDS});
            cur = seg;
        }
        const uint64_t S = sc[idx], C = so[idx];
        uint64_t Sd;
        if ((i + colour) & 1)
            Sd = (C >> 1) | (so[w == nw - 1 ? idx - w : idx + 1] << 63);
        else
            Sd = (C << 1) | (so[w == 0 ? idx + nw - 1 : idx - 1] >> 63);
        const uint64_t U = pb_row_word(g, c, oc, up, w, events), D = pb_row_word(g, c, oc, dn, w, events);
        int dB, dS;  // unused: the sums are recomputed where they are needed (pb_reduce)
        const uint64_t Snew = update_word<MODEL>(S, U, D, C, Sd, on ? ~0ull : 0ull, plan, (uint32_t)((SLAB ? i + c.row0 : i) * nw + w),
                                                 rep0 + (uint32_t)seg, step, *c.key, dB, dS);
        if (on) {
            sc[idx] = Snew;
            if (t < nexport)
                for (int e = 0; e < g.nexp; ++e)
                    if (g.exp_row[e] == lr)
                        pb_put((SLAB ? pb_export_buffer(c, g.exp_slot[e], c.nblocks) : c.xbuf) + ((size_t)colour * c.nslots + g.exp_slot[e]) * nw + w,
                               Snew, c.epoch + events + 1u);
        }
    }
    if (T >= items) return;

    // ---- later rounds: all neighbours local, incremental indexing -----------------------------------------------
    const bool strided = T % nw == 0;  // then a thread keeps its word column and moves down T / nw ordinal rows per round
    const int dr = strided ? T / nw : 0;
    const int wrap = (lenX - 1) * nw;  // from row 0 of a replica to its last row
    if (!KEEP) pb_items_later(g, c, t, T, f);
    int lr = f.lr2, w = f.w2, xk = f.xk2;
    int xn = xk < nxrow ? g.xrow[xk] : 0x7FFFFFFF;  // the next export row at or below which lr has not yet been moved past
    int left = w == 0 ? nw - 1 : -1, right = w == nw - 1 ? 1 - nw : 1;
    int seg = 0, seg_end = nseg > 1 ? c.segstart[1] : 0x7FFFFFFF, ioff = c.segoff[0];
#pragma unroll 1
    for (int base = T; base < items; base += T) {  // whole warps stay in the loop: the update votes
        const bool on = base + t < items;
        if (!strided && on) {
            const int ro = (base + t) / nw;
            w = (base + t) - ro * nw;
            left = w == 0 ? nw - 1 : -1;
            right = w == nw - 1 ? 1 - nw : 1;
            lr = ro - nxrow;
            for (xk = 0; xk < nxrow && lr >= g.xrow[xk]; ++xk) ++lr;
        }
        while (on && lr >= seg_end) {  // next replica (rows only ever increase within a thread)
            ++seg;
            ioff = c.segoff[seg];
            seg_end = seg + 1 < nseg ? c.segstart[seg + 1] : 0x7FFFFFFF;
        }
        if (on && seg != cur) {
            plan.init(SmemRow{c.splan + seg * PLAN_WORDS});
            cur = seg;
        }
        const int i = lr + ioff;
        const int o = on ? lr * nw + w : 0;  // idle lanes of the last round read word 0 of the block: always there
        const int ou = !on ? 0 : (i == 0 ? o + wrap : o - nw), od = !on ? 0 : (i == lenX - 1 ? o - wrap : o + nw);
        const uint64_t S = sc[o], C = so[o];
        uint64_t Sd;
        if ((i + colour) & 1)
            Sd = (C >> 1) | (so[on ? o + right : 0] << 63);
        else
            Sd = (C << 1) | (so[on ? o + left : 0] >> 63);
        int dB, dS;
        const uint64_t Snew = update_word<MODEL>(S, so[ou], so[od], C, Sd, on ? ~0ull : 0ull, plan,
                                                 (uint32_t)((SLAB ? i + c.row0 : i) * nw + w), rep0 + (uint32_t)seg, step, *c.key, dB, dS);
        if (on) sc[o] = Snew;
        if (strided) {
            lr += dr;
            while (lr >= xn) {  // moved past an export row: those were done in the first round
                ++lr;
                ++xk;
                xn = xk < nxrow ? g.xrow[xk] : 0x7FFFFFFF;
            }
        }
    }
}

// After the load: the export rows of both colours go to the exchange buffer as pass 1.
__device__ __forceinline__ void pb_publish_loaded(const PBGeom &g, const PBCtx &c, int t, int T) {
    for (int k = t; k < 2 * g.nexp * c.nw; k += T) {
        const int colour = k / (g.nexp * c.nw), r = k - colour * g.nexp * c.nw;
        const int e = r / c.nw, w = r - e * c.nw;
        pb_put(pb_export_buffer(c, g.exp_slot[e], c.nblocks) + ((size_t)colour * c.nslots + g.exp_slot[e]) * c.nw + w,
               c.s[((size_t)colour * g.nrows + g.exp_row[e]) * c.nw + w], c.epoch + 1u);
    }
}

// Exact bond sum and second observable (magnetisation / particle number) of the rows of this block, per replica,
// added into acc[seg][2] (shared memory, zeroed by the caller).  Every bond joins a site of colour 0 and a site of
// colour 1, so one walk over the words of colour 0 with their four neighbour words counts every bond once (the same
// integers as reduce_kernel, which walks both colours with two neighbours each); the same-row neighbour word is also
// the only visit of that colour-1 word, so the second observable is taken from both.  Warps take contiguous runs of
// items so that a warp changes replica at most nseg - 1 times.  `events` as in pb_pass: the rows of colour 1 that
// other blocks hold must be those of the last pass.
template <int MODEL>
__device__ __forceinline__ void pb_reduce(const PBGeom &g, const PBCtx &c, unsigned int events, long long *acc, int t, int T) {
    const int nw = c.nw, items = g.nrows * nw;
    const int nwarps = T / PMC_WARP, wid = t / PMC_WARP, lane = t % PMC_WARP;
    const int per_warp = ((items + nwarps - 1) / nwarps + PMC_WARP - 1) / PMC_WARP * PMC_WARP;
    const int lo = wid * per_warp, hi = lo + per_warp < items ? lo + per_warp : items;
    const uint64_t *const s0 = c.s, *const s1 = c.s + (size_t)items;
    int aB = 0, aS = 0, cur = 0;
    auto flush = [&]() {
#ifdef __CUDA_ARCH__
        const int b = __reduce_add_sync(0xFFFFFFFFu, aB), s = __reduce_add_sync(0xFFFFFFFFu, aS);
        if (lane == 0) {
            if (b) atomicAdd((unsigned long long *)&acc[2 * cur], (unsigned long long)(long long)b);
            if (s) atomicAdd((unsigned long long *)&acc[2 * cur + 1], (unsigned long long)(long long)s);
        }
#else
        acc[2 * cur] += aB;
        acc[2 * cur + 1] += aS;
#endif
        aB = aS = 0;
    };
#pragma unroll 1
    for (int base = lo; base < hi; base += PMC_WARP) {
        const bool on = base + lane < hi;
        const int o = on ? base + lane : hi - 1;
        int lr, w, seg, i, up, dn;
        pb_locate(g, c, o, false, lr, w, seg, i);
        pb_neighbours(g, c.lenX, lr, i, up, dn);
        int dB = 0, dS = 0;
        if (on) {
            const int idx = lr * nw + w;
            const uint64_t S = s0[idx], C = s1[idx];
            const uint64_t U = pb_row_word(g, c, 1, up, w, events), D = pb_row_word(g, c, 1, dn, w, events);
            uint64_t Sd;  // the other same-row neighbour: site k+1 (rows where colour 0 starts at column 1) or k-1
            if (i & 1)
                Sd = (C >> 1) | (s1[w == nw - 1 ? idx - w : idx + 1] << 63);
            else
                Sd = (C << 1) | (s1[w == 0 ? idx + nw - 1 : idx - 1] >> 63);
            if (MODEL == MODEL_ISING) {
                dB = 256 - 2 * (__popcll(S ^ U) + __popcll(S ^ D) + __popcll(S ^ C) + __popcll(S ^ Sd));
                dS = 2 * (__popcll(S) + __popcll(C)) - 128;
            } else {
                dB = __popcll(S & U) + __popcll(S & D) + __popcll(S & C) + __popcll(S & Sd);
                dS = __popcll(S) + __popcll(C);
            }
        }
#ifdef __CUDA_ARCH__
        const int seg0 = __shfl_sync(0xFFFFFFFFu, seg, 0);
        const bool uniform = __all_sync(0xFFFFFFFFu, seg == seg0);
#else
        const int seg0 = seg;
        const bool uniform = true;
#endif
        if (uniform) {
            if (seg0 != cur) {
                flush();
                cur = seg0;
            }
            aB += dB;
            aS += dS;
        } else {  // the 32 items straddle a replica boundary: every lane adds its own
            flush();
#ifdef __CUDA_ARCH__
            if (dB) atomicAdd((unsigned long long *)&acc[2 * seg], (unsigned long long)(long long)dB);
            if (dS) atomicAdd((unsigned long long *)&acc[2 * seg + 1], (unsigned long long)(long long)dS);
#endif
        }
    }
    flush();
}

// ---- replica exchange ----------------------------------------------------------------------------------------
// esum[global replica][2] = exact bond sum and second observable of every replica at this exchange.  The energy is
// evaluated like rx_energy_kernel (no FMA contraction) so that it equals the host expression J*B + (Eads-mu)*N bit
// for bit.
// The sums of a replica on another GPU arrive as packets (xpk, pass number = exchange round + 1: what pb_get waits for);
// xpk == nullptr: all replicas are here.
struct PBRemote {
    const PBPacket *xpk;  // [R][2] of this round's parity, or nullptr
    unsigned int need;
    int own_lo, own_hi;   // my replicas: [own_lo, own_hi)
    int *err;
    unsigned int *abort_word;
};
__device__ __forceinline__ double pb_energy(const long long *esum, const PBRemote &rm, int r, double c_bond, double c_second) {
#ifdef __CUDA_ARCH__
    long long Bs, Ss;
    if (rm.xpk && (r < rm.own_lo || r >= rm.own_hi)) {
        Bs = (long long)pb_get(rm.xpk + 2 * r, rm.need, rm.err, rm.abort_word, 20000000000LL);  // ten seconds: another GPU
        Ss = (long long)pb_get(rm.xpk + 2 * r + 1, rm.need, rm.err, rm.abort_word, 20000000000LL);
    } else {
        Bs = __ldcg(esum + 2 * r);
        Ss = __ldcg(esum + 2 * r + 1);
    }
    double e = __dmul_rn(c_bond, (double)Bs);
    if (c_second != 0.0) e = __dadd_rn(e, __dmul_rn((double)Ss, c_second));
#else
    (void)rm;
    const long long Bs = esum[2 * r], Ss = esum[2 * r + 1];
    double e = c_bond * (double)Bs;
    if (c_second != 0.0) e = e + (double)Ss * c_second;
#endif
    return e;
}

#ifdef __CUDACC__

struct PBatchArgs {
    long long nsweeps;
    unsigned long long sweep0;     // the handle's sweep counter at entry
    int xchg_every;                // exchange after every xchg_every sweeps; 0 = never
    int xchg_phase;                // sweeps already done since the last exchange
    int scheme0;                   // XCscheme of the first exchange of this launch (alternates from there)
    unsigned long long round0;     // the handle's exchange round at entry
    unsigned int *bar;             // [2] grid barrier arrivals and the abort word, zeroed by the host before the launch
    int *err;
    PBPacket *xbuf;                // [2][nslots][nw]; zeroed once, pass numbers = epoch + 1, 2, ... grow from launch to launch
    PBPacket *xbuf_up, *xbuf_dn;   // = xbuf, or the exchange buffers of the GPUs holding the slab above / below (CUDA IPC mappings)
    unsigned int epoch;            // first pass number of this launch minus 1: the same on every GPU of a slab run
    int row0;                      // slab: row of the whole lattice of my first owned row (0 otherwise)
    int cache_off;                 // KEEP instantiation: byte offset behind the rows of 32 bytes per thread of item data
    long long *esum;               // [2 parities][R][2] per-replica exact sums of an exchange, indexed by GLOBAL replica;
                                   // the local slice is zeroed by the host before the launch (blocks add deltas)
    // replicas spread over several GPUs (pymc_rx_connect): the all-gather of the energies is done by this kernel
    int nranks, rank;              // 1, 0 on a single GPU
    PBPacket *xpk;                        // [2 parities][R][2] mine: the sums of the other GPUs' replicas arrive here as packets
    PBPacket *peer_xpk[PB_MAXRANKS];      // the other GPUs' xpk arrays (CUDA IPC mappings); [rank] unused
    unsigned long long *out;       // [nrep][2] zeroed by the host: the exact sums at the end of the launch
    const double *betas;
    double c_bond, c_second;
    int32_t *rep_to_T[2], *T_to_rep[2];  // maps double-buffered by exchange parity; [0] holds the maps at entry
    unsigned long long *tried, *accepted;
    double *energies;              // [R] energies as of the last exchange (what pymc_rx_get_energies returns)
    int ntemps;
};

__device__ __forceinline__ void pb_grid_barrier(unsigned int *counter, unsigned int target, int *err) {
    __syncthreads();  // this block's stores before thread 0's fence
    if (threadIdx.x == 0) {
        pb_fence();
        atomicAdd(counter, 1u);
        if (pb_ld_acquire(counter) < target) {
            const long long t0 = clock64();
            unsigned int spins = 0;
            while (pb_ld_acquire(counter) < target) {
                __nanosleep(32);
                if ((++spins & 1023u) == 0) {
                    if (*(volatile unsigned int *)(counter + 1)) break;  // somebody has given up: so do I
                    if (clock64() - t0 > 4000000000LL) {
                        *(volatile unsigned int *)(counter + 1) = 1u;
                        *err = 1;
                        break;
                    }
                }
            }
        }
    }
    __syncthreads();
}

// grid = B blocks (<= number of SMs, cooperative launch: all resident), block = multiple of 32 threads with
// 4 * nw <= threads, dynamic shared memory = 2 * (largest nrows) * nw * 8 bytes.
// SLAB: one lattice in row slabs over several GPUs -- the owned rows of the local array are the "replica", row 0 / the last row
// are exported into the exchange buffers of the GPUs above / below, the random stream is indexed by whole-lattice rows.
template <int MODEL, int LAYOUT, bool KEEP, bool SLAB, class LatT>
__global__ void __launch_bounds__(PB_MAX_THREADS, 1) pbatch_kernel(const __grid_constant__ LatT L, const __grid_constant__ PBatchArgs a) {
    extern __shared__ __align__(16) uint64_t s_rows[];
    __shared__ uint32_t s_plan[PB_MAXSEG * PLAN_WORDS];
    __shared__ int s_segoff[PB_MAXSEG], s_segstart[PB_MAXSEG], s_T[PB_MAXSEG];
    __shared__ long long s_acc[PB_MAXSEG * 2], s_prev[2][PB_MAXSEG * 2];
    __shared__ PBGeom s_geom;
    const int b = blockIdx.x, B = gridDim.x, t = threadIdx.x, T = blockDim.x;
    // rows of a replica held here: all of them, or -- one lattice in row slabs over several GPUs -- the owned rows of the
    // local array (which carries halo rows for the per-sweep kernel: storage row = own_lo + row, lenS rows per colour)
    const int lenX = L.own_hi - L.own_lo, lenS = L.lenX, row_lo = L.own_lo, nw = L.nw, nk = L.nk, nrep = L.nrep;

    if (t == 0) pb_geometry(b, B, nrep, lenX, s_geom);
    __syncthreads();
    const PBGeom &g = s_geom;  // stays in shared memory: block-uniform, read where needed
    if (t < g.nseg) {
        s_segstart[t] = t == 0 ? 0 : (g.rep_first + t) * lenX - g.R0;
        s_segoff[t] = t == 0 ? g.i0 : -((g.rep_first + t) * lenX - g.R0);
        s_T[t] = L.single_plan ? 0 : a.rep_to_T[0][L.rep_offset + g.rep_first + t];
    }
    if (t < 2 * PB_MAXSEG) s_prev[0][t] = s_prev[1][t] = 0;
    PBCtx c;
    c.s = s_rows;
    c.segoff = s_segoff;
    c.segstart = s_segstart;
    c.splan = s_plan;
    c.xbuf = a.xbuf;
    c.xbuf_up = a.xbuf_up;
    c.xbuf_dn = a.xbuf_dn;
    c.epoch = a.epoch;
    c.row0 = a.row0;
    c.wait_cycles = a.xbuf_up != a.xbuf ? 20000000000LL : 4000000000LL;  // neighbours on other GPUs may launch later
    c.nblocks = B;
    c.abort_word = a.bar + 1;
    c.err = a.err;
    c.lenX = lenX;
    c.nw = nw;
    c.nw_shift = (nw & (nw - 1)) == 0 ? 31 - __clz(nw) : -1;
    c.nslots = pb_nslots(B, nrep);
    c.b = b;
    c.rep_offset = (uint32_t)L.rep_offset;
    c.key = &L.key;

    // ---- load my rows (both colours) ---------------------------------------------------------------------
    for (int colour = 0; colour < 2; ++colour) {
        if (LAYOUT == PYMC_LAYOUT_BITS) {
            for (int idx = t; idx < g.nrows * nw; idx += T) {
                const int lr = idx / nw, w = idx - lr * nw;
                const int gr = g.R0 + lr, rep = gr / lenX, i = gr - rep * lenX;
                s_rows[((size_t)colour * g.nrows + lr) * nw + w] = L.bits[((size_t)(rep * 2 + colour) * lenS + row_lo + i) * nw + w];
            }
        } else {
            uint16_t *s16 = reinterpret_cast<uint16_t *>(s_rows + (size_t)colour * g.nrows * nw);
            for (int lr = t >> 5; lr < g.nrows; lr += T >> 5) {  // a warp takes one row at a time: 16-byte chunks, 4 per word
                const int gr = g.R0 + lr, rep = gr / lenX, i = gr - rep * lenX;
                const uint4 *src = reinterpret_cast<const uint4 *>(L.i8 + ((size_t)(rep * 2 + colour) * lenS + row_lo + i) * nk);
                for (int cc = t & 31; cc < nw * 4; cc += 32) s16[lr * nw * 4 + cc] = (uint16_t)pack16<MODEL>(__ldg(src + cc));
            }
        }
    }
    __syncthreads();
    auto load_plans = [&]() {  // plan rows of my replicas at their current temperatures
        for (int k = t; k < g.nseg * PLAN_WORDS; k += T) {
            const int seg = k / PLAN_WORDS, j = k - seg * PLAN_WORDS;
            s_plan[k] = L.single_plan ? L.plan0[j] : __ldg(L.plans + (size_t)s_T[seg] * PLAN_WORDS + j);
        }
    };
    load_plans();
    pb_publish_loaded(g, c, t, T);
    __syncthreads();

    // pass-invariant per-thread item data in shared memory behind the rows, if the host found room for it
    PBI4 *const ca = KEEP ? reinterpret_cast<PBI4 *>(reinterpret_cast<unsigned char *>(s_rows) + a.cache_off) : nullptr;
    PBI4 *const cb = KEEP ? ca + T : nullptr;
    if (KEEP) {
        PBItems f;
        pb_items(g, c, t, T, f);
        pb_items_store(f, ca, cb, t);  // read back by this thread only
    }
    Plan<MODEL> plan;
    int plan_seg = -1;

    unsigned int events = 1, arrivals = 0;
    int scheme = a.scheme0, cur = 0;
    unsigned long long round = a.round0;
    for (long long sw = 0; sw < a.nsweeps; ++sw) {
#pragma unroll 1
        for (int colour = 0; colour < 2; ++colour) {
            pb_pass<MODEL, KEEP, SLAB>(g, c, colour, 2 * (a.sweep0 + sw) + colour, events, t, T, plan, plan_seg, ca, cb);
            ++events;
            __syncthreads();
        }
        const bool last = sw == a.nsweeps - 1;
        const bool xch = a.xchg_every > 0 && (a.xchg_phase + sw + 1) % a.xchg_every == 0;
        if (!last && !xch) continue;

        // ---- exact sums of my rows per replica ------------------------------------------------------------
        if (t < 2 * g.nseg) s_acc[t] = 0;
        __syncthreads();
        pb_reduce<MODEL>(g, c, events, s_acc, t, T);
        __syncthreads();
        if (xch) {
            const int par = (int)(round & 1ull), R = a.ntemps;
            long long *es = a.esum + (size_t)par * R * 2;
            if (t < 2 * g.nseg) {  // what my rows add to the replica's sums since I last wrote to this parity
                const long long d = s_acc[t] - s_prev[par][t];
                s_prev[par][t] = s_acc[t];
                if (d) atomicAdd((unsigned long long *)(es + 2 * (size_t)(L.rep_offset + g.rep_first + t / 2) + (t & 1)), (unsigned long long)d);
            }
            arrivals += (unsigned int)B;
            pb_grid_barrier(a.bar, arrivals, a.err);  // the sums of all replicas of this GPU are complete
            PBRemote rm;
            rm.xpk = nullptr;
            if (a.nranks > 1) {
                // All-gather over NVLink, written by hand: block 0 stores the sums of my replicas straight into every other
                // GPU's packet array (peer memory) as {32 data bits, exchange round + 1} packets -- the round rides in the
                // same atomic 8-byte store as the data, so there is no fence, no separate flag and no second trip over
                // NVLink; whoever needs the sums of a replica of another GPU (pb_energy) repeats the load of its packets
                // from local memory until they carry this round.  (First form: plain stores, __threadfence_system, a flag
                // per GPU: two dependent NVLink trips and two system-scope fences per exchange.)
                const unsigned int xr = (unsigned int)round + 1u;
                if (b == 0) {
                    for (int k = t; k < 2 * nrep * a.nranks; k += T) {
                        const int peer = k / (2 * nrep), j = k - peer * 2 * nrep;
                        if (peer != a.rank) {
                            const size_t at = (size_t)par * R * 2 + 2 * (size_t)L.rep_offset + j;
                            pb_put(a.peer_xpk[peer] + at, (uint64_t)__ldcg(a.esum + at), xr);
                        }
                    }
                }
                rm.xpk = a.xpk + (size_t)par * R * 2;
                rm.need = xr;
                rm.own_lo = (int)L.rep_offset;
                rm.own_hi = (int)L.rep_offset + nrep;
                rm.err = a.err;
                rm.abort_word = a.bar + 1;
            }
            // ---- the pair decisions of Xtrial (py_mc/mc_mpi.py:102-140) ---------------------------------------------
            // A decision is a pure function of (pair, round, the two energies): every block works out the pairs its own
            // replicas are in (it needs their new temperatures), block 0 works out all of them and keeps the books.
            const int nxt = cur ^ 1;
            const int32_t *t2r = a.T_to_rep[cur], *r2t = a.rep_to_T[cur];
            auto decide = [&](int tl, int &r_lo, int &r_hi) -> bool {  // pair (tl, tl + 1) of temperature indices
                r_lo = t2r[tl];
                r_hi = t2r[tl + 1];
                const double e_lo = pb_energy(es, rm, r_lo, a.c_bond, a.c_second);
                const double e_hi = pb_energy(es, rm, r_hi, a.c_bond, a.c_second);
                const double delta = __dmul_rn(__dsub_rn(a.betas[tl + 1], a.betas[tl]), __dsub_rn(e_lo, e_hi));
                if (delta < 0.0) return true;
                const uint4 o = philox4x32_10((uint32_t)tl, 0u, (uint32_t)round, ctr3(DOMAIN_RX, 0, round), L.key);
                return (double)o.x * (1.0 / 4294967296.0) <= exp(-delta);
            };
            if (b != 0) {
                if (t < g.nseg) {
                    const int r = (int)L.rep_offset + g.rep_first + t, Tc = r2t[r];
                    const int tl = Tc - (Tc >= scheme ? ((Tc - scheme) & 1) : 0);  // lower index of my pair
                    int Tn = Tc;
                    if (Tc >= scheme && tl + 1 < R) {
                        int r_lo, r_hi;
                        if (decide(tl, r_lo, r_hi)) Tn = r == r_lo ? tl + 1 : tl;
                    }
                    s_T[t] = Tn;
                }
            } else {
                for (int tt = t; tt < R; tt += T) {  // temperatures without a partner this round keep their replica
                    const bool paired = tt >= scheme && (((tt - scheme) & 1) ? true : tt + 1 < R);
                    if (!paired) {
                        const int r = t2r[tt];
                        a.T_to_rep[nxt][tt] = r;
                        a.rep_to_T[nxt][r] = tt;
                    }
                }
                for (int r = t; r < R; r += T) a.energies[r] = pb_energy(es, rm, r, a.c_bond, a.c_second);
                const int g_first = (int)L.rep_offset + g.rep_first;
                for (int p = t; 2 * p + scheme + 1 < R; p += T) {
                    const int tl = 2 * p + scheme;
                    int r_lo, r_hi;
                    const bool accept = decide(tl, r_lo, r_hi);
                    const int T_lo = accept ? tl + 1 : tl, T_hi = accept ? tl : tl + 1;  // new temperatures of r_lo, r_hi
                    if (r_lo >= g_first && r_lo < g_first + g.nseg) s_T[r_lo - g_first] = T_lo;
                    if (r_hi >= g_first && r_hi < g_first + g.nseg) s_T[r_hi - g_first] = T_hi;
                    a.rep_to_T[nxt][r_lo] = T_lo;
                    a.rep_to_T[nxt][r_hi] = T_hi;
                    a.T_to_rep[nxt][T_lo] = r_lo;
                    a.T_to_rep[nxt][T_hi] = r_hi;
                    a.tried[tl] += 1;
                    if (accept) a.accepted[tl] += 1;
                }
            }
            __syncthreads();
            load_plans();
            plan_seg = -1;  // the plan rows have changed
            __syncthreads();
            cur = nxt;
            scheme ^= 1;
            ++round;
        }
        if (last && t < 2 * g.nseg && s_acc[t])
            atomicAdd(a.out + 2 * (size_t)(g.rep_first + t / 2) + (t & 1), (unsigned long long)s_acc[t]);
    }
    __syncthreads();

    // ---- store my rows ---------------------------------------------------------------------------------------
    for (int colour = 0; colour < 2; ++colour) {
        if (LAYOUT == PYMC_LAYOUT_BITS) {
            for (int idx = t; idx < g.nrows * nw; idx += T) {
                const int lr = idx / nw, w = idx - lr * nw;
                const int gr = g.R0 + lr, rep = gr / lenX, i = gr - rep * lenX;
                L.bits[((size_t)(rep * 2 + colour) * lenS + row_lo + i) * nw + w] = s_rows[((size_t)colour * g.nrows + lr) * nw + w];
            }
        } else {
            const uint16_t *s16 = reinterpret_cast<const uint16_t *>(s_rows + (size_t)colour * g.nrows * nw);
            for (int lr = t >> 5; lr < g.nrows; lr += T >> 5) {
                const int gr = g.R0 + lr, rep = gr / lenX, i = gr - rep * lenX;
                uint4 *dst = reinterpret_cast<uint4 *>(L.i8 + ((size_t)(rep * 2 + colour) * lenS + row_lo + i) * nk);
                for (int cc = t & 31; cc < nw * 4; cc += 32) dst[cc] = unpack16<MODEL>(s16[lr * nw * 4 + cc]);
            }
        }
    }
}

#endif  // __CUDACC__

}  // namespace pmc
