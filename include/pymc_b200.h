/* pymc_b200.h -- C ABI of the B200-native Metropolis-sweep / replica-exchange engine.
 *
 * This is the drop-in boundary for py_mc's hot path.  The
 * reference is pure Python and has no FFI of its own; each entry point below names
 * the reference code whose work it takes over (paths relative to the reference tree).
 * INTEGRATION.md shows the ctypes stub a py_mc maintainer would add.
 *
 * Conventions
 *   - every function returns 0 on success and a negative pymc_status on failure;
 *     pymc_last_error() returns a thread-local human-readable message.
 *   - no exceptions, no C++/torch types: plain pointers and sizes only.
 *   - host pointers are caller-owned; all device memory belongs to the handle.
 *   - a handle is bound to one CUDA device and one stream; calls on one handle must be
 *     serialised by the caller.  pymc_sweep / pymc_rx_* only enqueue work; functions that
 *     return host data synchronise the handle's stream.
 *   - there is no CPU fallback: unsupported shapes return PYMC_ERR_UNSUPPORTED.
 */
#ifndef PYMC_B200_H
#define PYMC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PYMC_ABI_VERSION 1

typedef enum {
    PYMC_OK = 0,
    PYMC_ERR_BADARG = -1,      /* null pointer, out-of-range index, wrong size */
    PYMC_ERR_CUDA = -2,        /* a CUDA runtime call failed (message has the CUDA error) */
    PYMC_ERR_UNSUPPORTED = -3, /* odd lattice edge, width not packable for the layout, ... */
    PYMC_ERR_STATE = -4        /* call order problem, e.g. sweep before tables were set */
} pymc_status;

typedef enum {
    PYMC_MODEL_ISING = 0,  /* spins +1/-1,   examples/ising2D.py:7-52   */
    PYMC_MODEL_LATGAS = 1  /* occupancy 0/1, examples/latgas.py:7-57    */
} pymc_model;

typedef enum {
    PYMC_LAYOUT_INT8 = 0, /* one byte per site, colour-split [replica][colour][lenX][lenY/2]        */
    PYMC_LAYOUT_BITS = 1  /* one bit per site,  colour-split [replica][colour][lenX][lenY/128] u64; */
                          /* needs lenY % 128 == 0                                                  */
} pymc_layout;

typedef enum {
    PYMC_KERNEL_AUTO = 0,    /* fastest kernel that supports the shape                     */
    PYMC_KERNEL_GENERIC = 1, /* the simple one-thread-per-word kernel (any even lenX, lenY) */
    PYMC_KERNEL_TWO_PASS = 2, /* one launch per colour even where the fused int8 sweep kernel applies */
    PYMC_KERNEL_RESIDENT = 3, /* one block keeps one replica in shared memory for a whole call; needs    */
                              /* lenY % 128 == 0 and 2*lenX*lenY/16 bytes of shared memory.  AUTO takes */
                              /* it for batches of small lattices (>= one replica per SM, <= 32 KiB)    */
    PYMC_KERNEL_PERSISTENT = 4, /* one lattice spread over the shared memory of all SMs (one row slab per  */
                                /* block of a cooperative launch, grid barrier per colour pass) for all   */
                                /* sweeps of a call; needs lenY % 128 == 0 and lenX*lenY/8 bytes within    */
                                /* the GPU's shared memory                                                */
    PYMC_KERNEL_PBATCH = 5,     /* the rows of ALL replicas of the handle cut into one piece per SM and    */
                                /* kept in shared memory for all sweeps -- and, on one GPU, all replica   */
                                /* exchanges -- of a call; neighbour flags instead of grid barriers.      */
                                /* Needs lenY % 128 == 0 and nreplicas*lenX*lenY/8 bytes within the GPU's  */
                                /* shared memory.  AUTO takes it from 128 x 128 sites on (calls of >= 2 sweeps) */
    PYMC_KERNEL_ROWSTREAM = 6   /* int8 layout, one sweep per launch: a block walks down a piece of whole rows   */
                                /* that stream through shared memory (bulk asynchronous copies), both colours in   */
                                /* one pass.  Needs lenY = 128 * 2^k (1024 .. 65536).  AUTO uses it for the sweeps  */
                                /* the resident kernels do not take (single-sweep calls, lattices beyond the chip) */
} pymc_kernel;

/* Number of acceptance-table entries per temperature: Ising 5 (index = number of aligned
 * neighbours 0..4), lattice gas 10 (index = 5*c + number of occupied neighbours). */
#define PYMC_NCLASS_ISING 5
#define PYMC_NCLASS_LATGAS 10

typedef struct {
    int32_t abi_version;       /* PYMC_ABI_VERSION */
    int32_t device;            /* CUDA device ordinal */
    int32_t model;             /* pymc_model */
    int32_t layout;            /* pymc_layout */
    int32_t lenX, lenY;        /* lattice edges (config[x][y], y contiguous); both even; at most 2^34 sites per replica (a slab handle: per slab) */
    int32_t nreplicas;         /* replicas held by this handle (this GPU) */
    int32_t replica_offset;    /* global index of local replica 0 (multi-GPU partition) */
    int32_t nreplicas_global;  /* replicas over all GPUs (== nreplicas on one GPU) */
    int32_t ntemps;            /* rows of the acceptance table (temperatures / parameter sets) */
    uint64_t seed;             /* Philox key; the same on every GPU of one run */
    int32_t kernel;            /* pymc_kernel */
    int32_t reserved;
} pymc_params;

typedef struct pymc_handle pymc_handle;

/* ---- library ------------------------------------------------------------------- */
int pymc_abi_version(void);
const char *pymc_last_error(void);
int pymc_device_count(int *count);

/* ---- lifetime -------------------------------------------------------------------
 * Replaces CanonicalMonteCarlo.__init__ (py_mc/mc.py:163-168) for a whole batch of
 * replicas and ParallelMC.__init__'s one-engine-per-rank set-up (py_mc/mc_mpi.py:38-56). */
int pymc_create(const pymc_params *params, pymc_handle **out);
int pymc_destroy(pymc_handle *h);
/* Run on a caller-provided CUDA stream (cudaStream_t cast to void*; NULL = CUDA's legacy default
 * stream) so that the caller's own work on that stream -- e.g. an NCCL all-gather -- is ordered
 * with the kernels.  pymc_reset_stream goes back to the handle's private non-blocking stream. */
int pymc_set_stream(pymc_handle *h, void *cuda_stream);
int pymc_reset_stream(pymc_handle *h);
int pymc_synchronize(pymc_handle *h);
/* Barriers inside kernels (persistent sweep kernel, slab neighbour barrier) wait a bounded time.  If one gives up,
 * the lattices of the handle are no longer valid: the next call that synchronises and returns data
 * (pymc_synchronize, pymc_download_config, pymc_get_tracked, pymc_reduce_observables[_static], pymc_get_replica_to_T,
 * pymc_rx_get_*, pymc_stats_get, pymc_timer_stop) returns PYMC_ERR_CUDA with a message instead of data, once.
 * No counterpart in the reference.  pymc_debug_raise_async_error raises that condition artificially
 * (which = 0 persistent kernel, 1 slab barrier) so that the error path can be tested. */
int pymc_debug_raise_async_error(pymc_handle *h, int which);

/* ---- configurations ---------------------------------------------------------------
 * Host format: int8 row-major [lenX][lenY], +1/-1 (Ising) or 0/1 (lattice gas): the
 * integer image of the reference's float64 config array (examples/ising2D.py:56-70). */
int pymc_upload_config(pymc_handle *h, int replica, const int8_t *host, size_t nsites);
int pymc_download_config(pymc_handle *h, int replica, int8_t *host, size_t nsites);
/* The same two calls without the wait at the end: they only enqueue the copy and the layout kernel on the handle's
 * stream and return.  `host` must be page-locked (pymc_host_alloc) and must stay untouched (upload) / is not valid
 * (download) until pymc_synchronize on this handle returns; a failed in-kernel barrier is reported there.  With one
 * handle per lattice in flight -- each handle has its own stream -- the copies of one lattice overlap the sweeps of
 * another (py_mc_b200/pipeline.py).  No counterpart in the reference (its arrays never leave host memory). */
int pymc_upload_config_async(pymc_handle *h, int replica, const int8_t *host, size_t nsites);
int pymc_download_config_async(pymc_handle *h, int replica, int8_t *host, size_t nsites);
/* Hot start on the device, rule of ising2D_config.prepare_random (examples/ising2D.py:64-70)
 * applied to stream domain 1 (see oracle/philox.py); all local replicas. */
int pymc_init_random(pymc_handle *h);

/* ---- acceptance tables --------------------------------------------------------------
 * thr[t][c] = floor(exp(-dE_c/kT_t) * 2^32) clamped to 2^32-1, or 2^32-1 when dE_c < 0:
 * the integer form of the Metropolis test of MCstep (py_mc/mc.py:175-184).  The caller
 * computes it with the host libm so acceptance is bit-exact with the reference. */
int pymc_set_accept_tables(pymc_handle *h, const uint32_t *thr, int ntemps, int nclass);
/* replica -> table row.  Mirrors rank_to_T (py_mc/mc_mpi.py:79); n == nreplicas_global. */
int pymc_set_replica_to_T(pymc_handle *h, const int32_t *map, int n);
int pymc_get_replica_to_T(pymc_handle *h, int32_t *map, int n);

/* ---- the hot path -----------------------------------------------------------------
 * nsweeps full checkerboard sweeps of every local replica; one sweep = lenX*lenY
 * Metropolis trials (= that many MCstep() calls of py_mc/mc.py:170-185 driven in
 * checkerboard order).  Asynchronous. */
int pymc_sweep(pymc_handle *h, int64_t nsweeps);
/* Which kernel pymc_sweep uses for this handle and how many launches one sweep takes (1 for the
 * fused two-colour int8 kernel, otherwise 2); for benchmarks and logs. */
int pymc_sweep_kernel_info(pymc_handle *h, int *launches_per_sweep, const char **kernel_name);
/* The same for a call of exactly nsweeps sweeps: calls of a single sweep take a per-sweep kernel (the row-streaming or
 * the tiled int8 kernel, the bit-packed row kernel) even on handles whose longer calls stay on chip. */
int pymc_sweep_kernel_for(pymc_handle *h, int64_t nsweeps, int *launches_per_sweep, const char **kernel_name);
int pymc_get_sweep_counter(pymc_handle *h, uint64_t *counter);
int pymc_set_sweep_counter(pymc_handle *h, uint64_t counter);

/* ---- observables --------------------------------------------------------------------
 * Exact integers per local replica.  Ising: bond = sum s(i,j)(s(i-1,j)+s(i,j-1)) so that
 * E = J*bond (examples/ising2D.py:15-22), second = sum s (ising2D.py:24-26).
 * Lattice gas: bond = sum c(i,j)(c(i-1,j)+c(i,j-1)), second = sum c (latgas.py:17-29).
 * pymc_reduce_observables recomputes them from the lattice (what model.energy(config)
 * does at the top of run(), py_mc/mc.py:191) and re-bases the tracked values;
 * pymc_get_tracked returns the incrementally tracked values (energy += dE, mc.py:177). */
int pymc_reduce_observables(pymc_handle *h, int64_t *bond, int64_t *second);
int pymc_get_tracked(pymc_handle *h, int64_t *bond, int64_t *second);
/* pymc_get_tracked without the wait: pairs[2*r] = bond, pairs[2*r+1] = second of local replica r, in page-locked
 * host memory, valid once pymc_synchronize on this handle has returned. */
int pymc_get_tracked_async(pymc_handle *h, int64_t *pairs);

/* No counterpart in the reference (it has no device memory to manage).  Same results and same
 * side effect as pymc_reduce_observables, for callers that must not allocate while running: a
 * process that has given all device memory to another allocator or works under a fixed memory
 * budget (a late cudaMalloc can fail there), or a loop that cannot take the device-wide
 * synchronisation cudaMalloc/cudaFree imply.  It calls no cudaMalloc, cudaFree or other
 * allocating routine, host or device: the kernel's accumulators sit in a static 64 KiB device
 * array that each GPU holds from library load to process exit.
 * Limits and rules:
 *   - at most 4096 replicas per handle; more returns PYMC_ERR_UNSUPPORTED with a message.
 *   - the first call on each device looks up the array's device address (one
 *     cudaGetSymbolAddress); make that first call before any stream capture begins.
 *   - all handles on a GPU share the array.  The call clears the part it uses, so earlier contents
 *     never leak into a result, but two such calls whose GPU work overlaps (different streams or
 *     threads) race on it: run them one after the other. */
int pymc_reduce_observables_static(pymc_handle *h, int64_t *bond, int64_t *second);

/* ---- parity helper -------------------------------------------------------------------
 * For n sites (x[i], y[i]) of one replica return the site value and the sum of its four
 * neighbours: the integers from which trialstep's dE is formed (examples/ising2D.py:39-45,
 * examples/latgas.py:42-50).  The configuration is left unchanged. */
int pymc_propose(pymc_handle *h, int replica, const int32_t *x, const int32_t *y, int n,
                 int32_t *site_out, int32_t *nnsum_out);

/* ---- one lattice over several GPUs (row slabs) ----------------------------------------
 * No counterpart in the reference (it has no domain decomposition).
 * pymc_create_slab: params->lenX/lenY are the WHOLE lattice; this handle owns rows
 *   [slab_index*lenX/nslabs, +lenX/nslabs) plus two halo rows on either side.  int8 layout, fused
 *   sweep kernel; needs (lenX/nslabs) % 32 == 0 and lenY % 1024 == 0.  The random stream is indexed by
 *   whole-lattice rows, so the union of the slabs follows exactly the single-GPU trajectory.
 * pymc_slab_ipc_export / pymc_slab_connect: each process exports four CUDA IPC handles (its two lattice buffers, its
 *   flag words and the exchange buffer of the batch-resident kernel: 4 x 64 bytes) and is given those of the slab above
 *   and below (periodic).  After that a pymc_sweep call of ONE sweep stores the first/last two rows it computes straight
 *   into the neighbours' halo rows over NVLink from inside the tiled sweep kernel, with the neighbour barrier in the same
 *   kernel (bounded wait: pymc_slab_status reports 1 if a neighbour never arrived); a call of TWO OR MORE sweeps keeps the
 *   slab in the shared memory of the GPU for the whole call (pbatch_kernel) and hands the first / last row of every
 *   colour pass to the neighbouring GPUs as {data, pass number} packets stored into their exchange buffers (peer
 *   memory): one launch per GPU and call, no host and no NCCL in between.  All slabs must issue the same pymc_sweep calls.
 * pymc_upload_config / pymc_download_config move the owned rows only ((lenX/nslabs)*lenY sites);
 *   after uploads, every slab calls pymc_slab_push_halos once.  pymc_init_random needs no exchange.
 * pymc_reduce_observables / pymc_get_tracked return this slab's share: add them up over the slabs. */
int pymc_create_slab(const pymc_params *params, int slab_index, int nslabs, pymc_handle **out);
int pymc_slab_ipc_export(pymc_handle *h, void *handles, size_t nbytes);
int pymc_slab_connect(pymc_handle *h, const void *up_handles, const void *down_handles, size_t nbytes);
/* Neighbours inside one process (tests, one host thread driving several slabs): plain handles instead
 * of IPC, no flag barrier; the caller puts all slabs on one stream and sweeps them one sweep at a time in turn. */
int pymc_slab_connect_local(pymc_handle *h, pymc_handle *up, pymc_handle *down);
int pymc_slab_push_halos(pymc_handle *h);
int pymc_slab_status(pymc_handle *h, int *status);

/* ---- particle moves of the lattice gas ------------------------------------------------
 * Takes over the move branch of latticegas.trialstep (examples/latgas_move.py:56-83) with
 * nearest-neighbour particle-hole exchanges on 16 conflict-free bond classes and the physically
 * correct dE = J*d, d = (occupied neighbours gained) - (lost) in -3..3 (the reference's dE for moves
 * is defective).  thr[t][d+3] as in pymc_set_accept_tables.  One pair sweep tries every
 * nearest-neighbour bond once (2*lenX*lenY bonds); class_mask selects bond classes (bit c = class c;
 * 0xFFFF all, 0x00FF horizontal, 0xFF00 vertical).  Needs lenX % 4 == 0 and lenY % 4 == 0. */
int pymc_set_pair_tables(pymc_handle *h, const uint32_t *thr, int ntemps, int nclass);
int pymc_pair_sweep(pymc_handle *h, int64_t nsweeps, uint32_t class_mask);
/* nsteps times: one flip sweep (pymc_sweep), then one pair pass over the classes of mask_even or mask_odd,
 * whichever the parity of the pair counter selects (0 = no pair pass in that step).  The mixed flip / move
 * trial of latticegas.trialstep (examples/latgas_move.py:31-60) batched: (0x00FF, 0xFF00) alternates
 * horizontal and vertical bonds.  Same trajectory as the separate calls; one launch on resident handles. */
int pymc_mixed_sweep(pymc_handle *h, int64_t nsteps, uint32_t mask_even, uint32_t mask_odd);
int pymc_get_pair_counter(pymc_handle *h, uint64_t *counter);
int pymc_set_pair_counter(pymc_handle *h, uint64_t counter);

/* ---- hard discs -----------------------------------------------------------------------
 * Batched replicas of the continuum hard-disc model of examples/2D_hardcore/model_setup.py.
 * pymc_discs_run      ntrials single-disc trials per replica = that many MCstep() calls
 *                     (py_mc/mc.py:170-185) on HC_2D.trialstep / newconfig (model_setup.py:40-67);
 *                     isotropic = 0 reproduces the reference's diagonal displacement (:48).
 * pymc_discs_propose  the (id, new coordinates, overlap) the next trial of one replica would
 *                     propose, without applying it: the content of trialstep's (dconfig, dE).
 * pymc_discs_observe  overlap[r] = 1 iff HC_2D.energy would return inf (model_setup.py:24-37);
 *                     counts[r][b] = 2 * number of pairs with around(dist/dr) == b+1, the integer part
 *                     of g_r (model_setup.py:116-122); the prefactor is applied by the caller.
 * coors are float64 [ndisc][2]; at most 32 discs per replica (the shipped script uses 20). */
typedef struct pymc_discs pymc_discs;
int pymc_discs_create(int device, int nreplicas, int replica_offset, int ndisc, double Lcell, double rdisc,
                      double maxstep, uint64_t seed, int isotropic, pymc_discs **out);
int pymc_discs_destroy(pymc_discs *d);
int pymc_discs_upload(pymc_discs *d, int replica, const double *coors);
int pymc_discs_download(pymc_discs *d, int replica, double *coors);
int pymc_discs_run(pymc_discs *d, int64_t ntrials);
int pymc_discs_propose(pymc_discs *d, int replica, int *id, double *newcoor, int *overlap);
int pymc_discs_observe(pymc_discs *d, double dr, int nbins, int32_t *overlap, int64_t *counts);
int pymc_discs_counters(pymc_discs *d, uint64_t *trial_counter, uint64_t *accepted);
int pymc_discs_set_trial_counter(pymc_discs *d, uint64_t counter);
/* runs ntrials like pymc_discs_run and returns the device time of that launch (CUDA events) */
int pymc_discs_timer(pymc_discs *d, int64_t ntrials, float *elapsed_ms);

/* ---- replica exchange -----------------------------------------------------------------
 * Replaces TemperatureRX_MPI (py_mc/mc_mpi.py:75-209).  Needs ntemps == nreplicas_global
 * (the reference's own requirement, mc_mpi.py:47-52).
 * pymc_rx_setup      betas = 1/kTs (mc_mpi.py:78); energy = c_bond*bond + c_second*second
 *                    (J*B for Ising; J*B + (Eads-mu)*N for the lattice gas).
 * pymc_rx_energies   writes the energies of the local replicas into the shared energies[R] array
 *                    at [replica_offset, +nreplicas)  (mycalc.energy, mc_mpi.py:117,135).
 * pymc_rx_buffers    device pointers of energies[R] (float64) and replica_to_T[R] (int32): with
 *                    several GPUs the caller all-gathers energies in place (NCCL) between
 *                    pymc_rx_energies and pymc_rx_exchange; this replaces the Send/Recv of one
 *                    double per pair and the Allgather of the rank->T map (mc_mpi.py:106-138).
 * pymc_rx_exchange   Xtrial(XCscheme) (mc_mpi.py:102-140) for every pair at once, on the device;
 *                    identical on every GPU because it reads only all-gathered data.
 * pymc_rx_cycle      [sweeps, energies, exchange] x ncycles with alternating scheme (mc_mpi.py:165-169), no host
 *                    synchronisation inside; one kernel launch where the batch-resident kernel applies.  On a
 *                    handle that holds only part of the replicas it needs pymc_rx_connect (below). */
int pymc_rx_setup(pymc_handle *h, const double *betas, int ntemps, double c_bond, double c_second);
int pymc_rx_energies(pymc_handle *h);
int pymc_rx_buffers(pymc_handle *h, void **energies_all, void **replica_to_T);
int pymc_rx_exchange(pymc_handle *h, int xcscheme);
int pymc_rx_cycle(pymc_handle *h, int64_t ncycles, int64_t sweeps_per_exchange, int first_scheme);
/* The ladder spread over several GPUs with the exchange inside the sweep kernel (replaces the Send/Recv of one double
 * per pair and the Allgather of mc_mpi.py:106-138 by stores into peer memory over NVLink; no host, no NCCL in the loop).
 * pymc_rx_ipc_export  two CUDA IPC handles of this GPU (2 x 64 bytes): the array the other GPUs store the sums of their
 *                     replicas into as {data, round} packets, and a spare word array.
 * pymc_rx_connect     all_handles = the exports of all GPUs in rank order (nranks x 128 bytes).  Needs the block
 *                     partition rank g <-> replicas [g*R/G, (g+1)*R/G) and the batch-resident kernel.
 * Afterwards pymc_rx_cycle works on the partitioned handle: ONE launch per GPU for all sweeps and exchanges of the call;
 * every GPU must make the same calls.  A GPU that waits ten seconds for another one raises the handle's error
 * (see pymc_synchronize).  pymc_rx_set_round to an earlier round clears the packet array (packets carry the round): call it on every GPU
 * and synchronise the processes before the next cycle. */
int pymc_rx_ipc_export(pymc_handle *h, void *handles, size_t nbytes);
int pymc_rx_connect(pymc_handle *h, const void *all_handles, size_t nbytes, int nranks, int rank);
int pymc_rx_get_energies(pymc_handle *h, double *energies, int n);
/* attempts / acceptances per temperature pair, indexed by the lower temperature index */
int pymc_rx_get_stats(pymc_handle *h, uint64_t *tried, uint64_t *accepted, int n);
int pymc_rx_get_round(pymc_handle *h, uint64_t *round);
int pymc_rx_set_round(pymc_handle *h, uint64_t round);

/* ---- device-side statistics ----------------------------------------------------------
 * Per-temperature sums of the standard observables without a host round trip per sample: the
 * accumulation obs[rank_to_T[rank]] += observe(...) of TemperatureRX_MPI.run (py_mc/mc_mpi.py:170-177)
 * for E = c_bond*bond + c_second*second and M = second.  sums[t] = {sum E, sum E^2, sum |M|, sum M^2,
 * sum M^4, sum M}, counts[t] = samples taken at table row t.  From these: <E>, <|M|>, the specific
 * heat (<E^2>-<E>^2)/kT^2 and the Binder cumulant 1 - <M^4>/(3<M^2>^2).  With several GPUs each
 * handle holds the contribution of its own replicas; add them up (the reference's closing Allreduce,
 * mc_mpi.py:204). */
#define PYMC_NSTAT 6
int pymc_stats_reset(pymc_handle *h, double c_bond, double c_second);
int pymc_stats_accumulate(pymc_handle *h);
int pymc_stats_get(pymc_handle *h, double *sums, uint64_t *counts, int ntemps, int nstat);

/* Time series by temperature and their blocking analysis on the device: binning (py_mc/mc.py:47-61) and the
 * per-temperature regrouping of examples/analyze_ising2D_mpi.py:28-45 without obs_save.npy / Trank_hist.npy round
 * trips.  pymc_series_record appends one sample: series[sample][T] = (E, |M|) of the replica at temperature index T
 * (entries of replicas on other GPUs stay 0: add the GPUs' series up, pymc_series_get / pymc_series_put).
 * pymc_series_binning: for every (T, observable): mean over samples [discard, count) and the nlevels + 1 standard
 * errors of binning() over all samples (leading count % 2^nlevels dropped); mean[ntemps][2], err[ntemps][2][nlevels+1]. */
int pymc_series_reset(pymc_handle *h, int capacity, double c_bond, double c_second);
int pymc_series_record(pymc_handle *h);
int pymc_series_count(pymc_handle *h, int *count);
int pymc_series_get(pymc_handle *h, double *out, size_t n);
int pymc_series_put(pymc_handle *h, const double *in, int count);
int pymc_series_binning(pymc_handle *h, int discard, int nlevels, double *mean, double *err);

/* ---- pinned host memory ----------------------------------------------------------------
 * Page-locked buffers for upload/download at full PCIe speed (the reference keeps its arrays in
 * ordinary numpy memory, examples/ising2D.py:62; pageable buffers work too, only slower). */
int pymc_host_alloc(size_t bytes, void **out);
int pymc_host_free(void *ptr);

/* ---- timing helper (CUDA events on the handle's stream) ------------------------------ */
int pymc_timer_start(pymc_handle *h);
int pymc_timer_stop(pymc_handle *h, float *elapsed_ms); /* synchronises */

#ifdef __cplusplus
}
#endif
#endif /* PYMC_B200_H */
