/*
 * rhj.h -- C ABI of the B200 radix hash join library (librhj_b200.so).
 *
 * This is the drop-in boundary for ONE path of
 * lushl9301/Distributed-Radxi-Hash-Join-on-GPUs: the pipeline behind
 * hpcjoin::operators::HashJoin::join() (operators/HashJoin.cpp:45-220):
 * HistogramComputation -> Window -> NetworkPartitioning -> LocalPartitioning ->
 * BuildProbe.  Citations are file:line in the reference tree.
 *
 * Conventions
 *  - plain C types only; every pointer named *_dev is a device pointer owned by
 *    the caller unless stated otherwise; `stream` is a cudaStream_t passed as
 *    void* (NULL = the legacy default stream).
 *  - every function returns 0 on success or a negative rhj_status; the text of
 *    the last failure on the calling thread is rhj_last_error().  Nothing in
 *    the library calls exit() (the reference's JOIN_ASSERT does, utils/Debug.h:27-37).
 *  - one rhj_ctx per GPU per rank; a context is not thread-safe.
 *  - no CPU fallback exists: without a CUDA device every compute entry point
 *    fails with RHJ_ERR_CUDA.
 */
#ifndef RHJ_H_
#define RHJ_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rhj_ctx rhj_ctx;

typedef enum rhj_status {
  RHJ_OK = 0,
  RHJ_ERR_INVALID = -1,   /* bad argument / configuration */
  RHJ_ERR_CUDA = -2,      /* CUDA runtime failure (message has the CUDA text) */
  RHJ_ERR_NOMEM = -3,     /* pool exhausted */
  RHJ_ERR_CAPACITY = -4,  /* a receive window / workspace bound was exceeded on the device */
  RHJ_ERR_STATE = -5      /* stage called out of order */
} rhj_status;

/* Replaces the compile-time knobs of core/Configuration.h:19-38. */
typedef struct rhj_config {
  uint32_t network_fanout_bits; /* NETWORK_PARTITIONING_FANOUT (:30), 1..12 */
  uint32_t local_fanout_bits;   /* LOCAL_PARTITIONING_FANOUT (:31), 1..12 */
  uint32_t payload_bits;        /* PAYLOAD_BITS (:37); rid occupies network_fanout_bits+payload_bits bits */
  uint32_t enable_two_level;    /* ENABLE_TWO_LEVEL_PARTITIONING (:28) */
  uint32_t tuple_bytes;         /* 16 = data/Tuple.h:15-22 {u64 key; u64 rid};
                                   8 = packed input, key in bits [32,64), rid in bits [0,32) */
  uint32_t build_capacity;      /* tuples of one build slice held in shared memory (1024, 2048, 4096, 5120, 8192): 0 = default (5120) */
  double allocation_factor;     /* ALLOCATION_FACTOR (:35): receive-window head-room, >= 1.0 */
  uint32_t chunk_tuples;        /* tuples per scheduled chunk of the scatter passes, 0 = default */
  uint32_t flags;               /* RHJ_FLAG_* */
} rhj_config;

#define RHJ_FLAG_HIST_WARP_AGGREGATE 1u /* merge equal digits inside a warp before the shared-memory atomic (one hot key costs
                                           one atomic per warp instead of 32); slower on uniform data */
#define RHJ_FLAG_STAGE_HOST_INPUTS 2u    /* rhj_plan also reserves device staging for rhj_join_host */
/* Size-balanced partition -> rank assignment instead of AssignmentMap.cpp:41-43's p % n (which
 * receives the histograms but ignores them).  Rule, identical on every rank:
 * partitions in order of descending global inner+outer tuple count (ties: ascending id) each go to
 * the rank with the least tuples so far (ties: lowest rank) among the ranks that still have a free
 * slot; rank r has P/n + (r < P%n) slots, exactly as many partitions as p % n gives it.  Offsets
 * keep the reference's definition (OffsetMap.cpp:59-93) under the new owners.  All ranks of a join
 * must set the flag alike. */
#define RHJ_FLAG_BALANCED_ASSIGNMENT 4u
/* Development only: rhj_init also reads the RHJ_* ablation switches from the environment
 * (DESIGN.md section 4).  Without this flag the library never calls getenv. */
#define RHJ_FLAG_DEBUG_ENV 8u
/* Exchange of the network pass (all ranks alike).  Default: two ranks store full 128-byte lines
 * straight into the peer's window from the scatter kernel; three or more gather the remote
 * partitions in a local send buffer and push them to the owners as long sequential runs, group of
 * chunks by group of chunks, beside the scatter.  These flags force one or the other. */
#define RHJ_FLAG_STAGED_EXCHANGE 16u
#define RHJ_FLAG_DIRECT_EXCHANGE 32u

/* Join result: the reference only counts (HashJoin::RESULT_COUNTER, operators/HashJoin.cpp:28,
 * tasks/BuildProbe.cpp:97-115).  The sums are an order-independent checksum over
 * the matching (key, rid_R, rid_S) triples, all modulo 2^64. */
typedef struct rhj_result {
  uint64_t matches;
  uint64_t sum_key;
  uint64_t sum_rid_r;
  uint64_t sum_rid_s;
  uint64_t sum_mix;
} rhj_result;

/* One materialised match: the record ids of the build-side and the probe-side tuple. */
typedef struct rhj_pair {
  uint64_t rid_r;
  uint64_t rid_s;
} rhj_pair;

/* Device timings of the last join, microseconds, under the reference's own keys
 * (performance/Measurements.cpp:136-142, 246-268). */
typedef struct rhj_timings {
  double jtotal_us; /* JTOTAL */
  double jhist_us;  /* JHIST: histograms + offsets */
  double jmpi_us;   /* JMPI : network partitioning */
  double jproc_us;  /* JPROC: local partitioning + build/probe */
  double lp_us;     /* LPTASKTIME */
  double bp_us;     /* BPTASKTIME */
  double h2d_us;    /* host->device staging (rhj_join_host only), else 0 */
  double snetcompl_us; /* SNETCOMPL (performance/Measurements.cpp:157-165, 177; operators/HashJoin.cpp:119-121):
                          waiting for the network pass to complete on all ranks = sum over both relations of the
                          time between rhj_exchange_wait and the start of that relation's local pass (the launcher's
                          barrier lies in between); 0 with one rank.  Included in lp_us. */
  double push_us;      /* staged exchange: from the first push of the inner relation to the end of the last push of
                          the outer one on this rank (the time NVLink carries the join's remote tuples); else 0 */
} rhj_timings;

/* Peer-visible description of a rank's receive windows (replaces MPI_Win_create,
 * data/Window.cpp:35-45). */
typedef struct rhj_window_handle {
  unsigned char ipc[64]; /* cudaIpcMemHandle_t of the pool */
  uint64_t inner_offset; /* byte offset of the inner-relation window inside the pool */
  uint64_t outer_offset;
  uint64_t capacity_inner; /* tuples */
  uint64_t capacity_outer;
  uint64_t flags_offset;   /* byte offset of the rank's arrival flags (u32[2][64]) inside the pool */
} rhj_window_handle;

const char *rhj_last_error(void);
const char *rhj_version(void);
/* kernels of this library launched by the calling process so far */
uint64_t rhj_launch_count(void);

/* ---- context + memory pool ------------------------------------------------
 * rhj_init replaces HashJoin's constructor arguments (numberOfNodes, nodeId;
 * operators/HashJoin.cpp:31-39) plus the Configuration knobs. */
int rhj_init(rhj_ctx **out, int rank, int nranks, int device, const rhj_config *cfg);
int rhj_destroy(rhj_ctx *ctx);

/* memory/Pool.cpp:25-38 Pool::allocate -> one cudaMalloc arena. */
int rhj_pool_allocate(rhj_ctx *ctx, uint64_t bytes);
/* memory/Pool.cpp:40-64 Pool::getMemory: bump allocation, 256-byte aligned (reference: 64). */
int rhj_pool_get_memory(rhj_ctx *ctx, uint64_t bytes, void **out_dev);
/* memory/Pool.cpp:76-79 Pool::reset */
int rhj_pool_reset(rhj_ctx *ctx);
/* memory/Pool.cpp:72-74 Pool::freeAll */
int rhj_pool_free_all(rhj_ctx *ctx);
/* Bytes the pool needs for rhj_plan() with these sizes (inputs not included). */
uint64_t rhj_workspace_bytes(const rhj_config *cfg, int nranks, uint64_t max_local_inner,
                             uint64_t max_local_outer, uint64_t window_inner_cap, uint64_t window_outer_cap);

/* Carve windows and all intermediates out of the pool (allocating a pool of
 * exactly the right size when none exists).  Nothing is allocated afterwards.
 * window_*_cap = tuples this rank can receive per relation; 0 = allocation_factor * max_local_* + 1024.
 * Replaces Window's constructor sizing (data/Window.cpp:16-50).  The reference sizes each window
 * exactly, after the histogram phase (data/Window.cpp:160-178), and therefore never overflows; here
 * the capacity is fixed before the data is seen.  The default only holds for near-uniform partition
 * sizes: with more than one rank and skewed keys (a hot partition, a Zipf probe side) pass the exact
 * size -- rhj_host_offsets' window_sizes from a histogram of the relations, as the HashJoin launchers
 * do -- or a safe bound such as the global relation size.  A window that is too small is detected on
 * the device by rhj_compute_offsets, all later kernels of the join do nothing, and rhj_get_result (the
 * first call that synchronises) returns RHJ_ERR_CAPACITY. */
int rhj_plan(rhj_ctx *ctx, uint64_t max_local_inner, uint64_t max_local_outer, uint64_t window_inner_cap,
             uint64_t window_outer_cap);

/* ---- windows across ranks (MPI_Win_create / MPI_Put target resolution) --- */
int rhj_window_export(rhj_ctx *ctx, rhj_window_handle *out);
/* peer in another process: opens the CUDA IPC handle */
int rhj_window_import(rhj_ctx *ctx, int peer_rank, const rhj_window_handle *handle);
/* peer in this process (same or peer-enabled device): raw device pointers */
int rhj_window_set_peer(rhj_ctx *ctx, int peer_rank, void *inner_window_dev, void *outer_window_dev);
int rhj_window_get(rhj_ctx *ctx, void **inner_window_dev, void **outer_window_dev);
/* 1 when every peer's window was imported through rhj_window_import (one rank per GPU): the network
 * pass then ends by writing this join's number into an arrival flag inside every peer's pool, and
 * rhj_exchange_wait makes `stream` wait (a one-block kernel) until all peers' flags for the relation
 * show it -- the barrier of operators/HashJoin.cpp:119-121 without a collective; the launcher needs
 * no barrier of its own between rhj_exchange_wait and rhj_local_partition_rel.  All ranks must run
 * the same sequence of joins.  0 (ranks in one process via rhj_window_set_peer, or one rank): the
 * launcher provides the barrier. */
int rhj_exchange_has_flags(rhj_ctx *ctx);
/* 1 when this context exchanges through send buffers + pushes (see RHJ_FLAG_STAGED_EXCHANGE). */
int rhj_exchange_is_staged(rhj_ctx *ctx);

/* ---- the four tasks (tasks/Task.h:10-30), individually invocable ---------- */

/* TASK_HISTOGRAM part 1 -- histograms/LocalHistogram.cpp:35-53 for both relations.
 * Leaves u64[2][P] (inner, outer) on the device: rhj_local_histogram_dev(). */
int rhj_histogram(rhj_ctx *ctx, const void *inner_dev, uint64_t inner_n, const void *outer_dev, uint64_t outer_n,
                  void *stream);
uint64_t *rhj_local_histogram_dev(rhj_ctx *ctx);
/* The same phase one relation at a time (which: 0 inner, 1 outer; the inner relation's call starts a new
 * join), so that the stages of the inner relation can run while the outer relation is still on its way:
 * host->device in rhj_join_host, or -- with several ranks -- crossing NVLink.  rhj_compute_offsets_rel
 * takes that relation's gathered histograms, u64[nranks][P] in rank order (NULL with one rank); it is
 * refused under RHJ_FLAG_BALANCED_ASSIGNMENT, whose rule needs both relations' histograms. */
int rhj_histogram_rel(rhj_ctx *ctx, int which, const void *rel_dev, uint64_t n, void *stream);
int rhj_compute_offsets_rel(rhj_ctx *ctx, int which, const uint64_t *gathered_rel_dev, void *stream);

/* TASK_HISTOGRAM part 2 -- GlobalHistogram.cpp:31-48 (sum over ranks), AssignmentMap.cpp:34-49
 * (p % nranks), OffsetMap.cpp:59-93 (base / relative / absolute offsets), all on the device.
 * gathered_dev = u64[nranks][2][P], every rank's rhj_local_histogram_dev() in rank order
 * (an all-gather replaces MPI_Allreduce + MPI_Exscan); NULL is allowed when nranks == 1. */
int rhj_compute_offsets(rhj_ctx *ctx, const uint64_t *gathered_dev, void *stream);

/* tasks/HistogramComputation.cpp:78-130: the nine getters (host copies; synchronises `stream`).
 * which: 0 inner, 1 outer.  Arrays have P = 1 << network_fanout_bits entries. */
int rhj_get_assignment(rhj_ctx *ctx, uint32_t *out_host, void *stream);
int rhj_get_local_histogram(rhj_ctx *ctx, int which, uint64_t *out_host, void *stream);
int rhj_get_global_histogram(rhj_ctx *ctx, int which, uint64_t *out_host, void *stream);
int rhj_get_base_offsets(rhj_ctx *ctx, int which, uint64_t *out_host, void *stream);
int rhj_get_write_offsets(rhj_ctx *ctx, int which, uint64_t *out_host, void *stream);

/* TASK_NET_PARTITION -- tasks/NetworkPartitioning.cpp:74-222 + data/Window.cpp:86-144:
 * radix scatter on the low key bits with 16->8 byte compression through write-combining buffers
 * (cacheline_t, :28-40): whole 128-byte lines go into the owning rank's window -- directly (one or two
 * ranks: peer stores replace MPI_Put) or through a send buffer that is pushed as sequential runs (three and
 * more ranks; RHJ_FLAG_STAGED_EXCHANGE / RHJ_FLAG_DIRECT_EXCHANGE).  Uses the inputs given to rhj_histogram.
 * The local pass of a relation may start once every rank's stores of it have landed: rhj_exchange_wait. */
int rhj_network_partition(rhj_ctx *ctx, void *stream);
/* The same, one relation at a time (which: 0 inner, 1 outer), on a stream of the caller's choice, so that
 * a multi-rank launcher can start the inner relation's barrier + local pass while the outer relation is
 * still crossing NVLink. */
int rhj_network_partition_rel(rhj_ctx *ctx, int which, void *stream);
/* Number of SMs the kernels launched after this call may occupy (0 = all).  For launchers that run
 * two stages side by side on two streams (an NVLink-bound network pass next to an HBM-bound local pass). */
int rhj_set_sm_limit(rhj_ctx *ctx, int sms);

/* TASK_PARTITION -- tasks/LocalPartitioning.cpp:59-250 for every owned partition:
 * histogram, 8-tuple padded prefix, scatter.  No-op unless enable_two_level. */
int rhj_local_partition(rhj_ctx *ctx, void *stream);
/* The same, one relation at a time (which: 0 inner, 1 outer), so that a multi-rank launcher can
 * start the inner relation's local pass while the outer relation is still crossing NVLink:
 *   rhj_network_partition -> rhj_exchange_wait(0) -> [barrier] -> rhj_local_partition_rel(0)
 *                         -> rhj_exchange_wait(1) -> [barrier] -> rhj_local_partition_rel(1) -> rhj_build_probe
 * rhj_exchange_wait makes `stream` wait until THIS rank's stores of relation `which` into its
 * peers' windows are complete (MPI_Win_flush_local_all, data/Window.cpp:193-201) and -- when
 * rhj_exchange_has_flags() -- until every peer's arrival flag for the relation is set, which is the barrier
 * of operators/HashJoin.cpp:119-121; without flags the launcher's own barrier goes where [barrier] stands.
 * The time between this call and the relation's local pass is reported as rhj_timings.snetcompl_us. */
int rhj_exchange_wait(rhj_ctx *ctx, int which, void *stream);
/* Host-blocking form of rhj_exchange_wait, for launchers whose barrier is a host-side one
 * (MPI_Barrier, gloo). */
int rhj_exchange_sync(rhj_ctx *ctx, int which);
int rhj_local_partition_rel(rhj_ctx *ctx, int which, void *stream);
/* tasks/LocalPartitioning.cpp:138-192 results for owned partition `p` (host copies). */
int rhj_get_local_partition(rhj_ctx *ctx, int which, uint32_t p, uint64_t *hist_host, uint64_t *prefix_host,
                            void *stream);
/* device address + tuple count of what LocalPartitioning wrote for owned partition p
 * (count + 8 * L slots, as tasks/LocalPartitioning.cpp:79-81), or of the window partition itself. */
int rhj_get_partition_dev(rhj_ctx *ctx, int which, uint32_t p, int after_local_pass, void **out_dev,
                          uint64_t *out_tuples);

/* TASK_BUILD_PROBE -- tasks/BuildProbe.cpp:47-121 for every (sub-)partition pair this rank owns. */
int rhj_build_probe(rhj_ctx *ctx, void *stream);
/* The same, adding to the result of the build/probe calls before it instead of starting from zero: for a
 * probe side that is joined piece by piece against an inner relation partitioned once (SURVEY.md 8(f4);
 * the un-built kernel library planned this, data/data.hpp:17-20 MAX_CHUNK_SIZE, :83 iterCount).  Sequence:
 * the inner relation's rhj_histogram_rel(0) .. rhj_local_partition_rel(0) once, then per outer piece
 * rhj_histogram_rel(1, piece) -> rhj_compute_offsets_rel(1) -> rhj_network_partition_rel(1) ->
 * rhj_exchange_wait(1) -> rhj_local_partition_rel(1) -> rhj_build_probe (first piece) or
 * rhj_build_probe_accumulate (later pieces); rhj_get_result then holds the whole join.  Only the inner
 * relation and ONE outer piece have to fit the planned maxima and windows. */
int rhj_build_probe_accumulate(rhj_ctx *ctx, void *stream);
/* The same stage with its output materialised: where the reference's probe loop
 * does `++matches` (tasks/BuildProbe.cpp:101-104; the unbuilt GPU probe kernels write pairs,
 * operators/gpu/kernels.cu:237-241), every match also appends one rhj_pair to `pairs_dev`
 * (device memory of this rank, 16-byte aligned, room for `capacity` pairs).  The order of the
 * pairs is unspecified.  rhj_get_result reports the full count and checksum as after
 * rhj_build_probe; min(matches, capacity) pairs have been written, the rest are only counted.
 * Either build/probe entry point may be called again on the same partitions (count first, size
 * the buffer, then materialise). */
int rhj_build_probe_pairs(rhj_ctx *ctx, rhj_pair *pairs_dev, uint64_t capacity, void *stream);

/* Result of this rank (HashJoin::RESULT_COUNTER + checksum); synchronises `stream`
 * and reports device-side capacity errors. */
int rhj_get_result(rhj_ctx *ctx, rhj_result *out, void *stream);
int rhj_get_timings(rhj_ctx *ctx, rhj_timings *out);

/* ---- the operator ------------------------------------------------------------
 * HashJoin::join() for nranks == 1: all stages back to back on `stream`, no host
 * synchronisation until the result is read.  Inputs are device-resident. */
int rhj_join(rhj_ctx *ctx, const void *inner_dev, uint64_t inner_n, const void *outer_dev, uint64_t outer_n,
             void *stream, rhj_result *out);
/* Same, with HOST buffers: stages the relations through the pool (pinned-or-pageable
 * host memory -> device) inside the call.  This is the end-to-end entry a reference
 * user would bind in place of GPUWrapper / simple_hash_join_eth
 * (tasks/gpu/GPUWrapper.cu:38-64, operators/gpu/eth.cuh:26-34). */
int rhj_join_host(rhj_ctx *ctx, const void *inner_host, uint64_t inner_n, const void *outer_host, uint64_t outer_n,
                  void *stream, rhj_result *out);

/* The staging half of rhj_join_host for any number of ranks: copies this rank's host relations
 * (pinned or pageable) into the device staging buffers rhj_plan reserved (RHJ_FLAG_STAGE_HOST_INPUTS)
 * on `stream` and returns their device addresses, which the caller then hands to rhj_histogram.
 * Replaces the per-partition PCIe upload of the reference's GPU seam (tasks/gpu/GPUWrapper.cu:38-64 ->
 * operators/gpu/eth.cu:180-193) for multi-rank launchers; rhj_timings.h2d_us covers the copies. */
int rhj_stage_host_inputs(rhj_ctx *ctx, const void *inner_host, uint64_t inner_n, const void *outer_host,
                          uint64_t outer_n, void *stream, void **inner_dev, void **outer_dev);

/* ---- relation generators on the device (data/Relation.cpp:63-85 stand-ins for sizes
 * where host rand() Fisher-Yates is impractical; identical to oracle/orc_fill_perm) ---- */
int rhj_fill_perm(void *tuples_dev, uint32_t tuple_bytes, uint64_t n, uint64_t first, uint64_t domain,
                  uint64_t seed, uint64_t modulo, uint64_t start_key, uint64_t start_rid, void *stream);

/* The same bijection applied in place to values that are already on the device:
 * vals[i * stride_words] <- perm(vals[i * stride_words]), values in [0, domain).  Maps the ranks of a
 * skewed distribution to keys as oracle/orc_fill_zipf does. */
int rhj_apply_perm(uint64_t *vals_dev, uint64_t n, uint32_t stride_words, uint64_t domain, uint64_t seed, void *stream);

/* ---- local histogram without any allocation ----------------------------------------
 * No counterpart in the reference (its GPU seam mallocs and frees on every call,
 * operators/gpu/eth.cu:144-156, 228-230).
 *
 * Who needs it: an engine that has already handed all of the GPU's memory to its own
 * buffer manager (a later cudaMalloc from a library would simply fail), and callers on a
 * latency-critical path, where cudaMalloc/cudaFree may stall on everything else the
 * device is doing.  Neither can use rhj_init + rhj_plan + rhj_histogram, which take
 * their scratch from a cudaMalloc'ed pool.
 *
 * What it does: the same result as rhj_histogram for ONE relation -- u64[P] counts of
 * key & (P-1), P = 1 << network_fanout_bits (histograms/LocalHistogram.cpp:35-53) --
 * written to hist_out_dev (caller-owned device memory, P * 8 bytes).  The caller passes
 * inputs and output; the per-chunk counters the kernels need live in a fixed
 * 4 MiB + 256 B array inside the library, present on every GPU that loads the library
 * until the process exits.  No cudaMalloc, cudaFree or other allocating call is made.
 *
 * Limits and rules:
 *  - ceil(n / 65536) * P must not exceed 1,048,576 (e.g. 268,435,456 tuples at 8 radix
 *    bits, 16,777,216 at 12); larger inputs are refused with RHJ_ERR_CAPACITY.
 *  - runs on the calling thread's current device, on `stream`.
 *  - the first call on each device looks up the array's device address; make that call
 *    before starting any stream capture.
 *  - the array is part of the library's device code.  The CUDA runtime loads that code onto a
 *    device the first time the device uses this library (lazy module loading, the default since
 *    CUDA 12), and that load is the one moment this entry point does need free device memory:
 *    4 MiB for the array plus the kernels' code.  An engine that is about to take all of the
 *    GPU's memory should therefore make one call per device first (n = 0 is enough and launches
 *    almost nothing), or run with CUDA_MODULE_LOADING=EAGER; after that no call allocates.
 *  - all calls on a device share the array: calls whose GPU work can overlap (different
 *    streams or host threads) race with each other; order them on one stream or
 *    synchronise between them.  The array keeps the previous call's contents; the
 *    entry point itself resets what has to start from zero. */
int rhj_local_histogram_noalloc(const void *tuples_dev, uint32_t tuple_bytes, uint64_t n,
                                uint32_t network_fanout_bits, uint64_t *hist_out_dev, void *stream);

/* ---- small CUDA runtime conveniences so that a C / ctypes caller need not link cudart:
 * pinned host buffers (what MPI_Alloc_mem is to the reference's windows, data/Window.cpp:35)
 * and blocking copies. */
int rhj_host_alloc(void **out_host, uint64_t bytes);
int rhj_host_free(void *host);
int rhj_memcpy_d2h(void *dst_host, const void *src_dev, uint64_t bytes);
int rhj_memcpy_h2d(void *dst_dev, const void *src_host, uint64_t bytes);

/* ---- host-side planning logic, usable without a GPU (tests, multi-rank launchers) ----
 * AssignmentMap.cpp:34-49 and OffsetMap.cpp:59-93 on host arrays; locals = u64[nranks][P]. */
void rhj_host_assignment(uint32_t nranks, uint32_t parts, uint32_t *assignment);
void rhj_host_offsets(const uint64_t *locals, uint32_t nranks, uint32_t rank, uint32_t parts, uint64_t *global,
                      uint64_t *base, uint64_t *rel, uint64_t *abs, uint64_t *window_sizes);
/* The same under RHJ_FLAG_BALANCED_ASSIGNMENT: the assignment from the two global histograms
 * (u64[P] each), and the offsets of one relation for a given assignment. */
void rhj_host_assignment_balanced(const uint64_t *global_inner, const uint64_t *global_outer, uint32_t nranks,
                                  uint32_t parts, uint32_t *assignment);
void rhj_host_offsets_assigned(const uint64_t *locals, const uint32_t *assignment, uint32_t nranks, uint32_t rank,
                               uint32_t parts, uint64_t *global, uint64_t *base, uint64_t *rel, uint64_t *abs,
                               uint64_t *window_sizes);

#ifdef __cplusplus
}
#endif
#endif /* RHJ_H_ */
