// rhj_api.cu -- the extern "C" boundary declared in include/rhj.h: context, device
// memory pool (memory/Pool.cpp), receive windows (data/Window.cpp) and the launch
// sequences of the four tasks.  Host code here only sizes buffers and launches
// kernels; every per-tuple operation lives in rhj_kernels.cuh.
#include "../../include/rhj.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "rhj_kernels.cuh"

using namespace rhj;

// -----------------------------------------------------------------------------------
// errors
// -----------------------------------------------------------------------------------
static thread_local std::string g_last_error;
static unsigned long long g_kernel_launches = 0;  // kernels of this library launched by this process

static int fail(int code, const char *fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_last_error = buf;
  return code;
}

#define RHJ_CUDA(expr)                                                                         \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess)                                                                     \
      return fail(RHJ_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)

#define RHJ_CHECK_LAUNCH(name)                                                                  \
  do {                                                                                         \
    cudaError_t _e = cudaGetLastError();                                                       \
    if (_e != cudaSuccess)                                                                     \
      return fail(RHJ_ERR_CUDA, "launch of %s failed: %s", name, cudaGetErrorString(_e));      \
    ++g_kernel_launches;                                                                       \
  } while (0)

extern "C" const char *rhj_last_error(void) { return g_last_error.c_str(); }
extern "C" uint64_t rhj_launch_count(void) { return g_kernel_launches; }
extern "C" const char *rhj_version(void) { return "rhj_b200 0.1 (sm_100a)"; }

// -----------------------------------------------------------------------------------
// context
// -----------------------------------------------------------------------------------
static constexpr int kMaxRanks = 64;
static constexpr uint32_t kDefaultChunk = 1u << 16;
static constexpr uint32_t kDefaultBuildCap = 5120;  // 4096 + head-room: see bucket_slots()
static constexpr uint64_t kAlign = 256;
static constexpr uint64_t kItemSlack = 65536;  // work-list entries beyond one per sub-partition

enum Ev { EV_START, EV_HIST, EV_NET, EV_LOCAL, EV_BP, EV_H2D0, EV_COUNT };

struct rhj_ctx {
  int rank = 0, nranks = 1, device = 0, sm_count = 148;
  rhj_config cfg{};
  uint32_t P = 0, L = 1, key_shift = 0, shift_bits = 0, nowned = 0, NS = 0;
  uint32_t chunk1 = 0, chunk2 = 0, cap_r = 0, sslice = 0;

  // pool (memory/Pool.cpp)
  char *pool = nullptr;
  uint64_t pool_bytes = 0, pool_used = 0;

  // plan
  bool planned = false;
  uint64_t max_local[2] = {0, 0}, window_cap[2] = {0, 0};
  uint32_t max_chunks1[2] = {0, 0}, max_chunks2[2] = {0, 0};
  uint64_t *window[2] = {nullptr, nullptr};
  uint64_t *localout[2] = {nullptr, nullptr};
  void *staged_input[2] = {nullptr, nullptr};
  uint32_t *chunk1_hist[2] = {nullptr, nullptr};
  uint32_t *chunk2_hist[2] = {nullptr, nullptr};
  uint64_t *d_local_hist = nullptr, *d_global = nullptr, *d_base = nullptr, *d_rel = nullptr, *d_abs = nullptr,
           *d_dest = nullptr, *d_winsizes = nullptr;
  uint32_t *d_assign = nullptr;
  uint64_t *d_part_off = nullptr, *d_part_cnt = nullptr;
  uint32_t *d_cstart = nullptr;
  uint64_t *d_sub_tot = nullptr, *d_sub_off = nullptr, *d_sub_prefix = nullptr, *d_sub_addr = nullptr;
  uint32_t *d_sub_cnt = nullptr;
  uint64_t *d_item_lstart = nullptr, *d_item_jcount = nullptr, *d_item_jstart = nullptr;
  unsigned long long *d_work = nullptr, *d_result = nullptr;
  uint32_t *d_owned_list = nullptr, *d_slot_of = nullptr;  // partition <-> (owner, slot), see PlanNetworkArgs
  int *d_status = nullptr;
  ItemDesc *d_items = nullptr;
  uint64_t items_cap = 0;
  // staged exchange (nranks > 1)
  uint64_t *sendbuf[2] = {nullptr, nullptr};
  uint64_t *d_push_src = nullptr, *d_push_dst = nullptr, *d_push_cnt = nullptr;
  cudaStream_t side = nullptr;
  cudaEvent_t ev_scattered[2] = {nullptr, nullptr}, ev_pushed[2] = {nullptr, nullptr}, ev_push_begin[2] = {nullptr, nullptr};
  bool rel_partitioned[2] = {false, false}, rel_scattered[2] = {false, false}, wait_timed[2] = {false, false};
  bool hist_done[2] = {false, false}, offsets_done[2] = {false, false};
  cudaEvent_t ev_rel_net[2] = {nullptr, nullptr}, ev_rel_local[2] = {nullptr, nullptr};
  cudaEvent_t ev_wait0[2] = {nullptr, nullptr}, ev_wait1[2] = {nullptr, nullptr};
  bool staged_exchange = false;
  int reserve_sms = 0, tune_push_sms = 32;

  // peers (data/Window.cpp: target of MPI_Put)
  uint32_t *d_flags = nullptr;           // [2][64] arrival flags written by the peers
  uint64_t peer_flags[kMaxRanks] = {};   // the peers' flag arrays (0: none, e.g. ranks emulated on one device)
  uint32_t epoch = 0;                    // joins started on this context
  uint32_t epoch_rel[2] = {0, 0};        // network passes started per relation; the value that relation's flags carry
  bool use_flags = false;
  uint64_t peer_window[2][kMaxRanks] = {};
  void *ipc_base[kMaxRanks] = {};

  // per-join state
  const void *input[2] = {nullptr, nullptr};
  uint64_t input_n[2] = {0, 0};
  uint32_t nchunks1[2] = {0, 0};
  int stage = 0;  // 0 idle, 1 histogram done, 2 offsets, 3 network, 4 local, 5 build-probe
  bool local_planned = false;
  cudaEvent_t ev[EV_COUNT] = {};
  bool ev_valid[EV_COUNT] = {};
  bool staged_pending = false;  // rhj_stage_host_inputs ran since the last rhj_histogram
  bool h2d_pipelined = false;   // the last join's inputs came through rhj_join_host's copy stream
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_piece[128] = {};
  cudaEvent_t ev_h2d_done = nullptr;

  // kernel choice and ablation knobs (rhj_init; environment only with RHJ_FLAG_DEBUG_ENV)
  int tune_scatter = 3;  // 0 first generation (k_scatter), 3 write-combining rings (default)
  bool tune_legacy_bp = false;
  int tune_bp_tpb = 512;
  int tune_wc_variant = 0;
  int tune_push_priority = 1;
  uint64_t tune_piece_bytes = 256ull << 20;  // host->device piece of rhj_join_host
  int tune_wc_l2_net = 4, tune_wc_l2_loc = 4;  // L2 prefetch distance of the ring scatter, quarter batches
};

static void invalidate_plan(rhj_ctx *ctx);
static inline uint64_t align_up(uint64_t v, uint64_t a) { return (v + a - 1) / a * a; }

static int validate_config(const rhj_config *c, int rank, int nranks) {
  if (!c) return fail(RHJ_ERR_INVALID, "config is NULL");
  if (nranks < 1 || nranks > kMaxRanks) return fail(RHJ_ERR_INVALID, "nranks %d outside [1,%d]", nranks, kMaxRanks);
  if (rank < 0 || rank >= nranks) return fail(RHJ_ERR_INVALID, "rank %d outside [0,%d)", rank, nranks);
  if (c->network_fanout_bits < 1 || c->network_fanout_bits > 12)
    return fail(RHJ_ERR_INVALID, "network_fanout_bits %u outside [1,12]", c->network_fanout_bits);
  if (c->local_fanout_bits < 1 || c->local_fanout_bits > 12)
    return fail(RHJ_ERR_INVALID, "local_fanout_bits %u outside [1,12]", c->local_fanout_bits);
  uint32_t ks = c->network_fanout_bits + c->payload_bits;
  if (ks < 1 || ks + c->local_fanout_bits > 63)
    return fail(RHJ_ERR_INVALID, "network_fanout_bits + payload_bits + local_fanout_bits = %u must be in [2,63]",
                ks + c->local_fanout_bits);
  if (c->tuple_bytes != 16 && c->tuple_bytes != 8) return fail(RHJ_ERR_INVALID, "tuple_bytes must be 16 or 8");
  if (c->allocation_factor != 0 && c->allocation_factor < 1.0)
    return fail(RHJ_ERR_INVALID, "allocation_factor must be >= 1.0");
  uint32_t cap = c->build_capacity ? c->build_capacity : kDefaultBuildCap;
  if (cap != 1024 && cap != 2048 && cap != 4096 && cap != 5120 && cap != 8192)
    return fail(RHJ_ERR_INVALID, "build_capacity must be 1024, 2048, 4096, 5120 or 8192");
  uint32_t chunk = c->chunk_tuples ? c->chunk_tuples : kDefaultChunk;
  if (chunk % kChunkGranule != 0) return fail(RHJ_ERR_INVALID, "chunk_tuples must be a multiple of %d", kChunkGranule);
  return RHJ_OK;
}

static bool exchange_is_staged(const rhj_config *cfg, int nranks) {
  if (nranks < 2 || (cfg->flags & RHJ_FLAG_DIRECT_EXCHANGE)) return false;
  return nranks > 2 || (cfg->flags & RHJ_FLAG_STAGED_EXCHANGE) != 0;
}

// position of a rank's first partition in the owner-major partition list (k_plan_network)
static uint32_t first_owned_slot(uint32_t P, int rank, int nranks) {
  const uint32_t q = P / nranks, m = P % nranks;
  return (uint32_t)rank * q + ((uint32_t)rank < m ? (uint32_t)rank : m);
}

static uint32_t owned_partitions(uint32_t P, int rank, int nranks) {
  return (uint32_t)rank < P ? (P - rank + nranks - 1) / nranks : 0;
}

extern "C" int rhj_init(rhj_ctx **out, int rank, int nranks, int device, const rhj_config *cfg) {
  if (!out) return fail(RHJ_ERR_INVALID, "out is NULL");
  *out = nullptr;
  int rc = validate_config(cfg, rank, nranks);
  if (rc) return rc;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(RHJ_ERR_CUDA, "no CUDA device available (%s); this library has no CPU path",
                e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
  if (device < 0 || device >= ndev) return fail(RHJ_ERR_INVALID, "device %d outside [0,%d)", device, ndev);
  RHJ_CUDA(cudaSetDevice(device));
  rhj_ctx *c = new rhj_ctx();
  struct Guard {  // a failed rhj_init leaves nothing behind
    rhj_ctx *c;
    ~Guard() { if (c) rhj_destroy(c); }
  } guard{c};
  c->rank = rank;
  c->nranks = nranks;
  c->device = device;
  c->cfg = *cfg;
  if (c->cfg.allocation_factor == 0) c->cfg.allocation_factor = 1.1;  // core/Configuration.h:35
  c->P = 1u << cfg->network_fanout_bits;
  c->L = cfg->enable_two_level ? (1u << cfg->local_fanout_bits) : 1u;
  c->key_shift = cfg->network_fanout_bits + cfg->payload_bits;
  c->shift_bits = c->key_shift + cfg->local_fanout_bits;  // tasks/BuildProbe.cpp:55-57
  c->nowned = owned_partitions(c->P, rank, nranks);
  c->NS = c->nowned * c->L;
  c->chunk1 = c->chunk2 = cfg->chunk_tuples ? cfg->chunk_tuples : kDefaultChunk;
  c->cap_r = cfg->build_capacity ? cfg->build_capacity : kDefaultBuildCap;
  c->sslice = 4 * c->cap_r;
  cudaDeviceProp prop;
  RHJ_CUDA(cudaGetDeviceProperties(&prop, device));
  c->sm_count = prop.multiProcessorCount;
  for (int i = 0; i < EV_COUNT; ++i) RHJ_CUDA(cudaEventCreate(&c->ev[i]));
  for (int k = 0; k < 2; ++k) {
    RHJ_CUDA(cudaEventCreateWithFlags(&c->ev_rel_net[k], cudaEventDisableTiming));
    RHJ_CUDA(cudaEventCreateWithFlags(&c->ev_rel_local[k], cudaEventDisableTiming));
    RHJ_CUDA(cudaEventCreate(&c->ev_wait0[k]));
    RHJ_CUDA(cudaEventCreate(&c->ev_wait1[k]));
  }
  // Kernel choice by fan-out.  The write-combining scatter keeps one ring per bin in shared memory,
  // which stops fitting above 1024 bins; wider fan-outs run on the first-generation pair
  // (k_histogram + k_scatter: one staged tile, per-tile scan).
  c->tune_scatter = (c->P <= 1024 && (!cfg->enable_two_level || c->L <= 1024)) ? 3 : 0;
  // Exchange: with two ranks the scatter kernel's full-line stores go straight into the peer's window;
  // from three ranks on every SM would be storing single lines into hundreds of regions spread over all
  // peers at once, which the links sustain at a fraction of their rate (measured on 8 GPUs: 135 GB/s per
  // direction against 560 GB/s on two), so the remote partitions are gathered in a local send buffer and
  // pushed as long sequential runs.
  c->staged_exchange = exchange_is_staged(cfg, nranks);
  if (cfg->flags & RHJ_FLAG_DEBUG_ENV) {
    // Ablation switches; only read when the caller asks for them.
    if (const char *e = getenv("RHJ_SCATTER"))
      if (strcmp(e, "legacy") == 0) c->tune_scatter = 0;
    if (const char *e = getenv("RHJ_BP")) c->tune_legacy_bp = strcmp(e, "legacy") == 0;
    if (const char *e = getenv("RHJ_BP_TPB")) c->tune_bp_tpb = atoi(e) == 512 ? 512 : 1024;
    if (const char *e = getenv("RHJ_PUSH_SMS")) c->tune_push_sms = std::max(0, std::min(64, atoi(e)));
    if (const char *e = getenv("RHJ_WC")) c->tune_wc_variant = atoi(e);
    if (const char *e = getenv("RHJ_PUSH_PRIORITY")) c->tune_push_priority = atoi(e);
    if (const char *e = getenv("RHJ_H2D_PIECE_KB")) c->tune_piece_bytes = (uint64_t)std::max(1, atoi(e)) << 10;
    if (const char *e = getenv("RHJ_WC_L2_NET")) c->tune_wc_l2_net = std::max(0, std::min(64, atoi(e)));
    if (const char *e = getenv("RHJ_WC_L2_LOC")) c->tune_wc_l2_loc = std::max(0, std::min(64, atoi(e)));
  }
  if (c->staged_exchange) {
    // the pushes are what NVLink waits for: their CTAs are placed before those of a scatter or local-pass
    // launch that becomes ready at the same time
    int prio_lo = 0, prio_hi = 0;  // numerically lower = higher priority
    RHJ_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    RHJ_CUDA(cudaStreamCreateWithPriority(&c->side, cudaStreamNonBlocking, c->tune_push_priority ? prio_hi : prio_lo));
    for (int k = 0; k < 2; ++k) {
      RHJ_CUDA(cudaEventCreateWithFlags(&c->ev_scattered[k], cudaEventDisableTiming));
      RHJ_CUDA(cudaEventCreate(&c->ev_pushed[k]));
      RHJ_CUDA(cudaEventCreate(&c->ev_push_begin[k]));
    }
  }
  guard.c = nullptr;
  *out = c;
  return RHJ_OK;
}

// everything carved from the pool is gone: no stage of an earlier plan may run any more
static void invalidate_plan(rhj_ctx *ctx) {
  ctx->planned = false;
  ctx->stage = 0;
  ctx->local_planned = false;
  for (int k = 0; k < 2; ++k)
    ctx->hist_done[k] = ctx->offsets_done[k] = ctx->rel_scattered[k] = ctx->rel_partitioned[k] = false;
}

extern "C" int rhj_pool_free_all(rhj_ctx *ctx) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();  // kernels of the last join may still use the pool
  if (ctx->pool) cudaFree(ctx->pool);
  ctx->pool = nullptr;
  ctx->pool_bytes = ctx->pool_used = 0;
  invalidate_plan(ctx);
  return RHJ_OK;
}

extern "C" int rhj_destroy(rhj_ctx *ctx) {
  if (!ctx) return RHJ_OK;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  for (int r = 0; r < kMaxRanks; ++r)
    if (ctx->ipc_base[r]) cudaIpcCloseMemHandle(ctx->ipc_base[r]);
  rhj_pool_free_all(ctx);
  for (int i = 0; i < EV_COUNT; ++i)
    if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
  for (int k = 0; k < 2; ++k) {
    if (ctx->ev_rel_net[k]) cudaEventDestroy(ctx->ev_rel_net[k]);
    if (ctx->ev_rel_local[k]) cudaEventDestroy(ctx->ev_rel_local[k]);
    if (ctx->ev_wait0[k]) cudaEventDestroy(ctx->ev_wait0[k]);
    if (ctx->ev_wait1[k]) cudaEventDestroy(ctx->ev_wait1[k]);
    if (ctx->ev_scattered[k]) cudaEventDestroy(ctx->ev_scattered[k]);
    if (ctx->ev_pushed[k]) cudaEventDestroy(ctx->ev_pushed[k]);
    if (ctx->ev_push_begin[k]) cudaEventDestroy(ctx->ev_push_begin[k]);
  }
  if (ctx->side) cudaStreamDestroy(ctx->side);
  if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
  for (int i = 0; i < 128; ++i)
    if (ctx->ev_piece[i]) cudaEventDestroy(ctx->ev_piece[i]);
  if (ctx->ev_h2d_done) cudaEventDestroy(ctx->ev_h2d_done);
  delete ctx;
  return RHJ_OK;
}

extern "C" int rhj_pool_allocate(rhj_ctx *ctx, uint64_t bytes) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  if (ctx->pool) return fail(RHJ_ERR_STATE, "pool already allocated (%llu bytes)", (unsigned long long)ctx->pool_bytes);
  RHJ_CUDA(cudaSetDevice(ctx->device));
  bytes = align_up(bytes ? bytes : kAlign, kAlign);
  void *p = nullptr;
  cudaError_t e = cudaMalloc(&p, bytes);
  if (e != cudaSuccess)
    return fail(RHJ_ERR_NOMEM, "cudaMalloc of %llu pool bytes failed: %s", (unsigned long long)bytes, cudaGetErrorString(e));
  ctx->pool = static_cast<char *>(p);
  ctx->pool_bytes = bytes;
  ctx->pool_used = 0;
  return RHJ_OK;
}

extern "C" int rhj_pool_get_memory(rhj_ctx *ctx, uint64_t bytes, void **out_dev) {
  if (!ctx || !out_dev) return fail(RHJ_ERR_INVALID, "NULL argument");
  if (!ctx->pool) return fail(RHJ_ERR_STATE, "pool not allocated");
  uint64_t need = align_up(bytes ? bytes : 1, kAlign);
  if (ctx->pool_used + need > ctx->pool_bytes)
    return fail(RHJ_ERR_NOMEM, "pool exhausted: %llu used + %llu requested > %llu", (unsigned long long)ctx->pool_used,
                (unsigned long long)need, (unsigned long long)ctx->pool_bytes);
  *out_dev = ctx->pool + ctx->pool_used;
  ctx->pool_used += need;
  return RHJ_OK;
}

extern "C" int rhj_pool_reset(rhj_ctx *ctx) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  ctx->pool_used = 0;
  invalidate_plan(ctx);
  return RHJ_OK;
}

// -----------------------------------------------------------------------------------
// plan: sizes of every intermediate
// -----------------------------------------------------------------------------------
struct Layout {
  uint64_t window_cap[2], localout_tuples[2];
  uint32_t max_chunks1[2], max_chunks2[2];
  uint64_t total;
};

static uint64_t default_window_cap(const rhj_config *cfg, uint64_t max_local) {
  double f = cfg->allocation_factor == 0 ? 1.1 : cfg->allocation_factor;
  return (uint64_t)((double)max_local * f) + 1024;
}

static void compute_layout(const rhj_config *cfg, int rank, int nranks, const uint64_t max_local[2],
                           const uint64_t wcap_in[2], bool staged, Layout *lay) {
  const uint32_t P = 1u << cfg->network_fanout_bits;
  const uint32_t L = cfg->enable_two_level ? (1u << cfg->local_fanout_bits) : 1u;
  const uint32_t nowned = owned_partitions(P, rank, nranks);
  const uint64_t NS = (uint64_t)nowned * L;
  const uint32_t chunk = cfg->chunk_tuples ? cfg->chunk_tuples : kDefaultChunk;
  uint64_t total = 0;
  auto add = [&](uint64_t bytes) { total += align_up(bytes ? bytes : 1, kAlign); };
  for (int k = 0; k < 2; ++k) {
    lay->window_cap[k] = wcap_in[k] ? wcap_in[k] : default_window_cap(cfg, max_local[k]);
    lay->localout_tuples[k] = cfg->enable_two_level ? lay->window_cap[k] + NS * 8 + (uint64_t)nowned * 8 + 16 : 0;
    lay->max_chunks1[k] = (uint32_t)std::max<uint64_t>(1, (max_local[k] + chunk - 1) / chunk);
    lay->max_chunks2[k] = cfg->enable_two_level ? (uint32_t)((lay->window_cap[k] + chunk - 1) / chunk + nowned) : 0;
    add(lay->window_cap[k] * 8);
    if (cfg->enable_two_level) add(lay->localout_tuples[k] * 8);
    if (cfg->flags & RHJ_FLAG_STAGE_HOST_INPUTS) add(max_local[k] * cfg->tuple_bytes);
    add((uint64_t)lay->max_chunks1[k] * P * 4);
    if (cfg->enable_two_level) add((uint64_t)lay->max_chunks2[k] * L * 4);
  }
  for (int i = 0; i < 6; ++i) add(2ull * P * 8);   // local, global, base, rel, abs, dest
  add((uint64_t)P * 4);                            // assignment
  add((uint64_t)P * 4); add((uint64_t)P * 4);      // owned_list, slot_of
  add(2ull * nranks * 8);                          // window sizes
  add(2ull * nowned * 8); add(2ull * nowned * 8);  // part_off, part_cnt
  add(2ull * (nowned + 1) * 4);                    // cstart
  for (int i = 0; i < 4; ++i) add(2 * NS * 8);     // sub_tot, sub_off, sub_prefix, sub_addr
  add(2 * NS * 4);                                 // sub_cnt
  add(NS * 8); add((uint64_t)nowned * 8); add(((uint64_t)nowned + 1) * 8);  // item_lstart, jcount, jstart
  add(16); add(5 * 8); add(4);                     // work counter + pair cursor, result, status
  add(2 * 64 * 4);                                 // arrival flags
  add((NS + kItemSlack) * sizeof(ItemDesc));       // expanded build/probe work list
  if (staged) {                                    // send buffers + push lists of the staged exchange
    add(max_local[0] * 8); add(max_local[1] * 8);
    for (int i = 0; i < 3; ++i) add(2ull * P * 8);
  }
  lay->total = total;
}

extern "C" uint64_t rhj_workspace_bytes(const rhj_config *cfg, int nranks, uint64_t max_local_inner,
                                        uint64_t max_local_outer, uint64_t window_inner_cap, uint64_t window_outer_cap) {
  if (!cfg || validate_config(cfg, 0, nranks)) return 0;
  Layout lay;
  uint64_t ml[2] = {max_local_inner, max_local_outer}, wc[2] = {window_inner_cap, window_outer_cap};
  compute_layout(cfg, 0, nranks, ml, wc, exchange_is_staged(cfg, nranks), &lay);  // rank 0 owns the most partitions
  return lay.total;
}

template <typename T>
static int carve(rhj_ctx *ctx, uint64_t count, T **out) {
  void *p = nullptr;
  int rc = rhj_pool_get_memory(ctx, count * sizeof(T), &p);
  *out = static_cast<T *>(p);
  return rc;
}

extern "C" int rhj_plan(rhj_ctx *ctx, uint64_t max_local_inner, uint64_t max_local_outer, uint64_t window_inner_cap,
                        uint64_t window_outer_cap) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  if (ctx->planned) return fail(RHJ_ERR_STATE, "already planned; rhj_pool_reset first");
  RHJ_CUDA(cudaSetDevice(ctx->device));
  uint64_t ml[2] = {max_local_inner, max_local_outer}, wc[2] = {window_inner_cap, window_outer_cap};
  Layout lay;
  compute_layout(&ctx->cfg, ctx->rank, ctx->nranks, ml, wc, ctx->staged_exchange, &lay);
  for (int k = 0; k < 2; ++k) {
    if (ml[k] >= (1ull << 32) || lay.window_cap[k] >= (1ull << 32))
      return fail(RHJ_ERR_INVALID, "relation %d: per-rank sizes must stay below 2^32 tuples", k);
  }
  if (!ctx->pool) {
    int rc = rhj_pool_allocate(ctx, lay.total);
    if (rc) return rc;
  }
  if (ctx->pool_bytes - ctx->pool_used < lay.total)
    return fail(RHJ_ERR_NOMEM, "pool has %llu free bytes, plan needs %llu", (unsigned long long)(ctx->pool_bytes - ctx->pool_used),
                (unsigned long long)lay.total);
  const uint32_t P = ctx->P, L = ctx->L, nowned = ctx->nowned;
  const uint64_t NS = ctx->NS;
  int rc = 0;
  for (int k = 0; k < 2 && !rc; ++k) {
    ctx->max_local[k] = ml[k];
    ctx->window_cap[k] = lay.window_cap[k];
    ctx->max_chunks1[k] = lay.max_chunks1[k];
    ctx->max_chunks2[k] = lay.max_chunks2[k];
    rc = carve(ctx, lay.window_cap[k], &ctx->window[k]);
    if (!rc && ctx->cfg.enable_two_level) rc = carve(ctx, lay.localout_tuples[k], &ctx->localout[k]);
    if (!rc && (ctx->cfg.flags & RHJ_FLAG_STAGE_HOST_INPUTS)) {
      char *p = nullptr;
      rc = carve(ctx, ml[k] * ctx->cfg.tuple_bytes, &p);
      ctx->staged_input[k] = p;
    }
    if (!rc) rc = carve(ctx, (uint64_t)lay.max_chunks1[k] * P, &ctx->chunk1_hist[k]);
    if (!rc && ctx->cfg.enable_two_level) rc = carve(ctx, (uint64_t)lay.max_chunks2[k] * L, &ctx->chunk2_hist[k]);
  }
  if (!rc) rc = carve(ctx, 2ull * P, &ctx->d_local_hist);
  if (!rc) rc = carve(ctx, 2ull * P, &ctx->d_global);
  if (!rc) rc = carve(ctx, 2ull * P, &ctx->d_base);
  if (!rc) rc = carve(ctx, 2ull * P, &ctx->d_rel);
  if (!rc) rc = carve(ctx, 2ull * P, &ctx->d_abs);
  if (!rc) rc = carve(ctx, 2ull * P, &ctx->d_dest);
  if (!rc) rc = carve(ctx, (uint64_t)P, &ctx->d_assign);
  if (!rc) rc = carve(ctx, (uint64_t)P, &ctx->d_owned_list);
  if (!rc) rc = carve(ctx, (uint64_t)P, &ctx->d_slot_of);
  if (!rc) rc = carve(ctx, 2ull * ctx->nranks, &ctx->d_winsizes);
  if (!rc) rc = carve(ctx, 2ull * nowned, &ctx->d_part_off);
  if (!rc) rc = carve(ctx, 2ull * nowned, &ctx->d_part_cnt);
  if (!rc) rc = carve(ctx, 2ull * (nowned + 1), &ctx->d_cstart);
  if (!rc) rc = carve(ctx, 2 * NS, &ctx->d_sub_tot);
  if (!rc) rc = carve(ctx, 2 * NS, &ctx->d_sub_off);
  if (!rc) rc = carve(ctx, 2 * NS, &ctx->d_sub_prefix);
  if (!rc) rc = carve(ctx, 2 * NS, &ctx->d_sub_addr);
  if (!rc) rc = carve(ctx, 2 * NS, &ctx->d_sub_cnt);
  if (!rc) rc = carve(ctx, NS, &ctx->d_item_lstart);
  if (!rc) rc = carve(ctx, (uint64_t)nowned, &ctx->d_item_jcount);
  if (!rc) rc = carve(ctx, (uint64_t)nowned + 1, &ctx->d_item_jstart);
  if (!rc) rc = carve(ctx, 2ull, &ctx->d_work);  // [0] work counter, [1] output cursor of rhj_build_probe_pairs
  if (!rc) rc = carve(ctx, 5ull, &ctx->d_result);
  if (!rc) rc = carve(ctx, 1ull, &ctx->d_status);
  if (!rc) rc = carve(ctx, 2ull * 64, &ctx->d_flags);
  ctx->items_cap = NS + kItemSlack;
  if (!rc) rc = carve(ctx, ctx->items_cap, &ctx->d_items);
  if (ctx->staged_exchange) {
    for (int k = 0; k < 2 && !rc; ++k) rc = carve(ctx, ml[k], &ctx->sendbuf[k]);
    if (!rc) rc = carve(ctx, 2ull * P, &ctx->d_push_src);
    if (!rc) rc = carve(ctx, 2ull * P, &ctx->d_push_dst);
    if (!rc) rc = carve(ctx, 2ull * P, &ctx->d_push_cnt);
  }
  if (rc) return rc;
  RHJ_CUDA(cudaMemset(ctx->d_status, 0, sizeof(int)));
  RHJ_CUDA(cudaMemset(ctx->d_result, 0, 5 * 8));
  RHJ_CUDA(cudaMemset(ctx->d_flags, 0, 2 * 64 * 4));
  RHJ_CUDA(cudaDeviceSynchronize());  // the flags are zero before any peer can learn their address
  ctx->peer_window[0][ctx->rank] = reinterpret_cast<uint64_t>(ctx->window[0]);
  ctx->peer_window[1][ctx->rank] = reinterpret_cast<uint64_t>(ctx->window[1]);
  ctx->planned = true;
  ctx->stage = 0;
  return RHJ_OK;
}

// -----------------------------------------------------------------------------------
// windows across ranks
// -----------------------------------------------------------------------------------
extern "C" int rhj_window_export(rhj_ctx *ctx, rhj_window_handle *out) {
  if (!ctx || !out) return fail(RHJ_ERR_INVALID, "NULL argument");
  if (!ctx->planned) return fail(RHJ_ERR_STATE, "rhj_plan first");
  RHJ_CUDA(cudaSetDevice(ctx->device));
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "ipc handle size");
  cudaIpcMemHandle_t h;
  RHJ_CUDA(cudaIpcGetMemHandle(&h, ctx->pool));
  memcpy(out->ipc, &h, 64);
  out->inner_offset = reinterpret_cast<char *>(ctx->window[0]) - ctx->pool;
  out->outer_offset = reinterpret_cast<char *>(ctx->window[1]) - ctx->pool;
  out->capacity_inner = ctx->window_cap[0];
  out->capacity_outer = ctx->window_cap[1];
  out->flags_offset = reinterpret_cast<char *>(ctx->d_flags) - ctx->pool;
  return RHJ_OK;
}

extern "C" int rhj_window_import(rhj_ctx *ctx, int peer_rank, const rhj_window_handle *handle) {
  if (!ctx || !handle) return fail(RHJ_ERR_INVALID, "NULL argument");
  if (peer_rank < 0 || peer_rank >= ctx->nranks || peer_rank == ctx->rank)
    return fail(RHJ_ERR_INVALID, "peer rank %d invalid", peer_rank);
  RHJ_CUDA(cudaSetDevice(ctx->device));
  if (handle->capacity_inner != ctx->window_cap[0] || handle->capacity_outer != ctx->window_cap[1])
    return fail(RHJ_ERR_INVALID, "peer %d planned different window capacities", peer_rank);
  cudaIpcMemHandle_t h;
  memcpy(&h, handle->ipc, 64);
  if (ctx->ipc_base[peer_rank]) {  // re-import (the peer planned anew): drop the earlier mapping
    cudaIpcCloseMemHandle(ctx->ipc_base[peer_rank]);
    ctx->ipc_base[peer_rank] = nullptr;
  }
  void *base = nullptr;
  RHJ_CUDA(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
  ctx->ipc_base[peer_rank] = base;
  ctx->peer_window[0][peer_rank] = reinterpret_cast<uint64_t>(static_cast<char *>(base) + handle->inner_offset);
  ctx->peer_window[1][peer_rank] = reinterpret_cast<uint64_t>(static_cast<char *>(base) + handle->outer_offset);
  ctx->peer_flags[peer_rank] = reinterpret_cast<uint64_t>(static_cast<char *>(base) + handle->flags_offset);
  // arrival flags replace the launcher's barrier once every peer's flag array is mapped
  ctx->use_flags = true;
  for (int r = 0; r < ctx->nranks; ++r)
    if (r != ctx->rank && !ctx->peer_flags[r]) ctx->use_flags = false;
  return RHJ_OK;
}

extern "C" int rhj_window_set_peer(rhj_ctx *ctx, int peer_rank, void *inner_window_dev, void *outer_window_dev) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  if (peer_rank < 0 || peer_rank >= ctx->nranks) return fail(RHJ_ERR_INVALID, "peer rank %d invalid", peer_rank);
  ctx->peer_window[0][peer_rank] = reinterpret_cast<uint64_t>(inner_window_dev);
  ctx->peer_window[1][peer_rank] = reinterpret_cast<uint64_t>(outer_window_dev);
  return RHJ_OK;
}

extern "C" int rhj_exchange_is_staged(rhj_ctx *ctx) { return ctx && ctx->staged_exchange ? 1 : 0; }
extern "C" int rhj_exchange_has_flags(rhj_ctx *ctx) { return ctx && ctx->use_flags ? 1 : 0; }

extern "C" int rhj_window_get(rhj_ctx *ctx, void **inner_window_dev, void **outer_window_dev) {
  if (!ctx || !ctx->planned) return fail(RHJ_ERR_STATE, "rhj_plan first");
  if (inner_window_dev) *inner_window_dev = ctx->window[0];
  if (outer_window_dev) *outer_window_dev = ctx->window[1];
  return RHJ_OK;
}

// -----------------------------------------------------------------------------------
// launches
// -----------------------------------------------------------------------------------
static int record(rhj_ctx *ctx, int which, cudaStream_t st) {
  RHJ_CUDA(cudaEventRecord(ctx->ev[which], st));
  ctx->ev_valid[which] = true;
  return RHJ_OK;
}

template <class Source, class Map>
static int launch_histogram(rhj_ctx *ctx, const Source &src, const Map &cm, uint32_t grid_chunks, uint32_t nbins,
                            uint32_t *chunk_hist, cudaStream_t st) {
  if (grid_chunks == 0) return RHJ_OK;
  const bool agg = (ctx->cfg.flags & RHJ_FLAG_HIST_WARP_AGGREGATE) != 0;
  uint32_t nrep = std::max<uint32_t>(1, std::min<uint32_t>(kHistTPB / 32, 4096 / nbins));
  size_t smem = (size_t)nbins * nrep * 4;
  uint32_t grid = std::min<uint32_t>(grid_chunks, (uint32_t)std::max(1, ctx->sm_count - ctx->reserve_sms) * 4);
  if (agg)
    k_histogram<Source, Map, true><<<grid, kHistTPB, smem, st>>>(src, cm, nbins, nrep, chunk_hist, ctx->d_status);
  else
    k_histogram<Source, Map, false><<<grid, kHistTPB, smem, st>>>(src, cm, nbins, nrep, chunk_hist, ctx->d_status);
  RHJ_CHECK_LAUNCH("k_histogram");
  return RHJ_OK;
}

// Ring geometry of the write-combining scatter for a fan-out: CTA size, tuples per thread and
// batch, ring size per bin.  Two CTAs per SM where the rings allow it: while one waits at a
// barrier or for its loads the other one runs.
struct WcGeometry {
  int tpb, ipt, cfix;
  uint32_t ring;
};
static bool wc_geometry(uint32_t nbins, WcGeometry *g) {
  if (nbins > 1024) return false;
  if (nbins == 1024) *g = {1024, 4, 24, 24};
  else if (nbins == 512) *g = {512, 4, 24, 24};
  else if (nbins == 256) *g = {512, 8, 48, 48};
  else {
    uint32_t c = 16;
    while (c * 2 * nbins <= 12288 && c < 2048) c *= 2;  // power of two, at most 96 KB of rings
    *g = {512, 8, 0, c};
  }
  return true;
}

template <class Source, class Map, int TPB, int IPT, int CFIX>
static int launch_scatter_wc_t(rhj_ctx *ctx, const Source &src, const Map &cm, uint32_t grid_chunks, uint32_t nbins,
                               uint32_t ring, const uint32_t *chunk_off, const uint64_t *bin_base, cudaStream_t st,
                               uint32_t c_lo, uint32_t c_hi) {
  const size_t smem = scatter_wc_smem_bytes(nbins, ring, TPB);
  auto kern = k_scatter_wc<Source, Map, TPB, IPT, CFIX>;
  RHJ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  RHJ_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, TPB, smem));
  if (per_sm < 1) return fail(RHJ_ERR_INVALID, "scatter kernel does not fit (%zu bytes of shared memory at %u bins)", smem, nbins);
  per_sm = std::min(per_sm, TPB <= 256 ? 4 : TPB <= 512 ? 2 : 1);
  const uint32_t sms = (uint32_t)std::max(1, ctx->sm_count - ctx->reserve_sms);
  const uint32_t span = std::min(c_hi, grid_chunks) > c_lo ? std::min(c_hi, grid_chunks) - c_lo : 0u;
  if (span == 0) return RHJ_OK;
  const uint32_t grid = std::min<uint32_t>(span, sms * (uint32_t)per_sm);
  kern<<<grid, TPB, smem, st>>>(src, cm, nbins, ring, chunk_off, bin_base, ctx->d_status,
                                Source::kDigitFromValue ? ctx->tune_wc_l2_loc : ctx->tune_wc_l2_net, c_lo, c_hi);
  RHJ_CHECK_LAUNCH("k_scatter_wc");
  return RHJ_OK;
}

template <class Source, class Map>
static int launch_scatter_wc(rhj_ctx *ctx, const Source &src, const Map &cm, uint32_t grid_chunks, uint32_t nbins,
                             const uint32_t *chunk_off, const uint64_t *bin_base, cudaStream_t st, uint32_t c_lo,
                             uint32_t c_hi) {
  WcGeometry g;
  if (!wc_geometry(nbins, &g)) return fail(RHJ_ERR_INVALID, "no ring geometry for %u bins", nbins);
#ifdef RHJ_TUNE_VARIANTS
  // experiment builds only: RHJ_WC=<variant> picks another geometry for <= 256 bins
  if (nbins <= 256 && ctx->tune_wc_variant) {
    const uint32_t r32 = 32;
    switch (ctx->tune_wc_variant) {
      case 1: return launch_scatter_wc_t<Source, Map, 256, 8, 24>(ctx, src, cm, grid_chunks, nbins, 24, chunk_off, bin_base, st, c_lo, c_hi);
      case 2: return launch_scatter_wc_t<Source, Map, 256, 8, 0>(ctx, src, cm, grid_chunks, nbins, r32, chunk_off, bin_base, st, c_lo, c_hi);
      case 3: return launch_scatter_wc_t<Source, Map, 1024, 8, 96>(ctx, src, cm, grid_chunks, nbins, 96, chunk_off, bin_base, st, c_lo, c_hi);
      case 4: return launch_scatter_wc_t<Source, Map, 512, 4, 24>(ctx, src, cm, grid_chunks, nbins, 24, chunk_off, bin_base, st, c_lo, c_hi);
      case 5: return launch_scatter_wc_t<Source, Map, 256, 4, 24>(ctx, src, cm, grid_chunks, nbins, 24, chunk_off, bin_base, st, c_lo, c_hi);
      default: break;
    }
  }
#endif
  if (g.cfix == 48) return launch_scatter_wc_t<Source, Map, 512, 8, 48>(ctx, src, cm, grid_chunks, nbins, g.ring, chunk_off, bin_base, st, c_lo, c_hi);
  if (g.cfix == 24 && g.tpb == 512) return launch_scatter_wc_t<Source, Map, 512, 4, 24>(ctx, src, cm, grid_chunks, nbins, g.ring, chunk_off, bin_base, st, c_lo, c_hi);
  if (g.cfix == 24) return launch_scatter_wc_t<Source, Map, 1024, 4, 24>(ctx, src, cm, grid_chunks, nbins, g.ring, chunk_off, bin_base, st, c_lo, c_hi);
  return launch_scatter_wc_t<Source, Map, 512, 8, 0>(ctx, src, cm, grid_chunks, nbins, g.ring, chunk_off, bin_base, st, c_lo, c_hi);
}

template <class Source, class Map>
static int launch_scatter(rhj_ctx *ctx, const Source &src, const Map &cm, uint32_t grid_chunks, uint32_t nbins,
                          const uint32_t *chunk_off, const uint64_t *bin_base, cudaStream_t st, uint32_t c_lo = 0,
                          uint32_t c_hi = 0xFFFFFFFFu) {
  if (grid_chunks == 0) return RHJ_OK;
  if (ctx->tune_scatter == 3)
    return launch_scatter_wc(ctx, src, cm, grid_chunks, nbins, chunk_off, bin_base, st, c_lo, c_hi);
  size_t smem = scatter_smem_bytes(nbins);
  RHJ_CUDA(cudaFuncSetAttribute(k_scatter<Source, Map>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  uint32_t grid = std::min<uint32_t>(grid_chunks, (uint32_t)ctx->sm_count * 2);
  k_scatter<Source, Map><<<grid, kScatTPB, smem, st>>>(src, cm, nbins, chunk_off, bin_base, ctx->d_status);
  RHJ_CHECK_LAUNCH("k_scatter");
  return RHJ_OK;
}

// histogram of chunks [c_lo, c_hi) of relation k (tasks/HistogramComputation.cpp:63-76 -> LocalHistogram)
template <int TB>
static int histogram_chunks(rhj_ctx *ctx, int k, uint32_t c_lo, uint32_t c_hi, cudaStream_t st) {
  if (c_hi <= c_lo) return RHJ_OK;
  const uint64_t first = (uint64_t)c_lo * ctx->chunk1;
  InputSource<TB> src{static_cast<const char *>(ctx->input[k]) + first * TB, ctx->P - 1, ctx->cfg.network_fanout_bits,
                      ctx->key_shift};
  ChunkMapFlat cm{std::min<uint64_t>(ctx->input_n[k] - first, (uint64_t)(c_hi - c_lo) * ctx->chunk1), ctx->chunk1, c_hi - c_lo};
  return launch_histogram(ctx, src, cm, c_hi - c_lo, ctx->P, ctx->chunk1_hist[k] + (uint64_t)c_lo * ctx->P, st);
}

// per-chunk counts -> exclusive per-chunk offsets + the relation's local histogram
static int histogram_scan(rhj_ctx *ctx, int k, cudaStream_t st) {
  dim3 grid((ctx->P + 31) / 32, 1);
  k_colscan<32><<<grid, 32 * 32, 0, st>>>(ctx->chunk1_hist[k], ctx->P, nullptr, ctx->nchunks1[k],
                                           ctx->d_local_hist + (uint64_t)k * ctx->P, ctx->d_status);
  RHJ_CHECK_LAUNCH("k_colscan");
  ctx->hist_done[k] = true;
  if (ctx->hist_done[k ^ 1]) ctx->stage = std::max(ctx->stage, 1);
  return RHJ_OK;
}

// A new join starts with its inner relation: per-join state, start event, result and status words.
static int begin_join(rhj_ctx *ctx, cudaStream_t st) {
  ctx->local_planned = false;
  ctx->epoch += 1;
  ctx->epoch_rel[0] += 1;
  ctx->ev_valid[EV_H2D0] = ctx->staged_pending;  // h2d_us only when this join's inputs were staged
  ctx->staged_pending = false;
  ctx->h2d_pipelined = false;
  ctx->stage = 0;
  for (int k = 0; k < 2; ++k) {
    ctx->rel_partitioned[k] = ctx->rel_scattered[k] = ctx->wait_timed[k] = false;
    ctx->hist_done[k] = ctx->offsets_done[k] = false;
    ctx->input[k] = nullptr;
    ctx->input_n[k] = 0;
    ctx->nchunks1[k] = 0;
  }
  int rc = record(ctx, EV_START, st);
  if (rc) return rc;
  RHJ_CUDA(cudaMemsetAsync(ctx->d_status, 0, sizeof(int), st));
  RHJ_CUDA(cudaMemsetAsync(ctx->d_result, 0, 5 * 8, st));
  return RHJ_OK;
}

static int set_input(rhj_ctx *ctx, int k, const void *dev, uint64_t n) {
  if (n > ctx->max_local[k])
    return fail(RHJ_ERR_INVALID, "relation %d: %llu tuples exceed the planned maximum %llu", k, (unsigned long long)n,
                (unsigned long long)ctx->max_local[k]);
  if (n && !dev) return fail(RHJ_ERR_INVALID, "NULL relation pointer");
  if (reinterpret_cast<uintptr_t>(dev) & (ctx->cfg.tuple_bytes - 1))
    return fail(RHJ_ERR_INVALID, "relations must be %u-byte aligned", ctx->cfg.tuple_bytes);
  ctx->input[k] = dev;
  ctx->input_n[k] = n;
  ctx->nchunks1[k] = (uint32_t)((n + ctx->chunk1 - 1) / ctx->chunk1);
  return RHJ_OK;
}

extern "C" int rhj_histogram_rel(rhj_ctx *ctx, int which, const void *rel_dev, uint64_t n, void *stream) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  if (which < 0 || which > 1) return fail(RHJ_ERR_INVALID, "which must be 0 (inner) or 1 (outer)");
  if (!ctx->planned) return fail(RHJ_ERR_STATE, "rhj_plan first");
  if (which == 1 && !ctx->hist_done[0]) return fail(RHJ_ERR_STATE, "the inner relation's histogram starts a join: rhj_histogram_rel(0) first");
  RHJ_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = RHJ_OK;
  if (which == 0) {
    rc = begin_join(ctx, st);
  } else {
    // A further piece of the outer relation against the inner relation that is already partitioned
    // (probe sides larger than the pool): everything outer-side starts over.
    ctx->hist_done[1] = ctx->offsets_done[1] = ctx->rel_scattered[1] = ctx->rel_partitioned[1] = false;
    ctx->wait_timed[1] = false;
    ctx->local_planned = false;
    ctx->stage = std::min(ctx->stage, 1);
    ctx->epoch_rel[1] += 1;
  }
  if (!rc) rc = set_input(ctx, which, rel_dev, n);
  if (!rc) rc = ctx->cfg.tuple_bytes == 16 ? histogram_chunks<16>(ctx, which, 0, ctx->nchunks1[which], st)
                                           : histogram_chunks<8>(ctx, which, 0, ctx->nchunks1[which], st);
  if (!rc) rc = histogram_scan(ctx, which, st);
  return rc;
}

extern "C" int rhj_histogram(rhj_ctx *ctx, const void *inner_dev, uint64_t inner_n, const void *outer_dev,
                             uint64_t outer_n, void *stream) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  if (!ctx->planned) return fail(RHJ_ERR_STATE, "rhj_plan first");
  if (inner_n > ctx->max_local[0] || outer_n > ctx->max_local[1])
    return fail(RHJ_ERR_INVALID, "relation sizes (%llu, %llu) exceed the planned maxima (%llu, %llu)",
                (unsigned long long)inner_n, (unsigned long long)outer_n, (unsigned long long)ctx->max_local[0],
                (unsigned long long)ctx->max_local[1]);
  if ((inner_n && !inner_dev) || (outer_n && !outer_dev)) return fail(RHJ_ERR_INVALID, "NULL relation pointer");
  if ((reinterpret_cast<uintptr_t>(inner_dev) | reinterpret_cast<uintptr_t>(outer_dev)) & (ctx->cfg.tuple_bytes - 1))
    return fail(RHJ_ERR_INVALID, "relations must be %u-byte aligned", ctx->cfg.tuple_bytes);
  int rc = rhj_histogram_rel(ctx, 0, inner_dev, inner_n, stream);
  if (!rc) rc = rhj_histogram_rel(ctx, 1, outer_dev, outer_n, stream);
  return rc;
}

extern "C" uint64_t *rhj_local_histogram_dev(rhj_ctx *ctx) { return ctx ? ctx->d_local_hist : nullptr; }

// k_begin..k_end: relations planned by this launch; gathered: [nranks][2][P] (both) or [nranks][P] (one)
static int plan_network(rhj_ctx *ctx, int k_begin, int k_end, const uint64_t *gathered_dev, bool one_relation_layout,
                        cudaStream_t st) {
  for (int r = 0; r < ctx->nranks; ++r)
    if (!ctx->peer_window[0][r] || !ctx->peer_window[1][r])
      return fail(RHJ_ERR_STATE, "window of rank %d not imported (rhj_window_import / rhj_window_set_peer)", r);
  PlanNetworkArgs a{};
  if (gathered_dev) {
    a.gathered = gathered_dev;
    a.rank_stride = one_relation_layout ? ctx->P : 2 * ctx->P;
    a.rel_stride = one_relation_layout ? 0 : ctx->P;
  } else {  // one rank: its own histograms, u64[2][P]
    a.gathered = ctx->d_local_hist;
    a.rank_stride = 2 * ctx->P;
    a.rel_stride = ctx->P;
  }
  a.k_begin = k_begin;
  a.k_end = k_end;
  a.nranks = ctx->nranks;
  a.rank = ctx->rank;
  a.parts = ctx->P;
  for (int k = 0; k < 2; ++k) {
    a.window_cap[k] = ctx->window_cap[k];
    a.max_chunks2[k] = ctx->cfg.enable_two_level ? ctx->max_chunks2[k] : 0xFFFFFFFFu;
    for (int r = 0; r < ctx->nranks; ++r) a.peer_window[k][r] = ctx->peer_window[k][r];
  }
  a.global_hist = ctx->d_global;
  a.base = ctx->d_base;
  a.rel = ctx->d_rel;
  a.abs = ctx->d_abs;
  a.dest = ctx->d_dest;
  a.assignment = ctx->d_assign;
  a.balanced = (ctx->cfg.flags & RHJ_FLAG_BALANCED_ASSIGNMENT) ? 1u : 0u;
  a.owned_list = ctx->d_owned_list;
  a.slot_of = ctx->d_slot_of;
  a.window_sizes = ctx->d_winsizes;
  a.nowned = ctx->nowned;
  a.part_off = ctx->d_part_off;
  a.part_cnt = ctx->d_part_cnt;
  a.cstart = ctx->d_cstart;
  a.chunk2 = ctx->chunk2;
  a.status = ctx->d_status;
  for (int k = 0; k < 2; ++k) a.stage_base[k] = ctx->staged_exchange ? reinterpret_cast<uint64_t>(ctx->sendbuf[k]) : 0;
  a.push_src = ctx->d_push_src;
  a.push_dst = ctx->d_push_dst;
  a.push_cnt = ctx->d_push_cnt;
  k_plan_network<<<1, 1024, ((size_t)ctx->P + 1) * 8, st>>>(a);
  RHJ_CHECK_LAUNCH("k_plan_network");
  for (int k = k_begin; k < k_end; ++k) ctx->offsets_done[k] = true;
  if (ctx->offsets_done[0] && ctx->offsets_done[1]) {
    int rc = record(ctx, EV_HIST, st);
    if (rc) return rc;
    ctx->stage = std::max(ctx->stage, 2);
  }
  return RHJ_OK;
}

extern "C" int rhj_compute_offsets(rhj_ctx *ctx, const uint64_t *gathered_dev, void *stream) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  if (!ctx->hist_done[0] || !ctx->hist_done[1]) return fail(RHJ_ERR_STATE, "rhj_histogram first");
  if (!gathered_dev && ctx->nranks > 1) return fail(RHJ_ERR_INVALID, "gathered histograms required when nranks > 1");
  RHJ_CUDA(cudaSetDevice(ctx->device));
  return plan_network(ctx, 0, 2, gathered_dev, false, static_cast<cudaStream_t>(stream));
}

extern "C" int rhj_compute_offsets_rel(rhj_ctx *ctx, int which, const uint64_t *gathered_rel_dev, void *stream) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  if (which < 0 || which > 1) return fail(RHJ_ERR_INVALID, "which must be 0 (inner) or 1 (outer)");
  if (!ctx->hist_done[which]) return fail(RHJ_ERR_STATE, "rhj_histogram_rel first");
  if (!gathered_rel_dev && ctx->nranks > 1) return fail(RHJ_ERR_INVALID, "gathered histograms required when nranks > 1");
  if (ctx->cfg.flags & RHJ_FLAG_BALANCED_ASSIGNMENT)
    return fail(RHJ_ERR_STATE, "the size-balanced assignment needs both relations' histograms: use rhj_compute_offsets");
  RHJ_CUDA(cudaSetDevice(ctx->device));
  return plan_network(ctx, which, which + 1, gathered_rel_dev, true, static_cast<cudaStream_t>(stream));
}

// One relation of the network pass.  Direct exchange (default): the scatter kernel's full-line
// stores go straight into the owner's window, local or peer (MPI_Put, data/Window.cpp:108-115).
extern "C" int rhj_network_partition_rel(rhj_ctx *ctx, int which, void *stream) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  if (which < 0 || which > 1) return fail(RHJ_ERR_INVALID, "which must be 0 (inner) or 1 (outer)");
  if (!ctx->planned || !ctx->offsets_done[which]) return fail(RHJ_ERR_STATE, "rhj_compute_offsets first");
  RHJ_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int k = which;
  ChunkMapFlat cm{ctx->input_n[k], ctx->chunk1, ctx->nchunks1[k]};
  int rc = RHJ_OK;
  const int saved_reserve = ctx->reserve_sms;
  // Staged exchange: the relation is scattered in a few groups of chunks, and every group's share of the
  // remote partitions' runs is pushed to the owners on the side stream while the next group is scattered
  // (runs are contiguous per partition and group, so the peers see long sequential writes).  From the
  // second group on -- and for the outer relation from the start, its pushes queue behind the inner
  // relation's -- the scatter leaves the push its SMs.
  const uint32_t nch = ctx->nchunks1[k];
  const uint32_t groups = (ctx->staged_exchange && ctx->tune_scatter == 3) ? std::max(1u, std::min(4u, nch / 256)) : 1u;
  for (uint32_t g = 0; g < groups && !rc; ++g) {
    const uint32_t c_lo = (uint32_t)((uint64_t)nch * g / groups), c_hi = (uint32_t)((uint64_t)nch * (g + 1) / groups);
    if (ctx->staged_exchange && (k == 1 || g > 0)) ctx->reserve_sms = std::max(saved_reserve, ctx->tune_push_sms);
    if (ctx->cfg.tuple_bytes == 16) {
      InputSource<16> src{ctx->input[k], ctx->P - 1, ctx->cfg.network_fanout_bits, ctx->key_shift};
      rc = launch_scatter(ctx, src, cm, nch, ctx->P, ctx->chunk1_hist[k], ctx->d_dest + (uint64_t)k * ctx->P, st, c_lo, groups > 1 ? c_hi : 0xFFFFFFFFu);
    } else {
      InputSource<8> src{ctx->input[k], ctx->P - 1, ctx->cfg.network_fanout_bits, ctx->key_shift};
      rc = launch_scatter(ctx, src, cm, nch, ctx->P, ctx->chunk1_hist[k], ctx->d_dest + (uint64_t)k * ctx->P, st, c_lo, groups > 1 ? c_hi : 0xFFFFFFFFu);
    }
    if (rc) break;
    if (ctx->staged_exchange) {
      RHJ_CUDA(cudaEventRecord(ctx->ev_scattered[k], st));
      RHJ_CUDA(cudaStreamWaitEvent(ctx->side, ctx->ev_scattered[k], 0));
      if (g == 0) RHJ_CUDA(cudaEventRecord(ctx->ev_push_begin[k], ctx->side));
      const uint32_t push_ctas = (uint32_t)std::max(8, ctx->tune_push_sms * 8);
      const uint32_t *row_lo = g == 0 ? nullptr : ctx->chunk1_hist[k] + (uint64_t)c_lo * ctx->P;
      const uint32_t *row_hi = g + 1 == groups ? nullptr : ctx->chunk1_hist[k] + (uint64_t)c_hi * ctx->P;
      k_push<<<push_ctas, kPushTPB, 0, ctx->side>>>(ctx->d_push_src + (uint64_t)k * ctx->P,
                                                    ctx->d_push_dst + (uint64_t)k * ctx->P,
                                                    ctx->d_push_cnt + (uint64_t)k * ctx->P, ctx->P, ctx->d_status, row_lo,
                                                    row_hi);
      RHJ_CHECK_LAUNCH("k_push");
    }
  }
  ctx->reserve_sms = saved_reserve;
  if (rc) return rc;
  if (ctx->use_flags) {
    // after this rank's last store of the relation: scatter kernel (direct) or last push (staged)
    cudaStream_t sig = ctx->staged_exchange ? ctx->side : st;
    FlagPeers fp{};
    for (int r = 0; r < ctx->nranks; ++r) fp.addr[r] = ctx->peer_flags[r];
    k_signal_peers<<<1, 64, 0, sig>>>(fp, (uint32_t)ctx->nranks, (uint32_t)ctx->rank, (uint32_t)k, ctx->epoch_rel[k]);
    RHJ_CHECK_LAUNCH("k_signal_peers");
  }
  if (ctx->staged_exchange) RHJ_CUDA(cudaEventRecord(ctx->ev_pushed[k], ctx->side));
  RHJ_CUDA(cudaEventRecord(ctx->ev_rel_net[k], st));
  ctx->rel_scattered[k] = true;
  if (ctx->rel_scattered[k ^ 1]) {
    rc = record(ctx, EV_NET, st);
    if (rc) return rc;
    ctx->stage = 3;
  }
  return RHJ_OK;
}

extern "C" int rhj_network_partition(rhj_ctx *ctx, void *stream) {
  int rc = rhj_network_partition_rel(ctx, 0, stream);
  if (!rc) rc = rhj_network_partition_rel(ctx, 1, stream);
  return rc;
}

// SMs the kernels launched from now on may occupy (0 = all): lets a launcher run two stages side
// by side on different streams, e.g. the outer relation's NVLink-bound network pass next to the
// inner relation's HBM-bound local pass.
extern "C" int rhj_set_sm_limit(rhj_ctx *ctx, int sms) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  if (sms < 0) return fail(RHJ_ERR_INVALID, "sms must be >= 0");
  ctx->reserve_sms = (sms == 0 || sms >= ctx->sm_count) ? 0 : ctx->sm_count - sms;
  return RHJ_OK;
}

static int launch_plan_local(rhj_ctx *ctx, int k_begin, int k_end, bool do_items, cudaStream_t st) {
  if (ctx->nowned == 0) {
    RHJ_CUDA(cudaMemsetAsync(ctx->d_item_jstart, 0, 8, st));
    ctx->local_planned = true;
    return RHJ_OK;
  }
  PlanLocalArgs a{};
  a.nowned = ctx->nowned;
  a.L = ctx->L;
  a.two_level = ctx->cfg.enable_two_level ? 1 : 0;
  a.part_off = ctx->d_part_off;
  a.part_cnt = ctx->d_part_cnt;
  a.sub_tot = ctx->d_sub_tot;
  for (int k = 0; k < 2; ++k) {
    a.out_base[k] = reinterpret_cast<uint64_t>(ctx->localout[k]);
    a.win_base[k] = reinterpret_cast<uint64_t>(ctx->window[k]);
  }
  a.sub_off = ctx->d_sub_off;
  a.sub_prefix = ctx->d_sub_prefix;
  a.sub_addr = ctx->d_sub_addr;
  a.sub_cnt = ctx->d_sub_cnt;
  a.item_lstart = ctx->d_item_lstart;
  a.item_jcount = ctx->d_item_jcount;
  a.cap_r = ctx->cap_r;
  a.sslice = ctx->sslice;
  a.k_begin = k_begin;
  a.k_end = k_end;
  a.do_items = do_items ? 1 : 0;
  k_plan_local<<<ctx->nowned, 256, 0, st>>>(a);
  RHJ_CHECK_LAUNCH("k_plan_local");
  if (do_items) {
    k_plan_items<<<1, 1024, 0, st>>>(ctx->d_item_jcount, ctx->nowned, ctx->d_item_jstart, ctx->d_work);
    RHJ_CHECK_LAUNCH("k_plan_items");
    ctx->local_planned = true;
  }
  return RHJ_OK;
}

// Make `stream` wait until this rank's pushes of relation `which` have been issued and completed.
extern "C" int rhj_exchange_wait(rhj_ctx *ctx, int which, void *stream) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  if (which < 0 || which > 1) return fail(RHJ_ERR_INVALID, "which must be 0 (inner) or 1 (outer)");
  if (!ctx->rel_scattered[which]) return fail(RHJ_ERR_STATE, "rhj_network_partition first");
  RHJ_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // SNETCOMPL (performance/Measurements.cpp:157-165): from here -- the stream has nothing else to do
  // until this relation has landed everywhere -- to the start of the local pass
  RHJ_CUDA(cudaEventRecord(ctx->ev_wait0[which], st));
  // this rank's own stores (`stream` may be another one than the network pass ran on)
  RHJ_CUDA(cudaStreamWaitEvent(st, ctx->staged_exchange ? ctx->ev_pushed[which] : ctx->ev_rel_net[which], 0));
  ctx->wait_timed[which] = true;
  if (ctx->use_flags) {
    // every peer's stores of this relation have landed once its flag shows this join's epoch
    k_wait_peers<<<1, 64, 0, st>>>(ctx->d_flags, (uint32_t)ctx->nranks, (uint32_t)ctx->rank, (uint32_t)which, ctx->epoch_rel[which],
                                   5000000000ull, ctx->d_status);
    RHJ_CHECK_LAUNCH("k_wait_peers");
  }
  return RHJ_OK;
}

// Host-blocking form: returns when this rank's pushes of relation `which` are complete.
extern "C" int rhj_exchange_sync(rhj_ctx *ctx, int which) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  if (which < 0 || which > 1) return fail(RHJ_ERR_INVALID, "which must be 0 (inner) or 1 (outer)");
  if (!ctx->rel_scattered[which]) return fail(RHJ_ERR_STATE, "rhj_network_partition first");
  RHJ_CUDA(cudaSetDevice(ctx->device));
  RHJ_CUDA(cudaEventSynchronize(ctx->staged_exchange ? ctx->ev_pushed[which] : ctx->ev_rel_net[which]));
  return RHJ_OK;
}

extern "C" int rhj_local_partition_rel(rhj_ctx *ctx, int which, void *stream) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  if (which < 0 || which > 1) return fail(RHJ_ERR_INVALID, "which must be 0 (inner) or 1 (outer)");
  if (!ctx->rel_scattered[which]) return fail(RHJ_ERR_STATE, "rhj_network_partition first");
  RHJ_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = RHJ_OK;
  if (!ctx->wait_timed[which]) {
    rc = rhj_exchange_wait(ctx, which, stream);  // this rank's own stores of the relation are done
    if (rc) return rc;
  }
  RHJ_CUDA(cudaEventRecord(ctx->ev_wait1[which], st));
  const int k = which;
  // the other relation's local pass may have run on another stream; its counts are read below
  if (ctx->rel_partitioned[k ^ 1]) RHJ_CUDA(cudaStreamWaitEvent(st, ctx->ev_rel_local[k ^ 1], 0));
  const bool other_done = ctx->rel_partitioned[k ^ 1];
  if (ctx->cfg.enable_two_level && ctx->nowned > 0) {
    const uint32_t L = ctx->L;
    // while the other relation is still being pushed, leave it some SMs
    const int saved_reserve = ctx->reserve_sms;
    if (ctx->staged_exchange && !other_done) ctx->reserve_sms = std::max(ctx->reserve_sms, ctx->tune_push_sms);
    ValueSource src{ctx->window[k], ctx->key_shift, L - 1};
    ChunkMapSeg cm{ctx->d_part_off + (uint64_t)k * ctx->nowned, ctx->d_part_cnt + (uint64_t)k * ctx->nowned,
                   ctx->d_cstart + (uint64_t)k * (ctx->nowned + 1), ctx->nowned, ctx->chunk2};
    rc = launch_histogram(ctx, src, cm, ctx->max_chunks2[k], L, ctx->chunk2_hist[k], st);
    if (rc) return rc;
    dim3 grid((L + 31) / 32, ctx->nowned);
    k_colscan<8><<<grid, 8 * 32, 0, st>>>(ctx->chunk2_hist[k], L, ctx->d_cstart + (uint64_t)k * (ctx->nowned + 1), 0,
                                           ctx->d_sub_tot + (uint64_t)k * ctx->NS, ctx->d_status);
    RHJ_CHECK_LAUNCH("k_colscan(local)");
    rc = launch_plan_local(ctx, k, k + 1, other_done, st);
    if (rc) return rc;
    rc = launch_scatter(ctx, src, cm, ctx->max_chunks2[k], L, ctx->chunk2_hist[k],
                        ctx->d_sub_addr + (uint64_t)k * ctx->NS, st);
    ctx->reserve_sms = saved_reserve;
    if (rc) return rc;
  }
  ctx->rel_partitioned[k] = true;
  RHJ_CUDA(cudaEventRecord(ctx->ev_rel_local[k], st));
  if (other_done) {
    rc = record(ctx, EV_LOCAL, st);
    if (rc) return rc;
    ctx->stage = 4;
  }
  return RHJ_OK;
}

extern "C" int rhj_local_partition(rhj_ctx *ctx, void *stream) {
  int rc = rhj_local_partition_rel(ctx, 0, stream);
  if (!rc) rc = rhj_local_partition_rel(ctx, 1, stream);
  return rc;
}

template <int CAP, bool EMIT>
static int launch_build_probe(rhj_ctx *ctx, const BuildProbeArgs &a, cudaStream_t st) {
  // the pipelined kernel relies on the 8-tuple alignment/padding of the local pass; single-pass
  // mode (arbitrary partition offsets) and wide rids use the generic kernel
  if (!ctx->tune_legacy_bp && ctx->cfg.enable_two_level) {
    k_expand_items<CAP><<<(a.NS + 255) / 256, 256, 0, st>>>(a, ctx->d_items);
    RHJ_CHECK_LAUNCH("k_expand_items");
    size_t smem3 = build_probe3_smem_bytes<CAP>();
    const bool ks32 = ctx->key_shift == 32;
    auto launch = [&](auto kern, int tpb) -> int {
      RHJ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem3));
      int per_sm = 0;
      RHJ_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, tpb, smem3));
      if (per_sm < 1) per_sm = 1;
      kern<<<ctx->sm_count * per_sm, tpb, smem3, st>>>(a);
      return RHJ_OK;
    };
    int rc;
    if (ctx->tune_bp_tpb == 512) rc = ks32 ? launch(k_build_probe3<CAP, 512, true, EMIT>, 512) : launch(k_build_probe3<CAP, 512, false, EMIT>, 512);
    else rc = ks32 ? launch(k_build_probe3<CAP, 1024, true, EMIT>, 1024) : launch(k_build_probe3<CAP, 1024, false, EMIT>, 1024);
    if (rc) return rc;
    RHJ_CHECK_LAUNCH("k_build_probe3");
    // items beyond the expanded list (extreme skew) are left to the generic kernel below
    BuildProbeArgs rest = a;
    rest.item_begin = ctx->items_cap;
    size_t smem_g = build_probe_smem_bytes<CAP>();
    RHJ_CUDA(cudaFuncSetAttribute(k_build_probe<CAP, EMIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_g));
    k_build_probe<CAP, EMIT><<<ctx->sm_count, kBpTPB, smem_g, st>>>(rest);
    RHJ_CHECK_LAUNCH("k_build_probe(overflow)");
    return RHJ_OK;
  }
  size_t smem = build_probe_smem_bytes<CAP>();
  RHJ_CUDA(cudaFuncSetAttribute(k_build_probe<CAP, EMIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  RHJ_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_build_probe<CAP, EMIT>, kBpTPB, smem));
  if (per_sm < 1) per_sm = 1;
  k_build_probe<CAP, EMIT><<<ctx->sm_count * per_sm, kBpTPB, smem, st>>>(a);
  RHJ_CHECK_LAUNCH("k_build_probe");
  return RHJ_OK;
}

template <bool EMIT>
static int launch_build_probe_cap(rhj_ctx *ctx, const BuildProbeArgs &a, cudaStream_t st) {
  switch (ctx->cap_r) {
    case 1024: return launch_build_probe<1024, EMIT>(ctx, a, st);
    case 2048: return launch_build_probe<2048, EMIT>(ctx, a, st);
    case 5120: return launch_build_probe<5120, EMIT>(ctx, a, st);
    case 8192: return launch_build_probe<8192, EMIT>(ctx, a, st);
    default: return launch_build_probe<4096, EMIT>(ctx, a, st);
  }
}

static int build_probe_impl(rhj_ctx *ctx, void *pairs_dev, uint64_t pairs_cap, bool emit, void *stream,
                            bool keep_result = false) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  if (!ctx->planned || ctx->stage < 4) return fail(RHJ_ERR_STATE, "rhj_local_partition first");
  if (emit && pairs_cap > 0 && !pairs_dev) return fail(RHJ_ERR_INVALID, "pairs_dev is NULL");
  if (emit && (reinterpret_cast<uintptr_t>(pairs_dev) & 15)) return fail(RHJ_ERR_INVALID, "pairs_dev must be 16-byte aligned");
  RHJ_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (!ctx->local_planned) {
    int rc = launch_plan_local(ctx, 0, 2, true, st);
    if (rc) return rc;
  }
  // the stage may be repeated on the same partitions (count first, then materialise)
  if (!keep_result) RHJ_CUDA(cudaMemsetAsync(ctx->d_result, 0, 5 * 8, st));
  RHJ_CUDA(cudaMemsetAsync(ctx->d_work, 0, 2 * 8, st));
  if (ctx->nowned > 0) {
    BuildProbeArgs a{};
    for (int k = 0; k < 2; ++k) a.src[k] = ctx->cfg.enable_two_level ? ctx->localout[k] : ctx->window[k];
    a.sub_off = ctx->d_sub_off;
    a.sub_cnt = ctx->d_sub_cnt;
    a.item_lstart = ctx->d_item_lstart;
    a.item_jstart = ctx->d_item_jstart;
    a.work_counter = ctx->d_work;
    a.result = ctx->d_result;
    a.nowned = ctx->nowned;
    a.L = ctx->L;
    a.NS = ctx->NS;
    a.owned = ctx->d_owned_list + first_owned_slot(ctx->P, ctx->rank, ctx->nranks);
    a.net_bits = ctx->cfg.network_fanout_bits;
    a.key_shift = ctx->key_shift;
    a.shift_bits = ctx->shift_bits;
    a.sslice = ctx->sslice;
    a.status = ctx->d_status;
    a.items = ctx->d_items;
    a.items_cap = ctx->items_cap;
    a.pairs = static_cast<ulonglong2 *>(pairs_dev);
    a.pairs_cap = pairs_cap;
    a.pair_cursor = ctx->d_work + 1;
    int rc = emit ? launch_build_probe_cap<true>(ctx, a, st) : launch_build_probe_cap<false>(ctx, a, st);
    if (rc) return rc;
  }
  int rc = record(ctx, EV_BP, st);
  if (rc) return rc;
  ctx->stage = 5;
  return RHJ_OK;
}

extern "C" int rhj_build_probe(rhj_ctx *ctx, void *stream) { return build_probe_impl(ctx, nullptr, 0, false, stream); }

extern "C" int rhj_build_probe_accumulate(rhj_ctx *ctx, void *stream) {
  return build_probe_impl(ctx, nullptr, 0, false, stream, true);
}

extern "C" int rhj_build_probe_pairs(rhj_ctx *ctx, rhj_pair *pairs_dev, uint64_t capacity, void *stream) {
  return build_probe_impl(ctx, pairs_dev, capacity, true, stream);
}

// -----------------------------------------------------------------------------------
// getters
// -----------------------------------------------------------------------------------
static int copy_out(rhj_ctx *ctx, void *dst, const void *src, size_t bytes, void *stream) {
  RHJ_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  RHJ_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, st));
  RHJ_CUDA(cudaStreamSynchronize(st));
  return RHJ_OK;
}

extern "C" int rhj_get_assignment(rhj_ctx *ctx, uint32_t *out_host, void *stream) {
  if (!ctx || !out_host) return fail(RHJ_ERR_INVALID, "NULL argument");
  if (ctx->stage < 2) return fail(RHJ_ERR_STATE, "rhj_compute_offsets first");
  return copy_out(ctx, out_host, ctx->d_assign, (size_t)ctx->P * 4, stream);
}

static int get_u64(rhj_ctx *ctx, int which, const uint64_t *arr, int min_stage, uint64_t *out_host, void *stream) {
  if (!ctx || !out_host) return fail(RHJ_ERR_INVALID, "NULL argument");
  if (which < 0 || which > 1) return fail(RHJ_ERR_INVALID, "which must be 0 (inner) or 1 (outer)");
  if (ctx->stage < min_stage) return fail(RHJ_ERR_STATE, "stage not reached");
  return copy_out(ctx, out_host, arr + (uint64_t)which * ctx->P, (size_t)ctx->P * 8, stream);
}

extern "C" int rhj_get_local_histogram(rhj_ctx *ctx, int which, uint64_t *out_host, void *stream) {
  return get_u64(ctx, which, ctx ? ctx->d_local_hist : nullptr, 1, out_host, stream);
}
extern "C" int rhj_get_global_histogram(rhj_ctx *ctx, int which, uint64_t *out_host, void *stream) {
  return get_u64(ctx, which, ctx ? ctx->d_global : nullptr, 2, out_host, stream);
}
extern "C" int rhj_get_base_offsets(rhj_ctx *ctx, int which, uint64_t *out_host, void *stream) {
  return get_u64(ctx, which, ctx ? ctx->d_base : nullptr, 2, out_host, stream);
}
extern "C" int rhj_get_write_offsets(rhj_ctx *ctx, int which, uint64_t *out_host, void *stream) {
  return get_u64(ctx, which, ctx ? ctx->d_abs : nullptr, 2, out_host, stream);
}

// slot j of partition p among this rank's partitions, or an error if another rank owns it
static int owned_slot(rhj_ctx *ctx, uint32_t p, uint32_t *j) {
  if (p >= ctx->P) return fail(RHJ_ERR_INVALID, "partition %u outside [0,%u)", p, ctx->P);
  uint32_t owner = 0;
  if (!(ctx->cfg.flags & RHJ_FLAG_BALANCED_ASSIGNMENT) || ctx->nranks == 1) {
    owner = p % (uint32_t)ctx->nranks;  // AssignmentMap.cpp:41-43: known without asking the device
    *j = p / (uint32_t)ctx->nranks;
  } else {
    int rc = copy_out(ctx, &owner, ctx->d_assign + p, 4, nullptr);
    if (!rc) rc = copy_out(ctx, j, ctx->d_slot_of + p, 4, nullptr);
    if (rc) return rc;
  }
  if (owner != (uint32_t)ctx->rank) return fail(RHJ_ERR_INVALID, "partition %u not owned by rank %d", p, ctx->rank);
  return RHJ_OK;
}

extern "C" int rhj_get_local_partition(rhj_ctx *ctx, int which, uint32_t p, uint64_t *hist_host, uint64_t *prefix_host,
                                       void *stream) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  if (which < 0 || which > 1) return fail(RHJ_ERR_INVALID, "which must be 0 or 1");
  if (!ctx->cfg.enable_two_level) return fail(RHJ_ERR_STATE, "two-level partitioning is off");
  if (ctx->stage < 4) return fail(RHJ_ERR_STATE, "rhj_local_partition first");
  uint32_t j = 0;
  int rc = owned_slot(ctx, p, &j);
  if (rc) return rc;
  std::vector<uint32_t> cnt(ctx->L);
  rc = copy_out(ctx, cnt.data(), ctx->d_sub_cnt + (uint64_t)which * ctx->NS + (uint64_t)j * ctx->L, (size_t)ctx->L * 4, stream);
  if (rc) return rc;
  if (hist_host)
    for (uint32_t l = 0; l < ctx->L; ++l) hist_host[l] = cnt[l];
  if (prefix_host)
    rc = copy_out(ctx, prefix_host, ctx->d_sub_prefix + (uint64_t)which * ctx->NS + (uint64_t)j * ctx->L, (size_t)ctx->L * 8, stream);
  return rc;
}

extern "C" int rhj_get_partition_dev(rhj_ctx *ctx, int which, uint32_t p, int after_local_pass, void **out_dev,
                                     uint64_t *out_tuples) {
  if (!ctx || !out_dev || !out_tuples) return fail(RHJ_ERR_INVALID, "NULL argument");
  if (which < 0 || which > 1) return fail(RHJ_ERR_INVALID, "which must be 0 or 1");
  if (ctx->stage < (after_local_pass ? 4 : 3)) return fail(RHJ_ERR_STATE, "stage not reached");
  if (after_local_pass && !ctx->cfg.enable_two_level) return fail(RHJ_ERR_STATE, "two-level partitioning is off");
  uint32_t j = 0;
  int rc = owned_slot(ctx, p, &j);
  if (rc) return rc;
  uint64_t off = 0, cnt = 0;
  rc = copy_out(ctx, &off, ctx->d_part_off + (uint64_t)which * ctx->nowned + j, 8, nullptr);
  if (!rc) rc = copy_out(ctx, &cnt, ctx->d_part_cnt + (uint64_t)which * ctx->nowned + j, 8, nullptr);
  if (rc) return rc;
  if (after_local_pass) {
    *out_dev = ctx->localout[which] + ((off + (uint64_t)j * (ctx->L * 8 + 8) + 7) & ~7ull);
    *out_tuples = cnt + (uint64_t)ctx->L * 8;
  } else {
    *out_dev = ctx->window[which] + off;
    *out_tuples = cnt;
  }
  return RHJ_OK;
}

extern "C" int rhj_get_result(rhj_ctx *ctx, rhj_result *out, void *stream) {
  if (!ctx || !out) return fail(RHJ_ERR_INVALID, "NULL argument");
  if (ctx->stage < 5) return fail(RHJ_ERR_STATE, "rhj_build_probe first");
  unsigned long long r[5];
  int status = 0;
  RHJ_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  RHJ_CUDA(cudaMemcpyAsync(r, ctx->d_result, sizeof(r), cudaMemcpyDeviceToHost, st));
  RHJ_CUDA(cudaMemcpyAsync(&status, ctx->d_status, sizeof(int), cudaMemcpyDeviceToHost, st));
  RHJ_CUDA(cudaStreamSynchronize(st));
  if (status == kDevWindowOverflow)
    return fail(RHJ_ERR_CAPACITY, "a receive window is smaller than the tuples assigned to it; raise allocation_factor or the window capacity");
  if (status == kDevPeerTimeout)
    return fail(RHJ_ERR_STATE, "a peer rank did not finish its network pass within 5 s (arrival flag never set)");
  if (status != kDevOk) return fail(RHJ_ERR_CAPACITY, "device workspace bound exceeded (status %d)", status);
  out->matches = r[0];
  out->sum_key = r[1];
  out->sum_rid_r = r[2];
  out->sum_rid_s = r[3];
  out->sum_mix = r[4];
  return RHJ_OK;
}

extern "C" int rhj_get_timings(rhj_ctx *ctx, rhj_timings *out) {
  if (!ctx || !out) return fail(RHJ_ERR_INVALID, "NULL argument");
  memset(out, 0, sizeof(*out));
  if (ctx->stage < 5) return fail(RHJ_ERR_STATE, "no finished join");
  auto us = [&](int a, int b) -> double {
    float ms = 0;
    if (!ctx->ev_valid[a] || !ctx->ev_valid[b]) return 0.0;
    if (cudaEventElapsedTime(&ms, ctx->ev[a], ctx->ev[b]) != cudaSuccess) { cudaGetLastError(); return 0.0; }
    return (double)ms * 1000.0;
  };
  RHJ_CUDA(cudaSetDevice(ctx->device));
  RHJ_CUDA(cudaEventSynchronize(ctx->ev[EV_BP]));
  out->jtotal_us = us(EV_START, EV_BP);
  out->jhist_us = us(EV_START, EV_HIST);
  out->jmpi_us = us(EV_HIST, EV_NET);
  out->jproc_us = us(EV_NET, EV_BP);
  out->lp_us = us(EV_NET, EV_LOCAL);
  out->bp_us = us(EV_LOCAL, EV_BP);
  out->h2d_us = ctx->ev_valid[EV_H2D0] ? us(EV_H2D0, EV_START) : 0.0;
  if (ctx->h2d_pipelined) {  // copies overlap the join: from the first byte's start to the last byte's arrival
    float ms = 0;
    if (cudaEventElapsedTime(&ms, ctx->ev[EV_H2D0], ctx->ev_h2d_done) == cudaSuccess) out->h2d_us = (double)ms * 1000.0;
    else cudaGetLastError();
  }
  if (ctx->staged_exchange && ctx->rel_scattered[0] && ctx->rel_scattered[1]) {
    // NVLink window of the staged exchange: first push of the inner relation to last push of the outer one
    float ms = 0;
    if (cudaEventElapsedTime(&ms, ctx->ev_push_begin[0], ctx->ev_pushed[1]) == cudaSuccess) out->push_us = (double)ms * 1000.0;
    else cudaGetLastError();
  }
  for (int k = 0; k < 2; ++k) {
    if (!ctx->wait_timed[k]) continue;
    float ms = 0;
    if (cudaEventElapsedTime(&ms, ctx->ev_wait0[k], ctx->ev_wait1[k]) == cudaSuccess) out->snetcompl_us += (double)ms * 1000.0;
    else cudaGetLastError();
  }
  return RHJ_OK;
}

// -----------------------------------------------------------------------------------
// the operator (operators/HashJoin.cpp:45-220), single rank
// -----------------------------------------------------------------------------------
extern "C" int rhj_join(rhj_ctx *ctx, const void *inner_dev, uint64_t inner_n, const void *outer_dev, uint64_t outer_n,
                        void *stream, rhj_result *out) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  if (ctx->nranks != 1)
    return fail(RHJ_ERR_STATE, "rhj_join drives a single rank; with %d ranks call the stages and all-gather "
                "rhj_local_histogram_dev() between rhj_histogram and rhj_compute_offsets", ctx->nranks);
  if (!ctx->planned) {
    int rc = rhj_plan(ctx, inner_n, outer_n, 0, 0);
    if (rc) return rc;
  }
  int rc = rhj_histogram(ctx, inner_dev, inner_n, outer_dev, outer_n, stream);
  if (!rc) rc = rhj_compute_offsets(ctx, nullptr, stream);
  if (!rc) rc = rhj_network_partition(ctx, stream);
  if (!rc) rc = rhj_local_partition(ctx, stream);
  if (!rc) rc = rhj_build_probe(ctx, stream);
  if (!rc && out) rc = rhj_get_result(ctx, out, stream);
  return rc;
}

// Host relations -> the staging buffers rhj_plan reserved (RHJ_FLAG_STAGE_HOST_INPUTS), for any number of
// ranks: the multi-rank form of what rhj_join_host does before it joins.
extern "C" int rhj_stage_host_inputs(rhj_ctx *ctx, const void *inner_host, uint64_t inner_n, const void *outer_host,
                                     uint64_t outer_n, void *stream, void **inner_dev, void **outer_dev) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  if (!(ctx->cfg.flags & RHJ_FLAG_STAGE_HOST_INPUTS))
    return fail(RHJ_ERR_STATE, "rhj_stage_host_inputs needs RHJ_FLAG_STAGE_HOST_INPUTS in rhj_config.flags (device staging buffers)");
  if (!ctx->planned) return fail(RHJ_ERR_STATE, "rhj_plan first");
  if (inner_n > ctx->max_local[0] || outer_n > ctx->max_local[1])
    return fail(RHJ_ERR_INVALID, "relation sizes exceed the planned maxima");
  if ((inner_n && !inner_host) || (outer_n && !outer_host)) return fail(RHJ_ERR_INVALID, "NULL relation pointer");
  RHJ_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  RHJ_CUDA(cudaEventRecord(ctx->ev[EV_H2D0], st));
  ctx->staged_pending = true;
  RHJ_CUDA(cudaMemcpyAsync(ctx->staged_input[0], inner_host, inner_n * ctx->cfg.tuple_bytes, cudaMemcpyHostToDevice, st));
  RHJ_CUDA(cudaMemcpyAsync(ctx->staged_input[1], outer_host, outer_n * ctx->cfg.tuple_bytes, cudaMemcpyHostToDevice, st));
  if (inner_dev) *inner_dev = ctx->staged_input[0];
  if (outer_dev) *outer_dev = ctx->staged_input[1];
  return RHJ_OK;
}

// HashJoin::join() from host buffers, single rank, as a pipeline (first slice): the
// relations cross PCIe in pieces on a copy stream, the histogram of a piece runs as soon as it has
// arrived, and the inner relation is partitioned (network + local pass) while the outer relation is
// still being copied.  What is left after the last byte has arrived is the outer relation's last
// histogram piece, its two scatter passes and build/probe.
static constexpr int kMaxPieces = 64;  // per relation

extern "C" int rhj_join_host(rhj_ctx *ctx, const void *inner_host, uint64_t inner_n, const void *outer_host,
                             uint64_t outer_n, void *stream, rhj_result *out) {
  if (!ctx) return fail(RHJ_ERR_INVALID, "ctx is NULL");
  if (!(ctx->cfg.flags & RHJ_FLAG_STAGE_HOST_INPUTS))
    return fail(RHJ_ERR_STATE, "rhj_join_host needs RHJ_FLAG_STAGE_HOST_INPUTS in rhj_config.flags (device staging buffers)");
  if (ctx->nranks != 1)
    return fail(RHJ_ERR_STATE, "rhj_join_host drives a single rank; with %d ranks call rhj_stage_host_inputs and the stages", ctx->nranks);
  if (!ctx->planned) {
    int rc = rhj_plan(ctx, inner_n, outer_n, 0, 0);
    if (rc) return rc;
  }
  if (inner_n > ctx->max_local[0] || outer_n > ctx->max_local[1])
    return fail(RHJ_ERR_INVALID, "relation sizes exceed the planned maxima");
  if ((inner_n && !inner_host) || (outer_n && !outer_host)) return fail(RHJ_ERR_INVALID, "NULL relation pointer");
  RHJ_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (!ctx->copy_stream) {
    RHJ_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2 * kMaxPieces; ++i) RHJ_CUDA(cudaEventCreateWithFlags(&ctx->ev_piece[i], cudaEventDisableTiming));
    RHJ_CUDA(cudaEventCreate(&ctx->ev_h2d_done));
  }
  const void *host[2] = {inner_host, outer_host};
  const uint64_t n[2] = {inner_n, outer_n};
  const uint32_t tb = ctx->cfg.tuple_bytes;
  // the copies start once everything queued on `stream` so far is done (an earlier join may still read the buffers)
  RHJ_CUDA(cudaEventRecord(ctx->ev[EV_H2D0], st));
  RHJ_CUDA(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev[EV_H2D0], 0));
  uint32_t piece_chunks[2], npieces[2];
  for (int k = 0; k < 2; ++k) {
    const uint32_t nch = (uint32_t)((n[k] + ctx->chunk1 - 1) / ctx->chunk1);
    // pieces of about 256 MB, at most kMaxPieces of them
    const uint32_t want = std::max<uint32_t>(1, (uint32_t)(ctx->tune_piece_bytes / ((uint64_t)ctx->chunk1 * tb)));
    piece_chunks[k] = std::max(want, (nch + kMaxPieces - 1) / kMaxPieces);
    npieces[k] = (nch + piece_chunks[k] - 1) / piece_chunks[k];
    for (uint32_t j = 0; j < npieces[k]; ++j) {
      const uint64_t b = (uint64_t)j * piece_chunks[k] * ctx->chunk1;
      const uint64_t e = std::min<uint64_t>(n[k], b + (uint64_t)piece_chunks[k] * ctx->chunk1);
      RHJ_CUDA(cudaMemcpyAsync(static_cast<char *>(ctx->staged_input[k]) + b * tb, static_cast<const char *>(host[k]) + b * tb,
                               (e - b) * tb, cudaMemcpyHostToDevice, ctx->copy_stream));
      RHJ_CUDA(cudaEventRecord(ctx->ev_piece[k * kMaxPieces + j], ctx->copy_stream));
    }
  }
  RHJ_CUDA(cudaEventRecord(ctx->ev_h2d_done, ctx->copy_stream));
  ctx->staged_pending = true;
  int rc = begin_join(ctx, st);
  ctx->h2d_pipelined = true;
  for (int k = 0; k < 2 && !rc; ++k) {
    rc = set_input(ctx, k, ctx->staged_input[k], n[k]);
    for (uint32_t j = 0; j < npieces[k] && !rc; ++j) {
      RHJ_CUDA(cudaStreamWaitEvent(st, ctx->ev_piece[k * kMaxPieces + j], 0));
      const uint32_t c_lo = j * piece_chunks[k], c_hi = std::min(ctx->nchunks1[k], c_lo + piece_chunks[k]);
      rc = tb == 16 ? histogram_chunks<16>(ctx, k, c_lo, c_hi, st) : histogram_chunks<8>(ctx, k, c_lo, c_hi, st);
    }
    if (!rc) rc = histogram_scan(ctx, k, st);
    if (!rc) rc = plan_network(ctx, k, k + 1, nullptr, false, st);
    if (!rc) rc = rhj_network_partition_rel(ctx, k, stream);
    if (!rc) rc = rhj_local_partition_rel(ctx, k, stream);
  }
  if (!rc) rc = rhj_build_probe(ctx, stream);
  if (!rc && out) rc = rhj_get_result(ctx, out, stream);
  return rc;
}

// -----------------------------------------------------------------------------------
// device generator
// -----------------------------------------------------------------------------------
static PermParams make_perm(uint64_t domain, uint64_t seed) {
  PermParams pp{};
  pp.domain = domain;
  uint32_t b = 0;
  while (b < 64 && (1ull << b) < domain) ++b;
  pp.mask = b >= 64 ? ~0ull : ((1ull << b) - 1);
  pp.shift = b / 2 ? b / 2 : 1;
  pp.m0 = splitmix64(seed) | 1; pp.a0 = splitmix64(seed + 1);
  pp.m1 = splitmix64(seed + 2) | 1; pp.a1 = splitmix64(seed + 3);
  pp.m2 = splitmix64(seed + 4) | 1; pp.a2 = splitmix64(seed + 5);
  return pp;
}

extern "C" int rhj_apply_perm(uint64_t *vals_dev, uint64_t n, uint32_t stride_words, uint64_t domain, uint64_t seed,
                              void *stream) {
  if (n == 0) return RHJ_OK;
  if (!vals_dev || stride_words == 0) return fail(RHJ_ERR_INVALID, "NULL values or zero stride");
  uint32_t grid = (uint32_t)std::min<uint64_t>((n + 255) / 256, 148 * 16);
  k_apply_perm<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(vals_dev, n, stride_words, make_perm(domain, seed));
  RHJ_CHECK_LAUNCH("k_apply_perm");
  return RHJ_OK;
}

extern "C" int rhj_fill_perm(void *tuples_dev, uint32_t tuple_bytes, uint64_t n, uint64_t first, uint64_t domain,
                             uint64_t seed, uint64_t modulo, uint64_t start_key, uint64_t start_rid, void *stream) {
  if (tuple_bytes != 16 && tuple_bytes != 8) return fail(RHJ_ERR_INVALID, "tuple_bytes must be 16 or 8");
  if (n == 0) return RHJ_OK;
  if (!tuples_dev) return fail(RHJ_ERR_INVALID, "NULL tuples");
  const PermParams pp = make_perm(domain, seed);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  uint32_t grid = (uint32_t)std::min<uint64_t>((n + 255) / 256, 148 * 16);
  if (tuple_bytes == 16) k_fill_perm<16><<<grid, 256, 0, st>>>(tuples_dev, n, first, pp, modulo, start_key, start_rid);
  else k_fill_perm<8><<<grid, 256, 0, st>>>(tuples_dev, n, first, pp, modulo, start_key, start_rid);
  RHJ_CHECK_LAUNCH("k_fill_perm");
  return RHJ_OK;
}

// -----------------------------------------------------------------------------------
// runtime conveniences
// -----------------------------------------------------------------------------------
extern "C" int rhj_host_alloc(void **out_host, uint64_t bytes) {
  if (!out_host) return fail(RHJ_ERR_INVALID, "NULL argument");
  RHJ_CUDA(cudaHostAlloc(out_host, bytes ? bytes : 1, cudaHostAllocDefault));
  return RHJ_OK;
}
extern "C" int rhj_host_free(void *host) {
  if (host) RHJ_CUDA(cudaFreeHost(host));
  return RHJ_OK;
}
extern "C" int rhj_memcpy_d2h(void *dst_host, const void *src_dev, uint64_t bytes) {
  RHJ_CUDA(cudaMemcpy(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost));
  return RHJ_OK;
}
extern "C" int rhj_memcpy_h2d(void *dst_dev, const void *src_host, uint64_t bytes) {
  RHJ_CUDA(cudaMemcpy(dst_dev, src_host, bytes, cudaMemcpyHostToDevice));
  return RHJ_OK;
}

// -----------------------------------------------------------------------------------
// host-side planning logic (no GPU needed)
// -----------------------------------------------------------------------------------
extern "C" void rhj_host_assignment(uint32_t nranks, uint32_t parts, uint32_t *assignment) {
  for (uint32_t p = 0; p < parts; ++p) assignment[p] = p % nranks;  // histograms/AssignmentMap.cpp:41-43
}

// The device version is in k_plan_network; both follow the same rule, stated in include/rhj.h.
extern "C" void rhj_host_assignment_balanced(const uint64_t *global_inner, const uint64_t *global_outer, uint32_t nranks,
                                             uint32_t parts, uint32_t *assignment) {
  std::vector<std::pair<uint64_t, uint32_t>> order(parts);
  for (uint32_t p = 0; p < parts; ++p) order[p] = {global_inner[p] + global_outer[p], p};
  std::sort(order.begin(), order.end(), [](const std::pair<uint64_t, uint32_t> &a, const std::pair<uint64_t, uint32_t> &b) {
    return a.first != b.first ? a.first > b.first : a.second < b.second;
  });
  std::vector<uint64_t> load(nranks, 0);
  std::vector<uint32_t> room(nranks);
  for (uint32_t r = 0; r < nranks; ++r) room[r] = parts / nranks + (r < parts % nranks ? 1u : 0u);
  for (uint32_t i = 0; i < parts; ++i) {
    uint32_t best = nranks;
    for (uint32_t r = 0; r < nranks; ++r)
      if (room[r] && (best == nranks || load[r] < load[best])) best = r;
    assignment[order[i].second] = best;
    load[best] += order[i].first;
    room[best] -= 1;
  }
}

extern "C" void rhj_host_offsets_assigned(const uint64_t *locals, const uint32_t *assignment, uint32_t nranks, uint32_t rank,
                                          uint32_t parts, uint64_t *global, uint64_t *base, uint64_t *rel, uint64_t *abs,
                                          uint64_t *window_sizes) {
  std::vector<uint64_t> running(nranks, 0);
  for (uint32_t p = 0; p < parts; ++p) {
    uint64_t g = 0, before = 0;
    for (uint32_t r = 0; r < nranks; ++r) {
      uint64_t v = locals[(uint64_t)r * parts + p];
      if (r < rank) before += v;  // histograms/OffsetMap.cpp:78-83 (exclusive scan over ranks)
      g += v;                     // histograms/GlobalHistogram.cpp:37-42
    }
    uint32_t owner = assignment[p];
    if (global) global[p] = g;
    if (base) base[p] = running[owner];  // histograms/OffsetMap.cpp:59-73
    if (rel) rel[p] = before;
    if (abs) abs[p] = running[owner] + before;  // histograms/OffsetMap.cpp:87-93
    running[owner] += g;
  }
  if (window_sizes)
    for (uint32_t r = 0; r < nranks; ++r) window_sizes[r] = running[r];  // data/Window.cpp:168-178
}

extern "C" void rhj_host_offsets(const uint64_t *locals, uint32_t nranks, uint32_t rank, uint32_t parts, uint64_t *global,
                                 uint64_t *base, uint64_t *rel, uint64_t *abs, uint64_t *window_sizes) {
  std::vector<uint32_t> assignment(parts);
  rhj_host_assignment(nranks, parts, assignment.data());
  rhj_host_offsets_assigned(locals, assignment.data(), nranks, rank, parts, global, base, rel, abs, window_sizes);
}

// -----------------------------------------------------------------------------------
// Local histogram for callers that must not allocate (see include/rhj.h).
//
// The per-chunk count matrix that rhj_histogram() keeps in the pool lives here in a
// fixed array inside the library image instead, so this entry point never asks the
// CUDA runtime for memory.  Every GPU that loads the library carries the array
// (4 MiB + 256 B) until the process exits.  All calls on a device share it and it
// keeps whatever the previous call left behind: the status word is cleared on the
// caller's stream at the start of each call, the counts are overwritten before they
// are read.  Two calls whose GPU work overlaps (different streams, or different host
// threads) race on it.
// -----------------------------------------------------------------------------------
#include <mutex>

#define RHJ_NOALLOC_MAX_CELLS (1u << 20) /* chunks x bins */
#define RHJ_NOALLOC_HEADER_BYTES 256u
static __device__ __align__(256) unsigned char g_noalloc_scratch[RHJ_NOALLOC_HEADER_BYTES + RHJ_NOALLOC_MAX_CELLS * 4u];

static constexpr int kNoallocMaxDevices = 64;
static std::mutex g_noalloc_mutex;
static unsigned char *g_noalloc_dev_ptr[kNoallocMaxDevices] = {};
static int g_noalloc_sm_count[kNoallocMaxDevices] = {};

// In host code `g_noalloc_scratch` is only a handle for the runtime, not an address a
// kernel can use; the device address is looked up once per device, with that device current.
static int noalloc_scratch(unsigned char **out, int *sm_count) {
  int dev = -1;
  RHJ_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoallocMaxDevices) return fail(RHJ_ERR_INVALID, "device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_dev_ptr[dev]) {
    void *p = nullptr;
    RHJ_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    int sms = 0;
    RHJ_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    g_noalloc_sm_count[dev] = sms;
    g_noalloc_dev_ptr[dev] = static_cast<unsigned char *>(p);
  }
  *out = g_noalloc_dev_ptr[dev];
  *sm_count = g_noalloc_sm_count[dev];
  return RHJ_OK;
}

template <int TB>
static int noalloc_histogram_launch(const void *tuples_dev, uint64_t n, uint32_t net_bits, uint32_t nchunks,
                                    uint32_t *matrix, int *status, uint64_t *hist_out_dev, int sm_count,
                                    cudaStream_t st) {
  const uint32_t P = 1u << net_bits;
  if (nchunks > 0) {
    InputSource<TB> src{tuples_dev, P - 1, net_bits, 0};
    ChunkMapFlat cm{n, kDefaultChunk, nchunks};
    const uint32_t nrep = std::max<uint32_t>(1, std::min<uint32_t>(kHistTPB / 32, 4096 / P));
    const uint32_t grid = std::min<uint32_t>(nchunks, (uint32_t)sm_count * 4);
    k_histogram<InputSource<TB>, ChunkMapFlat, false>
        <<<grid, kHistTPB, (size_t)P * nrep * 4, st>>>(src, cm, P, nrep, matrix, status);
    RHJ_CHECK_LAUNCH("k_histogram(noalloc)");
  }
  dim3 grid((P + 31) / 32, 1);
  k_colscan<32><<<grid, 32 * 32, 0, st>>>(matrix, P, nullptr, nchunks, hist_out_dev, status);
  RHJ_CHECK_LAUNCH("k_colscan(noalloc)");
  return RHJ_OK;
}

extern "C" int rhj_local_histogram_noalloc(const void *tuples_dev, uint32_t tuple_bytes, uint64_t n,
                                           uint32_t network_fanout_bits, uint64_t *hist_out_dev, void *stream) {
  if (tuple_bytes != 16 && tuple_bytes != 8) return fail(RHJ_ERR_INVALID, "tuple_bytes must be 16 or 8");
  if (network_fanout_bits < 1 || network_fanout_bits > 12)
    return fail(RHJ_ERR_INVALID, "network_fanout_bits %u outside [1,12]", network_fanout_bits);
  if (!hist_out_dev) return fail(RHJ_ERR_INVALID, "hist_out_dev is NULL");
  if (n && !tuples_dev) return fail(RHJ_ERR_INVALID, "tuples_dev is NULL");
  if (reinterpret_cast<uintptr_t>(tuples_dev) & (tuple_bytes - 1))
    return fail(RHJ_ERR_INVALID, "tuples must be %u-byte aligned", tuple_bytes);
  const uint64_t nchunks = (n + kDefaultChunk - 1) / kDefaultChunk;
  const uint64_t cells = nchunks << network_fanout_bits;
  if (cells > RHJ_NOALLOC_MAX_CELLS)
    return fail(RHJ_ERR_CAPACITY,
                "rhj_local_histogram_noalloc: %llu tuples at %u radix bits need %llu scratch cells, the built-in "
                "array holds %u (at most %llu tuples at this fan-out); use rhj_histogram for larger inputs",
                (unsigned long long)n, network_fanout_bits, (unsigned long long)cells, RHJ_NOALLOC_MAX_CELLS,
                (unsigned long long)(RHJ_NOALLOC_MAX_CELLS >> network_fanout_bits) * kDefaultChunk);
  unsigned char *scratch = nullptr;
  int sm_count = 0;
  int rc = noalloc_scratch(&scratch, &sm_count);
  if (rc) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int *status = reinterpret_cast<int *>(scratch);
  uint32_t *matrix = reinterpret_cast<uint32_t *>(scratch + RHJ_NOALLOC_HEADER_BYTES);
  RHJ_CUDA(cudaMemsetAsync(status, 0, sizeof(int), st));  // the only word that must start at zero
  return tuple_bytes == 16
             ? noalloc_histogram_launch<16>(tuples_dev, n, network_fanout_bits, (uint32_t)nchunks, matrix, status,
                                            hist_out_dev, sm_count, st)
             : noalloc_histogram_launch<8>(tuples_dev, n, network_fanout_bits, (uint32_t)nchunks, matrix, status,
                                           hist_out_dev, sm_count, st);
}
