// rhj_kernels.cuh -- the four join stages as sm_100a kernels.
//
//   K1  k_histogram      histograms/LocalHistogram.cpp:35-53 (and LocalPartitioning.cpp:138-163)
//   Ks  k_colscan        per-chunk exclusive offsets (what lets the scatter write exact,
//                        write-once positions; OffsetMap.cpp:75-85 does this per rank only)
//   Kp  k_plan_network   GlobalHistogram.cpp:31-48, AssignmentMap.cpp:34-49, OffsetMap.cpp:59-93
//   K2/K3 k_scatter      NetworkPartitioning.cpp:116-173 / LocalPartitioning.cpp:194-250
//   Kq  k_plan_local     LocalPartitioning.cpp:165-192 (8-tuple padded prefix) + work list
//   K4  k_build_probe    BuildProbe.cpp:47-121
//
// All stages are HBM-bound integer work (no tensor cores).  Work is cut into
// fixed-size chunks of consecutive tuples; the histogram and the scatter of a pass
// use the same chunking, so a (chunk, bin) count from the histogram becomes an
// exact output offset for the scatter and every output slot is written once.
#pragma once

#include "rhj_device.cuh"

namespace rhj {

constexpr int kHistTPB = 512;
constexpr int kHistUnroll = 8;
constexpr int kScatTPB = 512;
constexpr int kScatIPT = 8;
constexpr int kScatTile = kScatTPB * kScatIPT;  // tuples staged in shared memory per tile
constexpr uint32_t kInvalidDigit = 0xFFFFFFFFu;

enum DevStatus : int {
  kDevOk = 0,
  kDevWindowOverflow = 1,
  kDevLocalOverflow = 2,
};

// ---------------------------------------------------------------------------------
// chunk -> tuple range
// ---------------------------------------------------------------------------------
struct ChunkMapFlat {  // one segment: the rank's whole input relation
  uint64_t n;
  uint32_t chunk;
  uint32_t nchunks;
  __device__ __forceinline__ uint32_t count() const { return nchunks; }
  __device__ __forceinline__ void range(uint32_t c, uint64_t &b, uint64_t &e, uint32_t &seg) const {
    b = (uint64_t)c * chunk;
    e = b + chunk < n ? b + chunk : n;
    seg = 0;
  }
};

struct ChunkMapSeg {  // one segment per owned network partition (chunks never straddle partitions)
  const uint64_t *part_off;  // [nparts] first tuple of the partition in the window
  const uint64_t *part_cnt;  // [nparts]
  const uint32_t *cstart;    // [nparts+1] chunk prefix
  uint32_t nparts;
  uint32_t chunk;
  __device__ __forceinline__ uint32_t count() const { return cstart[nparts]; }
  __device__ __forceinline__ void range(uint32_t c, uint64_t &b, uint64_t &e, uint32_t &seg) const {
    uint32_t j = upper_bound_idx(cstart, nparts, c);
    uint64_t lo = part_off[j], hi = lo + part_cnt[j];
    b = lo + (uint64_t)(c - cstart[j]) * chunk;
    e = b + chunk < hi ? b + chunk : hi;
    seg = j;
  }
};

// ---------------------------------------------------------------------------------
// tuple sources
// ---------------------------------------------------------------------------------
template <int TB>
struct InputSource {  // pass 1: Tuple -> (compressed value, key & (P-1))
  const void *data;
  uint32_t mask, net_bits, key_shift;
  __device__ __forceinline__ uint32_t digit(uint64_t i) const {
    return (uint32_t)InputTuple<TB>::load_key(data, i) & mask;
  }
  __device__ __forceinline__ void load(uint64_t i, uint64_t &value, uint32_t &dg) const {
    uint64_t key, rid;
    InputTuple<TB>::load(data, i, key, rid);
    dg = (uint32_t)key & mask;
    value = compress_tuple(key, rid, net_bits, key_shift);
  }
};

struct ValueSource {  // pass 2: CompressedTuple, digit = bits [key_shift, key_shift + loc_bits)
  const uint64_t *data;
  uint32_t shift, mask;
  __device__ __forceinline__ uint32_t digit(uint64_t i) const {
    return (uint32_t)(ld_stream_u64(data + i) >> shift) & mask;
  }
  __device__ __forceinline__ void load(uint64_t i, uint64_t &value, uint32_t &dg) const {
    value = ld_stream_u64(data + i);
    dg = (uint32_t)(value >> shift) & mask;
  }
};

// ---------------------------------------------------------------------------------
// K1: per-chunk histogram.  chunk_hist is [chunk][nbins] u32.
// Shared memory: nrep replicated histograms (one per group of warps) to spread the
// atomics; AGG additionally merges equal digits inside a warp (match.any) so that a
// single hot digit costs one atomic per warp instead of 32 serialised ones.
// ---------------------------------------------------------------------------------
template <class Source, class Map, bool AGG>
__global__ void __launch_bounds__(kHistTPB) k_histogram(Source src, Map cm, uint32_t nbins, uint32_t nrep,
                                                         uint32_t *__restrict__ chunk_hist,
                                                         const int *__restrict__ status) {
  extern __shared__ uint32_t sh_hist[];
  if (*status != kDevOk) return;
  const uint32_t tid = threadIdx.x, lane = tid & 31;
  uint32_t *my = sh_hist + ((tid >> 5) % nrep) * nbins;
  const uint32_t nchunks = cm.count();
  for (uint32_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
    for (uint32_t i = tid; i < nbins * nrep; i += kHistTPB) sh_hist[i] = 0;
    __syncthreads();
    uint64_t b, e;
    uint32_t seg;
    cm.range(c, b, e, seg);
    for (uint64_t base = b; base < e; base += (uint64_t)kHistTPB * kHistUnroll) {
      uint32_t dg[kHistUnroll];
#pragma unroll
      for (int u = 0; u < kHistUnroll; ++u) {
        uint64_t i = base + (uint64_t)u * kHistTPB + tid;
        dg[u] = i < e ? src.digit(i) : kInvalidDigit;
      }
#pragma unroll
      for (int u = 0; u < kHistUnroll; ++u) {
        if (AGG) {
          uint32_t peers = __match_any_sync(0xFFFFFFFFu, dg[u]);
          if (dg[u] != kInvalidDigit && lane == (uint32_t)(__ffs(peers) - 1)) atomicAdd(&my[dg[u]], __popc(peers));
        } else {
          if (dg[u] != kInvalidDigit) atomicAdd(&my[dg[u]], 1u);
        }
      }
    }
    __syncthreads();
    for (uint32_t i = tid; i < nbins; i += kHistTPB) {
      uint32_t s = 0;
      for (uint32_t r = 0; r < nrep; ++r) s += sh_hist[r * nbins + i];
      chunk_hist[(uint64_t)c * nbins + i] = s;
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------
// Ks: column-wise exclusive scan of a [rows][ncols] u32 matrix, per row segment.
// grid = (ceil(ncols/32), nsegs); block = WARPS*32.  Lane = column (coalesced rows),
// warps split the segment's rows; three phases: partial column sums, scan across
// warps in shared memory, rewrite with running prefix.  totals[seg][col] = column sum.
// seg_start == nullptr: a single segment of `nrows_single` rows.
// ---------------------------------------------------------------------------------
template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32) k_colscan(uint32_t *__restrict__ mat, uint32_t ncols,
                                                         const uint32_t *__restrict__ seg_start,
                                                         uint32_t nrows_single, uint64_t *__restrict__ totals,
                                                         const int *__restrict__ status) {
  __shared__ uint32_t part[WARPS][32];
  if (*status != kDevOk) return;
  const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint32_t col = blockIdx.x * 32 + lane;
  const uint32_t seg = blockIdx.y;
  uint32_t r0 = 0, r1 = nrows_single;
  if (seg_start) {
    r0 = seg_start[seg];
    r1 = seg_start[seg + 1];
  }
  const uint32_t nrows = r1 - r0;
  const uint32_t rpw = (nrows + WARPS - 1) / WARPS;
  uint32_t a = r0 + w * rpw, z = a + rpw;
  if (a > r1) a = r1;
  if (z > r1) z = r1;
  const bool live = col < ncols;
  uint32_t sum = 0;
  if (live) {
#pragma unroll 8
    for (uint32_t r = a; r < z; ++r) sum += mat[(uint64_t)r * ncols + col];
  }
  part[w][lane] = sum;
  __syncthreads();
  uint32_t excl = 0, total = 0;
#pragma unroll
  for (int k = 0; k < WARPS; ++k) {
    uint32_t v = part[k][lane];
    if (k < (int)w) excl += v;
    total += v;
  }
  if (live) {
    uint32_t run = excl;
#pragma unroll 8
    for (uint32_t r = a; r < z; ++r) {
      uint64_t idx = (uint64_t)r * ncols + col;
      uint32_t v = mat[idx];
      mat[idx] = run;
      run += v;
    }
    if (w == 0) totals[(uint64_t)seg * ncols + col] = total;
  }
}

// ---------------------------------------------------------------------------------
// Kp: everything tasks/HistogramComputation.cpp computes after the local histograms,
// for both relations, from the all-gathered [nranks][2][P] matrix.  One block.
// Also lays out the work of the later stages for the partitions this rank owns.
// ---------------------------------------------------------------------------------
struct PlanNetworkArgs {
  const uint64_t *gathered;  // [nranks][2][P]
  uint32_t nranks, rank, parts;
  uint64_t window_cap[2];
  uint64_t peer_window[2][64];  // byte address of every rank's window, per relation
  // outputs, per relation k: arrays of P entries at base + k*P
  uint64_t *global_hist, *base, *rel, *abs, *dest;
  uint32_t *assignment;    // [P]
  uint64_t *window_sizes;  // [2][nranks]
  // owned partitions (index j <-> partition rank + j*nranks)
  uint32_t nowned;
  uint64_t *part_off;   // [2][nowned]
  uint64_t *part_cnt;   // [2][nowned]
  uint32_t *cstart;     // [2][nowned+1] chunk prefix of the local pass
  uint32_t chunk2;      // tuples per local-pass chunk
  uint32_t max_chunks2[2];
  int *status;
};

__global__ void __launch_bounds__(1024) k_plan_network(PlanNetworkArgs a) {
  extern __shared__ uint64_t sh_scan[];  // [P + 1]
  __shared__ uint64_t tmp64[33];
  __shared__ uint32_t tmp32[33];
  const uint32_t P = a.parts, n = a.nranks, me = a.rank, tid = threadIdx.x;
  const uint32_t q = P / n, m = P % n;  // owner r holds q + (r < m) partitions
  for (int k = 0; k < 2; ++k) {
    // GlobalHistogram.cpp:37-42 (sum over ranks) and OffsetMap.cpp:78-83 (exclusive scan over ranks)
    for (uint32_t p = tid; p < P; p += blockDim.x) {
      uint64_t g = 0, before = 0;
      for (uint32_t r = 0; r < n; ++r) {
        uint64_t v = a.gathered[((uint64_t)r * 2 + k) * P + p];
        if (r < me) before += v;
        g += v;
      }
      a.global_hist[k * P + p] = g;
      a.rel[k * P + p] = before;
      if (k == 0) a.assignment[p] = p % n;  // AssignmentMap.cpp:41-43
    }
    __syncthreads();
    // OffsetMap.cpp:59-73: running sum per owner == exclusive scan in owner-major order
    uint64_t carry = 0;
    for (uint32_t i0 = 0; i0 < P; i0 += blockDim.x) {
      uint32_t i = i0 + tid;
      uint64_t v = 0;
      if (i < P) {
        uint32_t r, j;
        if (i < m * (q + 1)) { r = i / (q + 1); j = i % (q + 1); }
        else { r = m + (i - m * (q + 1)) / q; j = (i - m * (q + 1)) % q; }
        v = a.global_hist[k * P + r + j * n];
      }
      uint64_t total;
      uint64_t ex = block_exclusive_scan<uint64_t>(v, tmp64, &total);
      if (i < P) sh_scan[i] = carry + ex;
      carry += total;
    }
    if (tid == 0) sh_scan[P] = carry;
    __syncthreads();
    for (uint32_t p = tid; p < P; p += blockDim.x) {
      uint32_t r = p % n, j = p / n;
      uint32_t ostart = r * q + (r < m ? r : m);
      uint64_t bs = sh_scan[ostart + j] - sh_scan[ostart];
      uint64_t ab = bs + a.rel[k * P + p];
      a.base[k * P + p] = bs;
      a.abs[k * P + p] = ab;  // OffsetMap.cpp:87-93
      a.dest[k * P + p] = a.peer_window[k][r] + ab * 8;
    }
    for (uint32_t r = tid; r < n; r += blockDim.x) {
      uint32_t ostart = r * q + (r < m ? r : m);
      uint32_t oend = (r + 1) * q + (r + 1 < m ? r + 1 : m);
      uint64_t ws = sh_scan[oend] - sh_scan[ostart];  // data/Window.cpp:168-178
      a.window_sizes[k * n + r] = ws;
      if (ws > a.window_cap[k]) atomicMax(a.status, (int)kDevWindowOverflow);
    }
    __syncthreads();
    // owned partitions and their local-pass chunking
    uint32_t ccarry = 0;
    for (uint32_t j0 = 0; j0 < a.nowned; j0 += blockDim.x) {
      uint32_t j = j0 + tid;
      uint32_t nch = 0;
      if (j < a.nowned) {
        uint32_t p = me + j * n;
        uint64_t cnt = a.global_hist[k * P + p];
        uint32_t r = p % n, jj = p / n;
        uint32_t ostart = r * q + (r < m ? r : m);
        a.part_off[k * a.nowned + j] = sh_scan[ostart + jj] - sh_scan[ostart];
        a.part_cnt[k * a.nowned + j] = cnt;
        nch = (uint32_t)((cnt + a.chunk2 - 1) / a.chunk2);
      }
      uint32_t total;
      uint32_t ex = block_exclusive_scan<uint32_t>(nch, tmp32, &total);
      if (j < a.nowned) a.cstart[k * (a.nowned + 1) + j] = ccarry + ex;
      ccarry += total;
    }
    if (tid == 0) {
      a.cstart[k * (a.nowned + 1) + a.nowned] = ccarry;
      if (ccarry > a.max_chunks2[k]) atomicMax(a.status, (int)kDevLocalOverflow);
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------
// K2 / K3: shared-memory staged radix scatter.
// One tile = kScatTile tuples: load -> rank inside the tile with shared atomics ->
// scan bins -> place into the staging buffer ordered by bin -> write each bin's run
// with consecutive threads on consecutive addresses.  Runs of one bin from successive
// tiles of a chunk are adjacent in the destination, so partially filled 32-byte
// sectors are completed in L2 before they are evicted.  Destinations are absolute
// byte addresses and may live on a peer GPU (NVLink stores replace MPI_Put,
// data/Window.cpp:108-115).
//   bin_base  [nsegs][nbins] u64: byte address of this rank's first slot of (seg, bin)
//   chunk_off [chunk][nbins] u32: exclusive per-chunk offsets from k_colscan (tuples)
// ---------------------------------------------------------------------------------
struct ScatterSmem {
  uint64_t *stage;   // [kScatTile]
  uint64_t *delta;   // [nbins] destination address minus 8*start for the current tile
  uint64_t *cursor;  // [nbins] next destination address of the bin for this chunk
  uint32_t *cnt;     // [nbins]
  uint32_t *start;   // [nbins]
  uint16_t *dig;     // [kScatTile]
  uint32_t *tmp;     // [33]
};

__host__ __device__ inline size_t scatter_smem_bytes(uint32_t nbins) {
  return (size_t)kScatTile * 8 + (size_t)nbins * (8 + 8 + 4 + 4) + (size_t)kScatTile * 2 + 64 * 4;
}

template <class Source, class Map>
__global__ void __launch_bounds__(kScatTPB, 2) k_scatter(Source src, Map cm, uint32_t nbins,
                                                          const uint32_t *__restrict__ chunk_off,
                                                          const uint64_t *__restrict__ bin_base,
                                                          const int *__restrict__ status) {
  extern __shared__ __align__(16) unsigned char sh_raw[];
  if (*status != kDevOk) return;
  ScatterSmem s;
  s.stage = reinterpret_cast<uint64_t *>(sh_raw);
  s.delta = s.stage + kScatTile;
  s.cursor = s.delta + nbins;
  s.cnt = reinterpret_cast<uint32_t *>(s.cursor + nbins);
  s.start = s.cnt + nbins;
  s.tmp = s.start + nbins;
  s.dig = reinterpret_cast<uint16_t *>(s.tmp + 64);
  const uint32_t tid = threadIdx.x;
  const uint32_t nchunks = cm.count();

  for (uint32_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
    uint64_t b, e;
    uint32_t seg;
    cm.range(c, b, e, seg);
    for (uint32_t d = tid; d < nbins; d += kScatTPB) {
      s.cursor[d] = bin_base[(uint64_t)seg * nbins + d] + (uint64_t)chunk_off[(uint64_t)c * nbins + d] * 8;
      s.cnt[d] = 0;
    }
    __syncthreads();
    for (uint64_t t0 = b; t0 < e; t0 += kScatTile) {
      const uint32_t tile_n = (e - t0) < (uint64_t)kScatTile ? (uint32_t)(e - t0) : (uint32_t)kScatTile;
      uint64_t val[kScatIPT];
      uint32_t dg[kScatIPT], rk[kScatIPT];
#pragma unroll
      for (int u = 0; u < kScatIPT; ++u) {
        uint32_t i = u * kScatTPB + tid;
        dg[u] = kInvalidDigit;
        if (i < tile_n) src.load(t0 + i, val[u], dg[u]);
      }
#pragma unroll
      for (int u = 0; u < kScatIPT; ++u)
        if (dg[u] != kInvalidDigit) rk[u] = atomicAdd(&s.cnt[dg[u]], 1u);
      __syncthreads();
      // exclusive scan over bins; also advance the chunk cursors and clear the counters
      uint32_t carry = 0;
      for (uint32_t d0 = 0; d0 < nbins; d0 += kScatTPB) {
        uint32_t d = d0 + tid;
        uint32_t v = d < nbins ? s.cnt[d] : 0;
        uint32_t total;
        uint32_t ex = block_exclusive_scan<uint32_t>(v, s.tmp, &total);
        if (d < nbins) {
          uint32_t st = carry + ex;
          uint64_t cur = s.cursor[d];
          s.start[d] = st;
          s.delta[d] = cur - (uint64_t)st * 8;
          s.cursor[d] = cur + (uint64_t)v * 8;
          s.cnt[d] = 0;
        }
        carry += total;
      }
      __syncthreads();
#pragma unroll
      for (int u = 0; u < kScatIPT; ++u) {
        if (dg[u] != kInvalidDigit) {
          uint32_t pos = s.start[dg[u]] + rk[u];
          s.stage[pos] = val[u];
          s.dig[pos] = (uint16_t)dg[u];
        }
      }
      __syncthreads();
#pragma unroll
      for (int u = 0; u < kScatIPT; ++u) {
        uint32_t i = u * kScatTPB + tid;
        if (i < tile_n) {
          uint32_t d = s.dig[i];
          st_stream_u64(reinterpret_cast<uint64_t *>(s.delta[d] + (uint64_t)i * 8), s.stage[i]);
        }
      }
      __syncthreads();
    }
  }
}

// ---------------------------------------------------------------------------------
// Kq: after the local-pass histogram.  One block per owned partition j:
//   sub_cnt[k][j][l]  (from k_colscan totals)  ->  LocalPartitioning.cpp:165-192:
//   prefix[l] = sum_{l'<l} roundup8(cnt[l']);  output region of partition j starts at
//   part_off[j] + j*L*8 (each partition gets its size + L*8 slots, :79-81).
// Also emits the build/probe work list: a sub-partition pair with both sides
// non-empty (:116-124) becomes ceil(|R|/cap) * ceil(|S|/sslice) items.
// With two_level off the "sub-partitions" are the network partitions (L == 1).
// ---------------------------------------------------------------------------------
struct PlanLocalArgs {
  uint32_t nowned, L;
  uint32_t two_level;
  const uint64_t *part_off;  // [2][nowned]
  const uint64_t *part_cnt;  // [2][nowned]
  const uint64_t *sub_tot;   // [2][nowned*L] totals from k_colscan (two_level only)
  uint64_t out_base[2];      // byte address of the local-pass output buffers
  uint64_t win_base[2];      // byte address of the windows
  uint64_t *sub_off;         // [2][nowned*L] tuple index of the sub-partition in its buffer
  uint64_t *sub_prefix;      // [2][nowned*L] padded prefix inside the partition (LocalPartitioning::computePrefixSum)
  uint64_t *sub_addr;        // [2][nowned*L] byte address (scatter destination)
  uint32_t *sub_cnt;         // [2][nowned*L]
  uint64_t *item_lstart;     // [nowned*L] exclusive item prefix inside partition j
  uint64_t *item_jcount;     // [nowned]
  uint32_t cap_r, sslice;
};

__global__ void __launch_bounds__(256) k_plan_local(PlanLocalArgs a) {
  __shared__ uint64_t tmp64[33];
  const uint32_t j = blockIdx.x, tid = threadIdx.x, L = a.L, NS = a.nowned * L;
  uint64_t icarry = 0;
  uint64_t pcarry[2] = {0, 0};
  for (uint32_t l0 = 0; l0 < L; l0 += blockDim.x) {
    uint32_t l = l0 + tid;
    uint32_t cnt[2] = {0, 0};
    for (int k = 0; k < 2; ++k) {
      uint64_t c = 0;
      if (l < L) c = a.two_level ? a.sub_tot[(uint64_t)k * NS + (uint64_t)j * L + l] : a.part_cnt[k * a.nowned + j];
      cnt[k] = (uint32_t)c;
      uint64_t total;
      uint64_t ex = block_exclusive_scan<uint64_t>(a.two_level ? round_up8(c) : c, tmp64, &total);
      if (l < L) {
        uint64_t idx = (uint64_t)k * NS + (uint64_t)j * L + l;
        uint64_t prefix = pcarry[k] + ex;
        uint64_t off = a.two_level ? a.part_off[k * a.nowned + j] + (uint64_t)j * L * 8 + prefix
                                   : a.part_off[k * a.nowned + j];
        a.sub_prefix[idx] = prefix;
        a.sub_off[idx] = off;
        a.sub_addr[idx] = (a.two_level ? a.out_base[k] : a.win_base[k]) + off * 8;
        a.sub_cnt[idx] = cnt[k];
      }
      pcarry[k] += total;
    }
    uint64_t items = 0;
    if (l < L && cnt[0] > 0 && cnt[1] > 0)
      items = (uint64_t)((cnt[0] + a.cap_r - 1) / a.cap_r) * (uint64_t)((cnt[1] + a.sslice - 1) / a.sslice);
    uint64_t itotal;
    uint64_t iex = block_exclusive_scan<uint64_t>(items, tmp64, &itotal);
    if (l < L) a.item_lstart[(uint64_t)j * L + l] = icarry + iex;
    icarry += itotal;
  }
  if (tid == 0) a.item_jcount[j] = icarry;
}

// exclusive scan of item_jcount -> item_jstart[nowned+1]; resets the work counter and result
__global__ void __launch_bounds__(1024) k_plan_items(const uint64_t *__restrict__ item_jcount, uint32_t nowned,
                                                      uint64_t *__restrict__ item_jstart,
                                                      unsigned long long *__restrict__ work_counter) {
  __shared__ uint64_t tmp64[33];
  uint64_t carry = 0;
  for (uint32_t j0 = 0; j0 < nowned; j0 += blockDim.x) {
    uint32_t j = j0 + threadIdx.x;
    uint64_t v = j < nowned ? item_jcount[j] : 0, total;
    uint64_t ex = block_exclusive_scan<uint64_t>(v, tmp64, &total);
    if (j < nowned) item_jstart[j] = carry + ex;
    carry += total;
  }
  if (threadIdx.x == 0) {
    item_jstart[nowned] = carry;
    *work_counter = 0;
  }
}

// ---------------------------------------------------------------------------------
// K4: build + probe.  Persistent CTAs pull work items (sub-partition, build slice,
// probe slice) from a global counter.  The build slice (<= CAP tuples) is held in
// shared memory as the reference's bucket-chained table (BuildProbe.cpp:82-86):
// bucket[] = 1-based index of the newest tuple of the chain, next[] = the one before
// (atomicExch makes the LIFO insert safe for concurrent threads).  Hash bits start at
// shift_bits = F_net + PAYLOAD_BITS + F_loc as in :55-60.  The probe streams the S
// slice with 16-byte loads and walks the chain comparing value >> key_shift (:101-102);
// every duplicate counts.  Per-thread partial results are reduced once per CTA.
// ---------------------------------------------------------------------------------
constexpr int kBpTPB = 256;

struct BuildProbeArgs {
  const uint64_t *src[2];       // buffers the sub-partition offsets index into (R, S)
  const uint64_t *sub_off;      // [2][NS]
  const uint32_t *sub_cnt;      // [2][NS]
  const uint64_t *item_lstart;  // [NS]
  const uint64_t *item_jstart;  // [nowned+1]
  unsigned long long *work_counter;
  unsigned long long *result;   // [5]
  uint32_t nowned, L, NS;
  uint32_t rank, nranks;
  uint32_t net_bits, key_shift, shift_bits;
  uint32_t sslice;
  const int *status;
};

template <int CAP>
__host__ __device__ constexpr size_t build_probe_smem_bytes() {
  return (size_t)CAP * (8 + 4 + 2);
}

template <int CAP>
__global__ void __launch_bounds__(kBpTPB) k_build_probe(BuildProbeArgs a) {
  extern __shared__ __align__(16) unsigned char sh_raw[];
  __shared__ unsigned long long s_item;
  __shared__ unsigned long long s_red[5][kBpTPB / 32];
  if (*a.status != kDevOk) return;
  uint64_t *tup = reinterpret_cast<uint64_t *>(sh_raw);
  uint32_t *bucket = reinterpret_cast<uint32_t *>(tup + CAP);
  uint16_t *next = reinterpret_cast<uint16_t *>(bucket + CAP);
  const uint32_t tid = threadIdx.x;
  const uint64_t rid_mask = a.key_shift >= 64 ? ~0ull : ((1ull << a.key_shift) - 1);
  const uint64_t total_items = a.item_jstart[a.nowned];

  uint64_t acc_n = 0, acc_key = 0, acc_rr = 0, acc_rs = 0, acc_mix = 0;
  uint64_t item = blockIdx.x;
  while (item < total_items) {
    // decode item -> (j, l, build slice, probe slice)
    const uint32_t j = warp_upper_bound(a.item_jstart, a.nowned, item);
    const uint64_t local = item - a.item_jstart[j];
    const uint32_t l = warp_upper_bound(a.item_lstart + (uint64_t)j * a.L, a.L, local);
    const uint32_t sp = j * a.L + l;
    const uint32_t q = (uint32_t)(local - a.item_lstart[sp]);
    const uint32_t cnt_r = a.sub_cnt[sp], cnt_s = a.sub_cnt[a.NS + sp];
    const uint32_t s_slices = (cnt_s + a.sslice - 1) / a.sslice;
    const uint32_t ri = q / s_slices, si = q % s_slices;
    const uint32_t r_begin = ri * CAP, r_n = min((uint32_t)CAP, cnt_r - r_begin);
    const uint32_t s_begin = si * a.sslice, s_n = min(a.sslice, cnt_s - s_begin);
    const uint64_t *rp = a.src[0] + a.sub_off[sp] + r_begin;
    const uint64_t *sp_ptr = a.src[1] + a.sub_off[a.NS + sp] + s_begin;
    const uint32_t partition = a.rank + j * a.nranks;

    // table size: next power of two of the slice (BuildProbe.cpp:59-60 uses the partition size)
    uint32_t nb = 1;
    while (nb < r_n) nb <<= 1;
    const uint32_t hmask = nb - 1;
    for (uint32_t i = tid; i < nb; i += kBpTPB) bucket[i] = 0;
    __syncthreads();

    // ---- build (BuildProbe.cpp:82-86) ----
    {
      const uint32_t head = (reinterpret_cast<uintptr_t>(rp) & 8) ? 1u : 0u;
      const uint32_t h = head < r_n ? head : r_n;
      auto insert = [&](uint64_t v, uint32_t t) {
        tup[t] = v;
        uint32_t idx = (uint32_t)(v >> a.shift_bits) & hmask;
        next[t] = (uint16_t)atomicExch(&bucket[idx], t + 1);
      };
      if (tid < h) insert(ld_stream_u64(rp), 0);
      const ulonglong2 *v2 = reinterpret_cast<const ulonglong2 *>(rp + h);
      const uint32_t nv = (r_n - h) >> 1;
      for (uint32_t i = tid; i < nv; i += kBpTPB * 4) {
        ulonglong2 x[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (i + u * kBpTPB < nv) x[u] = ld_stream_v2(v2 + i + u * kBpTPB);
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (i + u * kBpTPB < nv) {
            uint32_t t = h + 2 * (i + u * kBpTPB);
            insert(x[u].x, t);
            insert(x[u].y, t + 1);
          }
      }
      if (((r_n - h) & 1) && tid == 0) insert(ld_stream_u64(rp + r_n - 1), r_n - 1);
    }
    __syncthreads();

    // ---- probe (BuildProbe.cpp:97-106) ----
    {
      auto probe = [&](uint64_t v) {
        const uint64_t kpart = v >> a.key_shift;
        uint32_t hit = bucket[(uint32_t)(v >> a.shift_bits) & hmask];
        while (hit) {
          const uint64_t r = tup[hit - 1];
          if ((r >> a.key_shift) == kpart) {
            const uint64_t key = (kpart << a.net_bits) | partition;
            const uint64_t rid_r = r & rid_mask, rid_s = v & rid_mask;
            acc_n += 1;
            acc_key += key;
            acc_rr += rid_r;
            acc_rs += rid_s;
            acc_mix += mix_triple(key, rid_r, rid_s);
          }
          hit = next[hit - 1];
        }
      };
      const uint32_t head = (reinterpret_cast<uintptr_t>(sp_ptr) & 8) ? 1u : 0u;
      const uint32_t h = head < s_n ? head : s_n;
      if (tid < h) probe(ld_stream_u64(sp_ptr));
      const ulonglong2 *v2 = reinterpret_cast<const ulonglong2 *>(sp_ptr + h);
      const uint32_t nv = (s_n - h) >> 1;
      for (uint32_t i = tid; i < nv; i += kBpTPB * 4) {
        ulonglong2 x[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (i + u * kBpTPB < nv) x[u] = ld_stream_v2(v2 + i + u * kBpTPB);
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (i + u * kBpTPB < nv) {
            probe(x[u].x);
            probe(x[u].y);
          }
      }
      if (((s_n - h) & 1) && tid == 0) probe(ld_stream_u64(sp_ptr + s_n - 1));
    }
    // next item (the barrier also protects the table against the next build)
    if (tid == 0) s_item = atomicAdd(a.work_counter, 1ull) + gridDim.x;
    __syncthreads();
    item = s_item;
    __syncthreads();
  }

  // one reduction per CTA
  uint64_t v[5] = {acc_n, acc_key, acc_rr, acc_rs, acc_mix};
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    uint64_t w = warp_sum(v[k]);
    if ((tid & 31) == 0) s_red[k][tid >> 5] = w;
  }
  __syncthreads();
  if (tid < 5) {
    unsigned long long s = 0;
    for (int w = 0; w < kBpTPB / 32; ++w) s += s_red[tid][w];
    if (s) atomicAdd(&a.result[tid], s);
  }
}

// ---------------------------------------------------------------------------------
// device generator, bit-identical to oracle orc_perm / orc_fill_perm
// ---------------------------------------------------------------------------------
__host__ __device__ inline uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

struct PermParams {
  uint64_t domain, mask, m0, a0, m1, a1, m2, a2;
  uint32_t shift;
};

__host__ __device__ inline uint64_t perm_apply(uint64_t x, const PermParams &p) {
  if (p.domain <= 1) return 0;
  do {
    x = (x * p.m0 + p.a0) & p.mask;
    x ^= x >> p.shift;
    x = (x * p.m1 + p.a1) & p.mask;
    x ^= x >> p.shift;
    x = (x * p.m2 + p.a2) & p.mask;
    x ^= x >> p.shift;
  } while (x >= p.domain);
  return x;
}

template <int TB>
__global__ void k_fill_perm(void *tuples, uint64_t n, uint64_t first, PermParams pp, uint64_t modulo,
                            uint64_t start_key, uint64_t start_rid) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t k = perm_apply(first + i, pp);
    if (modulo) k %= modulo;
    uint64_t key = start_key + k, rid = start_rid + i;
    if (TB == 16) {
      reinterpret_cast<ulonglong2 *>(tuples)[i] = make_ulonglong2(key, rid);
    } else {
      reinterpret_cast<uint64_t *>(tuples)[i] = (key << 32) | (rid & 0xFFFFFFFFull);
    }
  }
}

}  // namespace rhj
