// rhj_kernels.cuh -- the four join stages as sm_100a kernels.
//
//   K1  k_histogram        histograms/LocalHistogram.cpp:35-53 (and LocalPartitioning.cpp:138-163), per chunk
//   Ks  k_colscan          per-chunk exclusive offsets (what lets the scatter write exact,
//                          write-once positions; OffsetMap.cpp:75-85 does this per rank only)
//   Kp  k_plan_network     GlobalHistogram.cpp:31-48, AssignmentMap.cpp:34-49, OffsetMap.cpp:59-93
//   K2/K3 k_scatter_wc     NetworkPartitioning.cpp:116-173 / LocalPartitioning.cpp:194-250: write-combining
//                          rings in shared memory, only whole 128-byte lines leave the SM
//   Kx  k_push             the staged exchange (three and more ranks): contiguous per-destination runs ->
//                          peer windows (data/Window.cpp:108-115, MPI_Put); k_signal_peers / k_wait_peers:
//                          the barrier of operators/HashJoin.cpp:119-121 as arrival flags
//   Kq  k_plan_local/items LocalPartitioning.cpp:165-192 (8-tuple padded prefix) + work list
//   K4  k_build_probe3     BuildProbe.cpp:47-121
//   k_scatter and k_build_probe are the first-generation kernels.  They run fan-outs of more than 1024
//   bins per pass (the rings stop fitting shared memory), single-pass mode and work items beyond the
//   expanded list (k_build_probe), and stay selectable as the ablation baseline (RHJ_SCATTER=legacy,
//   RHJ_BP=legacy with RHJ_FLAG_DEBUG_ENV).
//
// All stages are HBM-bound integer work (no tensor cores).  Work is cut into
// fixed-size chunks of consecutive tuples; the histogram and the scatter of a pass
// use the same chunking, so a (chunk, bin) count from the histogram becomes an
// exact output offset for the scatter and every output slot is written once.
#pragma once

#include <cooperative_groups.h>

#include "rhj_device.cuh"

namespace rhj {

constexpr int kHistTPB = 512;
constexpr int kHistUnroll = 8;
constexpr int kScatTPB = 512;
constexpr int kScatIPT = 8;
constexpr int kScatTile = kScatTPB * kScatIPT;  // tuples staged in shared memory per tile
constexpr int kChunkGranule = 8192;              // chunk_tuples must be a multiple of this
constexpr uint32_t kInvalidDigit = 0xFFFFFFFFu;

enum DevStatus : int {
  kDevOk = 0,
  kDevWindowOverflow = 1,
  kDevLocalOverflow = 2,
  kDevPeerTimeout = 3,
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
struct RawOf { using type = uint64_t; };
template <>
struct RawOf<16> { using type = ulonglong2; };

template <int TB>
struct InputSource {  // pass 1: Tuple -> (compressed value, key & (P-1))
  const void *data;
  uint32_t mask, net_bits, key_shift;
  using Raw = typename RawOf<TB>::type;  // what one input element looks like in memory
  static constexpr bool kDigitFromValue = false;
  __device__ __forceinline__ const void *base() const { return data; }
  __device__ __forceinline__ void decode(const ulonglong2 &r, uint64_t &value, uint32_t &dg) const {
    dg = (uint32_t)r.x & mask;
    value = compress_tuple(r.x, r.y, net_bits, key_shift);
  }
  __device__ __forceinline__ void decode(const uint64_t &r, uint64_t &value, uint32_t &dg) const {
    const uint64_t key = r >> 32, rid = r & 0xFFFFFFFFull;
    dg = (uint32_t)key & mask;
    value = compress_tuple(key, rid, net_bits, key_shift);
  }
  __device__ __forceinline__ uint32_t digit_of_value(uint64_t) const { return 0; }
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
  using Raw = uint64_t;
  static constexpr bool kDigitFromValue = true;
  __device__ __forceinline__ const void *base() const { return data; }
  __device__ __forceinline__ void decode(const uint64_t &r, uint64_t &value, uint32_t &dg) const {
    value = r;
    dg = (uint32_t)(r >> shift) & mask;
  }
  __device__ __forceinline__ uint32_t digit_of_value(uint64_t v) const { return (uint32_t)(v >> shift) & mask; }
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
// K2 / K3, write-combining: the reference's software write-combining buffers
// (cacheline_t, tasks/NetworkPartitioning.cpp:28-40, 138-141, remainders :186-209;
// tasks/LocalPartitioning.cpp:194-250) restated for an SM.  Every bin owns a small ring
// of C tuples in shared memory.  A tuple takes the next position q of its bin with one
// shared atomic; position q of bin d belongs at byte address base[d] + 8 q, and base[d]
// is 128-byte aligned, so q % 16 == 0 is the start of a destination line.  Between two
// batches of input the warps that own a bin copy every COMPLETE line of it out of the
// ring with 16-byte stores -- eight lanes per 128-byte line, four lines per store
// instruction -- and keep the remainder (< 16 tuples) for the next batch.  Only whole,
// aligned lines leave the SM: no partial sectors for L2 to merge, no read-modify-write
// in DRAM, and a destination in a peer GPU's window sees full-line NVLink writes.  Heads
// and tails (partial lines) are written once per bin and CTA, because a CTA works on a
// CONTIGUOUS run of chunks and the chunk offsets of consecutive chunks are adjacent.
//
// One 32-bit word per bin holds the bin's next position (low half) and the first
// position that no longer fits the ring (high half), so placing a tuple costs a single
// shared-memory transaction besides the store of the tuple itself.  A tuple that does not
// fit (a hot bin under skew) goes straight to its final address with an 8-byte store:
// exact offsets make every position's address known, so overflow needs no second pass.
// Positions are kept small by moving multiples of the rebase unit from the counters into
// base[] (the unit is a multiple of the ring size and of a line, so neither ring slots nor
// line boundaries move): with a compile-time ring size the positions still in the ring are
// below unit + ring, and "position modulo ring size" is one or two conditional subtractions.
//
// No per-tile metadata: the histogram pass only has to deliver per-chunk offsets.
//   bin_base  [nsegs][nbins] u64: byte address of this rank's first slot of (seg, bin)
//   chunk_off [chunk][nbins] u32: exclusive per-chunk offsets from k_colscan (tuples)
// ---------------------------------------------------------------------------------
template <int CFIX>
__host__ __device__ constexpr uint32_t wc_rebase() { return CFIX == 0 ? 16384u : (CFIX % 16 == 0 ? (uint32_t)CFIX : 2u * CFIX); }

__host__ __device__ inline size_t scatter_wc_smem_bytes(uint32_t nbins, uint32_t ring, uint32_t tpb) {
  (void)tpb;
  return (size_t)nbins * ring * 8 + (size_t)nbins * (8 + 4 + 2) + 16;
}

__device__ __forceinline__ void st_stream_v2(void *p, const ulonglong2 &v) {
  asm volatile("st.global.L1::no_allocate.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(v.x), "l"(v.y) : "memory");
}
// shared-memory accesses through pinned 32-bit shared-window addresses (see pin_u32)
__device__ __forceinline__ uint32_t atoms_inc(uint32_t addr) {
  uint32_t old;
  asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(old) : "r"(addr) : "memory");
  return old;
}
__device__ __forceinline__ void sts_u32(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void sts_u16(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"((uint16_t)v) : "memory");
}
__device__ __forceinline__ void sts_u64(uint32_t addr, uint64_t v) {
  asm volatile("st.shared.u64 [%0], %1;" ::"r"(addr), "l"(v) : "memory");
}
__device__ __forceinline__ ulonglong2 lds_v2(uint32_t addr) {
  ulonglong2 v;
  asm volatile("ld.shared.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "r"(addr));
  return v;
}

// CFIX: ring size as a compile-time constant (24 or 48 tuples), or 0 = a power of two given at run time
template <class Source, class Map, int TPB, int IPT, int CFIX>
__global__ void __launch_bounds__(TPB, TPB <= 256 ? 4 : TPB <= 512 ? 2 : 1)
    k_scatter_wc(const Source src, const Map cm, const uint32_t nbins, const uint32_t ring_tuples,
                 const uint32_t *__restrict__ chunk_off, const uint64_t *__restrict__ bin_base,
                 const int *__restrict__ status, const int l2_ahead, const uint32_t c_lo, const uint32_t c_hi) {
  using Raw = typename Source::Raw;
  constexpr uint32_t BATCH = TPB * IPT, NW = TPB / 32, M = wc_rebase<CFIX>();
  constexpr uint32_t FULL = 0xFFFFFFFFu;
  extern __shared__ __align__(128) unsigned char sh_raw[];
  if (*status != kDevOk) return;
  const uint32_t C = CFIX ? (uint32_t)CFIX : ring_tuples;
  uint64_t *const ring = reinterpret_cast<uint64_t *>(sh_raw);      // [nbins][C]
  uint64_t *const base = ring + (size_t)nbins * C;                  // [nbins] line-aligned byte address of position 0
  uint32_t *const word = reinterpret_cast<uint32_t *>(base + nbins);  // [nbins] (limit << 16) | next position
  uint16_t *const head = reinterpret_cast<uint16_t *>(word + nbins);  // [nbins] first position still in the ring
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t cmask = C - 1;
  // ring slot of a position that is still in the ring (q < rebase unit + ring size when CFIX != 0)
  auto slot = [&](uint32_t q) -> uint32_t {
    if constexpr (CFIX == 0) {
      return q & cmask;
    } else {
      if (q >= (uint32_t)CFIX) q -= CFIX;
      if (M > (uint32_t)CFIX && q >= (uint32_t)CFIX) q -= CFIX;
      return q;
    }
  };
  const uint32_t sa_ring = pin_u32(smem_u32(ring)), sa_word = pin_u32(smem_u32(word));
  const uint32_t sa_head = pin_u32(smem_u32(head)), sa_base = pin_u32(smem_u32(base));

  // this launch's chunks [c_lo, min(c_hi, all)) are cut into one contiguous run per CTA
  const uint32_t c_all = cm.count();
  const uint32_t c_top = c_hi < c_all ? c_hi : c_all;
  const uint32_t nchunks = c_top > c_lo ? c_top - c_lo : 0u;
  const uint32_t c_begin = c_lo + (uint32_t)((uint64_t)blockIdx.x * nchunks / gridDim.x);
  const uint32_t c_end = c_lo + (uint32_t)((uint64_t)(blockIdx.x + 1) * nchunks / gridDim.x);
  if (c_begin >= c_end) return;
  for (uint32_t d = tid; d < nbins; d += TPB) {
    word[d] = 0;
    head[d] = 0;
    base[d] = 0;
  }

  // Copy the complete lines of every bin out of the rings (mode 0); at the start of chunk c
  // also bind bins whose destination does not continue where they stand (mode 1: their
  // remainder goes out as a partial line first); mode 2 empties everything.
  // Runs between two barriers: nobody places tuples meanwhile.
  const uint32_t bpw = nbins >= NW ? nbins / NW : 1u;  // bins a warp owns
  const uint32_t first = warp * bpw;
  const uint32_t sub = lane >> 3, sub2 = 2 * (lane & 7);
  auto flush = [&](const int mode, const uint32_t c, const uint32_t seg) {
    for (uint32_t g0 = 0; g0 < bpw; g0 += 32) {
      // lane = bin: what leaves the ring, and the bin's state afterwards
      const uint32_t idx = g0 + lane, bin = first + idx;
      const bool valid = idx < bpw && bin < nbins;
      uint32_t t = 0, fa = 0, fb = 0, nh = 0;
      uint64_t b = 0;
      if (valid) {
        const uint32_t w32 = lds_u32(sa_word + bin * 4);
        t = w32 & 0xFFFFu;
        const uint32_t lim = w32 >> 16;
        const uint32_t h = lds_u16(sa_head + bin * 2);
        const bool all = mode == 2 || t >= lim;
        bool rebind = false;
        uint64_t new_addr = 0;
        if (mode == 1) {
          b = lds_u64(sa_base + bin * 8);
          new_addr = bin_base[(size_t)seg * nbins + bin] + (uint64_t)chunk_off[(size_t)c * nbins + bin] * 8;
          rebind = new_addr != b + (uint64_t)t * 8;  // the remainder goes out first, then the bin is bound anew
        }
        fa = h;
        if (all || rebind) {
          fb = t < lim ? t : lim;  // what the ring holds; later positions went straight to memory
          nh = t;
        } else {
          const uint32_t e = t & ~15u;
          fb = e > h ? e : h;
          nh = fb;
        }
        if (rebind) {  // applied when the state is written back below
          b = new_addr & ~uint64_t(127);
          t = nh = (uint32_t)(new_addr >> 3) & 15u;
        }
      }
      const uint32_t range = fa | (fb << 16);
      // four bins with something to copy at a time: eight lanes per bin, 16 bytes per lane and line
      uint32_t pending = __ballot_sync(FULL, fb > fa);
      while (pending) {
        uint32_t m = pending;
        uint32_t owner = __ffs(m) - 1;
        m &= m - 1;
        if (sub >= 1) owner = __ffs(m) - 1;
        m &= m - 1;
        if (sub >= 2) owner = __ffs(m) - 1;
        m &= m - 1;
        if (sub >= 3) owner = __ffs(m) - 1;
        m &= m - 1;
        pending = m;
        const uint32_t r = __shfl_sync(FULL, range, owner & 31);
        if (owner < 32) {
          const uint32_t fa_i = r & 0xFFFFu, fb_i = r >> 16;
          const uint32_t bin_i = first + g0 + owner;
          const uint64_t b_i = lds_u64(sa_base + bin_i * 8);  // still the value the ring's positions refer to
          const uint32_t sa_bin = sa_ring + bin_i * C * 8;
          for (uint32_t ql = fa_i & ~15u; ql < fb_i; ql += 16) {
            const uint32_t q0 = ql + sub2;
            if (fa_i <= ql && ql + 16 <= fb_i) {
              st_stream_v2(reinterpret_cast<void *>(b_i + (uint64_t)q0 * 8), lds_v2(sa_bin + slot(q0) * 8));
            } else {
              if (fa_i <= q0 && q0 < fb_i)
                st_stream_u64(reinterpret_cast<uint64_t *>(b_i + (uint64_t)q0 * 8), lds_u64(sa_bin + slot(q0) * 8));
              if (fa_i <= q0 + 1 && q0 + 1 < fb_i)
                st_stream_u64(reinterpret_cast<uint64_t *>(b_i + (uint64_t)(q0 + 1) * 8), lds_u64(sa_bin + slot(q0 + 1) * 8));
            }
          }
        }
      }
      __syncwarp();
      if (valid && (mode == 1 || nh != fa)) {
        if (mode != 1) b = lds_u64(sa_base + bin * 8);
        if (nh >= M) {
          const uint32_t k = (nh / M) * M;
          nh -= k;
          t -= k;
          b += (uint64_t)k * 8;
        }
        const uint32_t lim = (nh & ~15u) + C;
        sts_u32(sa_word + bin * 4, (lim << 16) | t);
        sts_u16(sa_head + bin * 2, nh);
        sts_u64(sa_base + bin * 8, b);
      }
    }
  };

  // flush(0) for narrow fan-outs (at most two bins per warp, i.e. the reference's shipped 32 partitions):
  // every bin has several lines to give per batch, so eight lanes simply take one bin, read its state
  // themselves (broadcast loads) and copy its lines -- no votes, no shuffles (-8 % on a 32-bin pass).
  // With more bins per warp the vote of flush() wins: at 256 bins this form was measured 5 % slower,
  // although it executes a third fewer instructions (bins without a full line idle their eight lanes).
  auto flush_fast = [&]() {
    const uint32_t gmask = 0xFFu << (sub * 8);
    for (uint32_t i0 = 0; i0 < bpw; i0 += 4) {
      const uint32_t idx = i0 + sub, bin = first + idx;
      if (idx < bpw && bin < nbins) {
        const uint32_t w32 = lds_u32(sa_word + bin * 4);
        const uint32_t h = lds_u16(sa_head + bin * 2);
        uint32_t t = w32 & 0xFFFFu;
        const uint32_t lim = w32 >> 16;
        uint32_t fb, nh;
        if (t >= lim) {  // ring full (or overflowed): everything it holds goes out
          fb = lim;
          nh = t;
        } else {
          fb = t & ~15u;
          if (fb < h) fb = h;
          nh = fb;
        }
        if (nh != h) {
          uint64_t b = lds_u64(sa_base + bin * 8);
          const uint32_t sa_bin = sa_ring + bin * C * 8;
          for (uint32_t ql = h & ~15u; ql < fb; ql += 16) {
            const uint32_t q0 = ql + sub2;
            if (h <= ql && ql + 16 <= fb) {
              st_stream_v2(reinterpret_cast<void *>(b + (uint64_t)q0 * 8), lds_v2(sa_bin + slot(q0) * 8));
            } else {
              if (h <= q0 && q0 < fb)
                st_stream_u64(reinterpret_cast<uint64_t *>(b + (uint64_t)q0 * 8), lds_u64(sa_bin + slot(q0) * 8));
              if (h <= q0 + 1 && q0 + 1 < fb)
                st_stream_u64(reinterpret_cast<uint64_t *>(b + (uint64_t)(q0 + 1) * 8), lds_u64(sa_bin + slot(q0 + 1) * 8));
            }
          }
          __syncwarp(gmask);  // every lane of the group has read the bin's state before one of them replaces it
          if ((lane & 7) == 0) {
            if (nh >= M) {
              const uint32_t k = (nh / M) * M;
              nh -= k;
              t -= k;
              b += (uint64_t)k * 8;
              sts_u64(sa_base + bin * 8, b);
            }
            sts_u32(sa_word + bin * 4, (((nh & ~15u) + C) << 16) | t);
            sts_u16(sa_head + bin * 2, nh);
          }
        }
      }
    }
  };

  auto load_raw = [&](const Raw *p, uint32_t n, Raw (&raw)[IPT]) {
    p += tid;
    if (n >= BATCH) {
#pragma unroll
      for (int u = 0; u < IPT; ++u) raw[u] = ld_stream(p + u * TPB);
    } else {
#pragma unroll
      for (int u = 0; u < IPT; ++u)
        if (u * TPB + tid < n) raw[u] = ld_stream(p + u * TPB);
    }
  };
  auto put = [&](uint64_t v, uint32_t d, uint32_t w32) {
    const uint32_t q = w32 & 0xFFFFu;
    if (__builtin_expect(q < (w32 >> 16), 1)) sts_u64(sa_ring + (d * C + slot(q)) * 8, v);
    else st_stream_u64(reinterpret_cast<uint64_t *>(base[d] + (uint64_t)q * 8), v);
  };
  auto place = [&](const Raw (&raw)[IPT], uint32_t n) {
    constexpr int G = IPT < 4 ? IPT : 4;  // atomics in flight per thread
    if (n >= BATCH) {
#pragma unroll
      for (int u0 = 0; u0 < IPT; u0 += G) {
        uint64_t v[G];
        uint32_t d[G], w[G];
#pragma unroll
        for (int g = 0; g < G; ++g) {
          src.decode(raw[u0 + g], v[g], d[g]);
          w[g] = atoms_inc(sa_word + d[g] * 4);
        }
#pragma unroll
        for (int g = 0; g < G; ++g) put(v[g], d[g], w[g]);
      }
    } else {
#pragma unroll
      for (int u = 0; u < IPT; ++u) {
        if (u * TPB + tid < n) {
          uint64_t v;
          uint32_t d;
          src.decode(raw[u], v, d);
          put(v, d, atoms_inc(sa_word + d * 4));
        }
      }
    }
  };

  const Raw *const in0 = static_cast<const Raw *>(src.base());
  uint64_t cb, ce;
  uint32_t seg;
  cm.range(c_end - 1, cb, ce, seg);
  const uint64_t in_end = ce;  // the chunks of one CTA are contiguous in the input
  uint32_t c = c_begin;
  cm.range(c, cb, ce, seg);
  uint64_t pos = cb;
  Raw raw[IPT];
  load_raw(in0 + pos, (uint32_t)min((uint64_t)BATCH, ce - pos), raw);
  // L2 prefetch window: [pos + dist, pos + dist + BATCH) with pos = start of the batch in registers;
  // l2_ahead is the distance in quarter batches (measured optimum: one batch; two or more thrash L2)
  const uint64_t l2_dist = (uint64_t)l2_ahead * (BATCH / 4);
  if (tid == 0 && l2_ahead > 0) {
    const uint64_t p = pos + BATCH;
    if (p < in_end) prefetch_l2_range(in0 + p, min(l2_dist, in_end - p) * sizeof(Raw));
  }
  bool chunk_start = true;
  while (true) {
    const uint32_t n = (uint32_t)min((uint64_t)BATCH, ce - pos);
    __syncthreads();
    if (!chunk_start && bpw <= 2) flush_fast();
    else flush(chunk_start ? 1 : 0, c, seg);
    __syncthreads();
    place(raw, n);
    pos += n;
    chunk_start = false;
    if (pos >= ce) {
      if (++c >= c_end) break;
      cm.range(c, cb, ce, seg);
      pos = cb;
      chunk_start = true;
    }
    load_raw(in0 + pos, (uint32_t)min((uint64_t)BATCH, ce - pos), raw);
    if (tid == 0 && l2_ahead > 0) {
      const uint64_t p = pos + l2_dist;
      if (p < in_end) prefetch_l2_range(in0 + p, min((uint64_t)BATCH, in_end - p) * sizeof(Raw));
    }
  }
  __syncthreads();
  flush(2, 0, 0);
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
  const uint64_t *gathered;  // [nranks][2][P], or [nranks][P] of one relation (rank_stride, rel_stride below)
  uint32_t rank_stride, rel_stride;  // element (r, k, p) = gathered[r * rank_stride + k * rel_stride + p]
  int k_begin, k_end;                // relations planned by this launch
  uint32_t nranks, rank, parts;
  uint64_t window_cap[2];
  uint64_t peer_window[2][64];  // byte address of every rank's window, per relation
  // outputs, per relation k: arrays of P entries at base + k*P
  uint64_t *global_hist, *base, *rel, *abs, *dest;
  uint32_t *assignment;    // [P]
  uint64_t *window_sizes;  // [2][nranks]
  // Partition -> owner.  0: p % n (AssignmentMap.cpp:41-43).  1: size-balanced (no counterpart in
  // the reference's built code): partitions by descending global inner+outer count go to the
  // least-loaded rank that still has a free slot; every rank keeps exactly the number of
  // partitions p % n would give it, so all workspace sizes stay static.
  uint32_t balanced;
  uint32_t *owned_list;  // [P] partitions in owner-major order, ascending inside one owner
  uint32_t *slot_of;     // [P] index of the partition among its owner's partitions
  // owned partitions (index j <-> partition owned_list[first slot of this rank + j])
  uint32_t nowned;
  uint64_t *part_off;   // [2][nowned]
  uint64_t *part_cnt;   // [2][nowned]
  uint32_t *cstart;     // [2][nowned+1] chunk prefix of the local pass
  uint32_t chunk2;      // tuples per local-pass chunk
  uint32_t max_chunks2[2];
  int *status;
  // staged exchange (nranks > 1): tuples for remote partitions are first gathered in a local
  // send buffer, contiguous per partition, and pushed to the owner with coalesced stores
  uint64_t stage_base[2];                  // byte address of the send buffers, 0 = store directly into peers
  uint64_t *push_src, *push_dst, *push_cnt;  // [2][P]; cnt == 0 for partitions this rank owns
};

__global__ void __launch_bounds__(1024) k_plan_network(PlanNetworkArgs a) {
  extern __shared__ uint64_t sh_scan[];  // [P + 1]
  __shared__ uint64_t tmp64[33];
  __shared__ uint32_t tmp32[33];
  const uint32_t P = a.parts, n = a.nranks, me = a.rank, tid = threadIdx.x;
  const uint32_t q = P / n, m = P % n;  // owner r holds q + (r < m) partitions
  auto first_slot = [&](uint32_t r) { return r * q + (r < m ? r : m); };
  // GlobalHistogram.cpp:37-42 (sum over ranks) and OffsetMap.cpp:78-83 (exclusive scan over ranks)
  auto local_count = [&](uint32_t r, int k, uint32_t p) -> uint64_t {
    return a.gathered[(uint64_t)r * a.rank_stride + (uint64_t)k * a.rel_stride + p];
  };
  for (int k = a.k_begin; k < a.k_end; ++k)
    for (uint32_t p = tid; p < P; p += blockDim.x) {
      uint64_t g = 0, before = 0;
      for (uint32_t r = 0; r < n; ++r) {
        uint64_t v = local_count(r, k, p);
        if (r < me) before += v;
        g += v;
      }
      a.global_hist[k * P + p] = g;
      a.rel[k * P + p] = before;
    }
  __syncthreads();
  if (!a.balanced || n == 1) {
    for (uint32_t p = tid; p < P; p += blockDim.x) {
      a.assignment[p] = p % n;  // AssignmentMap.cpp:41-43
      a.slot_of[p] = p / n;
      a.owned_list[first_slot(p % n) + p / n] = p;
    }
  } else {
    // sort (weight descending, partition ascending): ascending bitonic sort of (~weight, p) keys
    constexpr uint64_t kWMax = 0xFFFFFFFFFFFFull;
    for (uint32_t p = tid; p < P; p += blockDim.x) {
      const uint64_t w = a.global_hist[p] + a.global_hist[P + p];
      sh_scan[p] = ((kWMax - (w < kWMax ? w : kWMax)) << 16) | p;
    }
    __syncthreads();
    for (uint32_t size = 2; size <= P; size <<= 1)
      for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
        for (uint32_t t = tid; t < P / 2; t += blockDim.x) {
          const uint32_t lo = 2 * t - (t & (stride - 1)), hi = lo + stride;
          const bool up = (lo & size) == 0;
          const uint64_t x = sh_scan[lo], y = sh_scan[hi];
          if ((x > y) == up) {
            sh_scan[lo] = y;
            sh_scan[hi] = x;
          }
        }
        __syncthreads();
      }
    if (tid < 32) {  // greedy placement; lane l keeps the load and free slots of ranks l and l + 32
      const uint32_t lane = tid;
      uint64_t load0 = 0, load1 = 0;
      uint32_t room0 = lane < n ? q + (lane < m ? 1u : 0u) : 0u;
      uint32_t room1 = lane + 32 < n ? q + (lane + 32 < m ? 1u : 0u) : 0u;
      for (uint32_t i = 0; i < P; ++i) {
        const uint64_t key = sh_scan[i];
        const uint32_t p = (uint32_t)(key & 0xFFFF);
        const uint64_t w = kWMax - (key >> 16);
        uint64_t c = ~0ull;
        if (room0) c = (load0 << 6) | lane;
        if (room1 && ((load1 << 6) | (lane + 32)) < c) c = (load1 << 6) | (lane + 32);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
          const uint64_t o = __shfl_xor_sync(0xFFFFFFFFu, c, off);
          c = o < c ? o : c;
        }
        const uint32_t best = (uint32_t)(c & 63);
        if (best == lane) {
          load0 += w;
          room0 -= 1;
        } else if (best == lane + 32) {
          load1 += w;
          room1 -= 1;
        }
        if (lane == 0) a.assignment[p] = best;
      }
    }
    __syncthreads();
    if (tid < n) {  // slots in ascending partition order
      uint32_t c = 0;
      const uint32_t s0 = first_slot(tid);
      for (uint32_t p = 0; p < P; ++p)
        if (a.assignment[p] == tid) {
          a.slot_of[p] = c;
          a.owned_list[s0 + c] = p;
          ++c;
        }
    }
  }
  __syncthreads();
  for (int k = a.k_begin; k < a.k_end; ++k) {
    // OffsetMap.cpp:59-73: running sum per owner == exclusive scan in owner-major order
    uint64_t carry = 0;
    for (uint32_t i0 = 0; i0 < P; i0 += blockDim.x) {
      uint32_t i = i0 + tid;
      uint64_t v = i < P ? a.global_hist[k * P + a.owned_list[i]] : 0;
      uint64_t total;
      uint64_t ex = block_exclusive_scan<uint64_t>(v, tmp64, &total);
      if (i < P) sh_scan[i] = carry + ex;
      carry += total;
    }
    if (tid == 0) sh_scan[P] = carry;
    __syncthreads();
    for (uint32_t p = tid; p < P; p += blockDim.x) {
      const uint32_t r = a.assignment[p], ostart = first_slot(r);
      uint64_t bs = sh_scan[ostart + a.slot_of[p]] - sh_scan[ostart];
      uint64_t ab = bs + a.rel[k * P + p];
      a.base[k * P + p] = bs;
      a.abs[k * P + p] = ab;  // OffsetMap.cpp:87-93
      a.dest[k * P + p] = a.peer_window[k][r] + ab * 8;
    }
    for (uint32_t r = tid; r < n; r += blockDim.x) {
      uint64_t ws = sh_scan[first_slot(r + 1)] - sh_scan[first_slot(r)];  // data/Window.cpp:168-178
      a.window_sizes[k * n + r] = ws;
      if (ws > a.window_cap[k]) atomicMax(a.status, (int)kDevWindowOverflow);
    }
    __syncthreads();
    // owned partitions and their local-pass chunking
    uint32_t ccarry = 0;
    const uint32_t mine = first_slot(me);
    for (uint32_t j0 = 0; j0 < a.nowned; j0 += blockDim.x) {
      uint32_t j = j0 + tid;
      uint32_t nch = 0;
      if (j < a.nowned) {
        const uint32_t p = a.owned_list[mine + j];
        uint64_t cnt = a.global_hist[k * P + p];
        a.part_off[k * a.nowned + j] = sh_scan[mine + j] - sh_scan[mine];
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
    if (a.stage_base[k]) {
      // send-buffer offset of partition p = my tuples of the remote partitions before it
      uint64_t scarry = 0;
      for (uint32_t i0 = 0; i0 < P; i0 += blockDim.x) {
        const uint32_t i = i0 + tid;
        const uint64_t v = (i < P && a.assignment[i] != me) ? local_count(me, k, i) : 0;
        uint64_t total;
        const uint64_t ex = block_exclusive_scan<uint64_t>(v, tmp64, &total);
        if (i < P) sh_scan[i] = scarry + ex;
        scarry += total;
      }
      __syncthreads();
      for (uint32_t p = tid; p < P; p += blockDim.x) {
        const bool remote = a.assignment[p] != me;
        const uint64_t src = a.stage_base[k] + sh_scan[p] * 8;
        a.push_cnt[k * P + p] = remote ? local_count(me, k, p) : 0;
        a.push_src[k * P + p] = src;
        a.push_dst[k * P + p] = a.dest[k * P + p];
        if (remote) a.dest[k * P + p] = src;  // the scatter kernel fills the send buffer instead
      }
      __syncthreads();
    }
  }
}

// ---------------------------------------------------------------------------------
// Push: copy every remote partition's run from the local send buffer into the owner's
// window (replaces MPI_Put, data/Window.cpp:108-115).  Runs are
// megabytes long and contiguous on both sides, so each warp store covers whole 128-byte
// lines of the peer's memory: NVLink sees full-line writes instead of the 8-byte-granular
// runs the scatter kernel would produce.
// ---------------------------------------------------------------------------------
// Many small CTAs (256 threads, 8 stores in flight per thread), eight per SM's worth of the SMs the
// concurrent scatter or local-pass kernel leaves free (rhj_ctx::tune_push_sms).  What was measured on
// 8 GPUs (256M x 256M per GPU, ms per join): this form 11.7; 32 CTAs of 1024 threads that own 32 SMs
// 12.4 (the push itself drops to ~620 GB/s: NVLink wants more stores in flight than 32 SMs hold);
// 148 CTAs of 128 threads beside a scatter kernel capped at 56 registers 13.9.
constexpr int kPushTPB = 256;
constexpr int kPushInFlight = 8;
__global__ void __launch_bounds__(kPushTPB) k_push(const uint64_t *__restrict__ src_addr,
                                                    const uint64_t *__restrict__ dst_addr,
                                                    const uint64_t *__restrict__ cnt, const uint32_t nparts,
                                                    const int *__restrict__ status,
                                                    const uint32_t *__restrict__ row_lo,
                                                    const uint32_t *__restrict__ row_hi) {
  if (*status != kDevOk) return;
  constexpr uint32_t kSlice = 8192;  // tuples per work item (64 KB)
  // row_lo / row_hi: exclusive per-chunk offsets (rows of the chunk matrix) that bound the part of every
  // partition's run to copy -- the tuples of one group of chunks; nullptr = from the start / to the end
  for (uint32_t p = 0; p < nparts; ++p) {
    if (cnt[p] == 0) continue;
    const uint64_t lo = row_lo ? row_lo[p] : 0, hi = row_hi ? row_hi[p] : cnt[p];
    const uint64_t n = hi - lo;
    if (n == 0) continue;
    const uint64_t src_a = src_addr[p] + lo * 8, dst_a = dst_addr[p] + lo * 8;
    const uint64_t *src = reinterpret_cast<const uint64_t *>(src_a);
    uint64_t *dst = reinterpret_cast<uint64_t *>(dst_a);
    // head: up to the first 128-byte boundary of the destination, then slices of whole lines
    uint64_t head = ((128 - (dst_a & 127)) & 127) / 8;
    if (head > n) head = n;
    if (blockIdx.x == p % gridDim.x && threadIdx.x < head) dst[threadIdx.x] = ld_stream_u64(src + threadIdx.x);
    const uint64_t nslices = (n - head + kSlice - 1) / kSlice;
    // start each partition at a different CTA so that short runs do not all land on CTA 0
    for (uint64_t sl = (blockIdx.x + gridDim.x - p % gridDim.x) % gridDim.x; sl < nslices; sl += gridDim.x) {
      const uint64_t b = head + sl * kSlice;
      const uint64_t e = b + kSlice < n ? b + kSlice : n;
      for (uint64_t i = b + threadIdx.x; i < e; i += (uint64_t)kPushTPB * kPushInFlight) {
        uint64_t v[kPushInFlight];
#pragma unroll
        for (int u = 0; u < kPushInFlight; ++u)
          if (i + (uint64_t)u * kPushTPB < e) v[u] = ld_stream_u64(src + i + (uint64_t)u * kPushTPB);
#pragma unroll
        for (int u = 0; u < kPushInFlight; ++u)
          if (i + (uint64_t)u * kPushTPB < e) st_stream_u64(dst + i + (uint64_t)u * kPushTPB, v[u]);
      }
    }
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
  extern __shared__ __align__(128) unsigned char sh_raw[];
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
//   roundup8(part_off[j] + j*(L*8+8)): each partition gets its size + L*8 slots (:79-81) and
//   starts on an 8-tuple (64-byte) boundary like the reference's pool allocation.
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
  int k_begin, k_end;  // relations laid out by this launch
  int do_items;        // also emit the work list (needs sub_cnt of BOTH relations in memory)
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
      if (k < a.k_begin || k >= a.k_end) {  // laid out by an earlier launch: only its counts are needed
        if (l < L && a.do_items) cnt[k] = a.sub_cnt[(uint64_t)k * NS + (uint64_t)j * L + l];
        continue;
      }
      uint64_t c = 0;
      if (l < L) c = a.two_level ? a.sub_tot[(uint64_t)k * NS + (uint64_t)j * L + l] : a.part_cnt[k * a.nowned + j];
      cnt[k] = (uint32_t)c;
      uint64_t total;
      uint64_t ex = block_exclusive_scan<uint64_t>(a.two_level ? round_up8(c) : c, tmp64, &total);
      if (l < L) {
        uint64_t idx = (uint64_t)k * NS + (uint64_t)j * L + l;
        uint64_t prefix = pcarry[k] + ex;
        uint64_t off = a.two_level ? round_up8(a.part_off[k * a.nowned + j] + (uint64_t)j * (L * 8 + 8)) + prefix
                                   : a.part_off[k * a.nowned + j];
        a.sub_prefix[idx] = prefix;
        a.sub_off[idx] = off;
        a.sub_addr[idx] = (a.two_level ? a.out_base[k] : a.win_base[k]) + off * 8;
        a.sub_cnt[idx] = cnt[k];
      }
      pcarry[k] += total;
    }
    if (!a.do_items) continue;
    uint64_t items = 0;
    if (l < L && cnt[0] > 0 && cnt[1] > 0)
      items = (uint64_t)((cnt[0] + a.cap_r - 1) / a.cap_r) * (uint64_t)((cnt[1] + a.sslice - 1) / a.sslice);
    uint64_t itotal;
    uint64_t iex = block_exclusive_scan<uint64_t>(items, tmp64, &itotal);
    if (l < L) a.item_lstart[(uint64_t)j * L + l] = icarry + iex;
    icarry += itotal;
  }
  if (tid == 0 && a.do_items) a.item_jcount[j] = icarry;
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
  const uint32_t *owned;  // [nowned] network partition id of the j-th owned partition
  uint32_t net_bits, key_shift, shift_bits;
  uint32_t sslice;
  const int *status;
  const struct ItemDesc *items;  // [items_cap] expanded work list (k_expand_items)
  uint64_t items_cap;
  uint64_t item_begin;  // generic kernel: first item it is responsible for
  // materialised output (EMIT variants): matches append (rid_R, rid_S) at a position reserved
  // from `pair_cursor`; pairs at positions >= pairs_cap are counted but not stored
  ulonglong2 *pairs;
  uint64_t pairs_cap;
  unsigned long long *pair_cursor;
};

// One build/probe work item, fully resolved: what a CTA needs to start its loads.
struct __align__(16) ItemDesc {
  uint64_t r_off, s_off;  // tuple index into src[0] / src[1]
  uint32_t r_n, s_n;
  uint32_t partition, pad;
};

// Resolve item -> (sub-partition, build slice, probe slice) by searching the two-level
// item prefix (used to expand the list, and directly for items beyond the list's capacity).
template <int CAP, class Search>
__device__ __forceinline__ ItemDesc resolve_item(const BuildProbeArgs &a, uint64_t item, Search search) {
  const uint32_t j = search(a.item_jstart, a.nowned, item);
  const uint64_t local = item - a.item_jstart[j];
  const uint32_t l = search(a.item_lstart + (uint64_t)j * a.L, a.L, local);
  const uint32_t sp = j * a.L + l;
  const uint32_t q = (uint32_t)(local - a.item_lstart[sp]);
  const uint32_t cnt_r = a.sub_cnt[sp], cnt_s = a.sub_cnt[a.NS + sp];
  const uint32_t s_slices = (cnt_s + a.sslice - 1) / a.sslice;
  const uint32_t ri = q / s_slices, si = q % s_slices;
  ItemDesc d;
  d.r_off = a.sub_off[sp] + (uint64_t)ri * CAP;
  d.s_off = a.sub_off[a.NS + sp] + (uint64_t)si * a.sslice;
  d.r_n = min((uint32_t)CAP, cnt_r - ri * CAP);
  d.s_n = min(a.sslice, cnt_s - si * a.sslice);
  d.partition = a.owned[j];
  d.pad = 0;
  return d;
}

// Expand the work list: one thread per sub-partition writes the descriptors of its items.
template <int CAP>
__global__ void __launch_bounds__(256) k_expand_items(BuildProbeArgs a, ItemDesc *__restrict__ out) {
  const uint32_t sp = blockIdx.x * blockDim.x + threadIdx.x;
  if (sp >= a.NS) return;
  const uint32_t j = sp / a.L;
  const uint32_t cnt_r = a.sub_cnt[sp], cnt_s = a.sub_cnt[a.NS + sp];
  if (cnt_r == 0 || cnt_s == 0) return;
  const uint32_t r_slices = (cnt_r + CAP - 1) / CAP, s_slices = (cnt_s + a.sslice - 1) / a.sslice;
  uint64_t idx = a.item_jstart[j] + a.item_lstart[sp];
  for (uint32_t ri = 0; ri < r_slices; ++ri)
    for (uint32_t si = 0; si < s_slices; ++si, ++idx) {
      if (idx >= a.items_cap) return;
      ItemDesc d;
      d.r_off = a.sub_off[sp] + (uint64_t)ri * CAP;
      d.s_off = a.sub_off[a.NS + sp] + (uint64_t)si * a.sslice;
      d.r_n = min((uint32_t)CAP, cnt_r - ri * CAP);
      d.s_n = min(a.sslice, cnt_s - si * a.sslice);
      d.partition = a.owned[j];
      d.pad = 0;
      out[idx] = d;
    }
}

// Bucket heads of the shared-memory table: the largest power of two <= CAP.  A build slice of up
// to CAP tuples hashes into min(nextpow2(slice), bucket_slots) chains, so a capacity such as 5120
// keeps sub-partitions that come out a little above 4096 tuples in one work item (load factor <=
// 1.25) instead of splitting them and probing their S side twice.
template <int CAP>
__host__ __device__ constexpr uint32_t bucket_slots() {
  uint32_t n = 1;
  while (n * 2 <= (uint32_t)CAP) n *= 2;
  return n;
}

template <int CAP>
__host__ __device__ constexpr size_t build_probe_smem_bytes() {
  return (size_t)CAP * (8 + 2) + (size_t)bucket_slots<CAP>() * 4;
}

// EMIT: each match also reserves one output slot.  Lanes that find a match in the same step
// share one reservation (one atomic per converged group), which is enough for the paths this
// kernel serves: single-pass mode and work items beyond the expanded list.
__device__ __forceinline__ void emit_pair_grouped(const BuildProbeArgs &a, uint64_t rid_r, uint64_t rid_s) {
  const cooperative_groups::coalesced_group g = cooperative_groups::coalesced_threads();
  unsigned long long base = 0;
  if (g.thread_rank() == 0) base = atomicAdd(a.pair_cursor, (unsigned long long)g.size());
  base = g.shfl(base, 0);
  const uint64_t pos = base + g.thread_rank();
  if (pos < a.pairs_cap) a.pairs[pos] = make_ulonglong2(rid_r, rid_s);
}

template <int CAP, bool EMIT = false>
__global__ void __launch_bounds__(kBpTPB) k_build_probe(BuildProbeArgs a) {
  extern __shared__ __align__(128) unsigned char sh_raw[];
  __shared__ unsigned long long s_item;
  __shared__ unsigned long long s_red[5][kBpTPB / 32];
  if (*a.status != kDevOk) return;
  uint64_t *tup = reinterpret_cast<uint64_t *>(sh_raw);
  uint32_t *bucket = reinterpret_cast<uint32_t *>(tup + CAP);
  uint16_t *next = reinterpret_cast<uint16_t *>(bucket + bucket_slots<CAP>());
  const uint32_t tid = threadIdx.x;
  const uint64_t rid_mask = a.key_shift >= 64 ? ~0ull : ((1ull << a.key_shift) - 1);
  const uint64_t total_items = a.item_jstart[a.nowned];

  uint64_t acc_n = 0, acc_key = 0, acc_rr = 0, acc_rs = 0, acc_mix = 0;
  uint64_t item = a.item_begin + blockIdx.x;
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
    const uint32_t partition = a.owned[j];

    // table size: next power of two of the slice (BuildProbe.cpp:59-60 uses the partition size)
    uint32_t nb = 1;
    while (nb < r_n && nb < bucket_slots<CAP>()) nb <<= 1;
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
            if (EMIT) emit_pair_grouped(a, rid_r, rid_s);
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
    if (tid == 0) s_item = a.item_begin + atomicAdd(a.work_counter, 1ull) + gridDim.x;
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
// K4, pipelined (two-level mode: sub-partitions start on 8-tuple boundaries and are
// padded to them, LocalPartitioning.cpp:165-192, so slices are 16-byte aligned and may be
// over-read up to the padding).  Items are assigned to CTAs round-robin from the expanded
// work list.  While item i is built and probed out of shared memory / registers, the
// loads of item i+1 are already in flight: its build slice by one bulk asynchronous copy
// (TMA) straight into the other table buffer, its first probe batch into registers; the
// descriptor of item i+2 is fetched at the same time.  Straight-line code per item is
// kept minimal -- the kernel is bound by instruction issue, not by HBM.
// ---------------------------------------------------------------------------------
constexpr int kBp3SV = 2;  // 16-byte probe vectors per thread per batch

template <int CAP>
__host__ __device__ constexpr size_t build_probe3_smem_bytes() {
  return (size_t)CAP * (8 + 8 + 2) + (size_t)bucket_slots<CAP>() * 4 + 64;
}

template <int CAP, int TPB, bool KS32, bool EMIT = false>
__global__ void __launch_bounds__(TPB, 2048 / TPB >= 4 ? 2 : 1) k_build_probe3(const BuildProbeArgs a) {
  constexpr int BATCH = TPB * kBp3SV * 2;  // probe tuples per batch
  extern __shared__ __align__(128) unsigned char sh_raw[];
  __shared__ unsigned long long s_red[5][TPB / 32];
  __shared__ uint32_t s_wtot[EMIT ? 32 : 1];             // EMIT: matches per warp in this batch
  __shared__ unsigned long long s_wbase[EMIT ? 32 : 1];  //       first output slot of the warp
  if (*a.status != kDevOk) return;
  uint64_t *const tup0 = reinterpret_cast<uint64_t *>(sh_raw);
  uint32_t *const bucket = reinterpret_cast<uint32_t *>(tup0 + 2 * CAP);
  uint16_t *const next = reinterpret_cast<uint16_t *>(bucket + bucket_slots<CAP>());
  uint64_t *const bars = reinterpret_cast<uint64_t *>(next + CAP);
  static_assert(CAP % 4 == 0, "mbarriers behind next[] must stay 8-byte aligned");
  const uint32_t tid = threadIdx.x;
  const uint64_t rid_mask = a.key_shift >= 64 ? ~0ull : ((1ull << a.key_shift) - 1);
  const uint32_t hash_shift32 = a.shift_bits - 32;  // KS32: hash bits start inside the high word
  const uint64_t total = a.item_jstart[a.nowned] < a.items_cap ? a.item_jstart[a.nowned] : a.items_cap;
  const uint32_t stride = gridDim.x;

  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_fence_init();
  }
  __syncthreads();

  // descriptor = two 16-byte words {r_off, s_off}, {r_n | s_n << 32, partition}
  auto fetch = [&](uint64_t item, ulonglong2 &lo, ulonglong2 &hi) {
    const ulonglong2 *p = reinterpret_cast<const ulonglong2 *>(a.items + item);
    lo = p[0];
    hi = p[1];
  };
  auto issue = [&](const ulonglong2 &lo, const ulonglong2 &hi, uint32_t slot, ulonglong2 (&xs)[kBp3SV]) {
    const uint32_t r_n = (uint32_t)hi.x, s_n = (uint32_t)(hi.x >> 32);
    if (tid == 0) {
      const uint32_t bytes = ((r_n + 1) & ~1u) * 8;  // padded slice: even tuple count
      mbar_arrive_expect_tx(&bars[slot], bytes);
      bulk_g2s(tup0 + slot * CAP, a.src[0] + lo.x, bytes, &bars[slot]);
      // probe batches after the first are loaded right before use: have them in L2 by then
      if (s_n > (uint32_t)BATCH) prefetch_l2_range(a.src[1] + lo.y + BATCH, (uint64_t)(s_n - BATCH) * 8);
    }
    const ulonglong2 *v2 = reinterpret_cast<const ulonglong2 *>(a.src[1] + lo.y) + tid;
#pragma unroll
    for (int u = 0; u < kBp3SV; ++u)
      if (2 * (u * TPB + tid) < s_n) xs[u] = ld_stream_v2(v2 + u * TPB);
  };

  uint64_t id = blockIdx.x;
  ulonglong2 c_lo, c_hi, n_lo, n_hi;
  ulonglong2 xs_cur[kBp3SV], xs_nxt[kBp3SV];
  bool v_cur = id < total, v_nxt = id + stride < total;
  if (v_cur) {
    fetch(id, c_lo, c_hi);
    issue(c_lo, c_hi, 0, xs_cur);
  }
  if (v_nxt) fetch(id + stride, n_lo, n_hi);

  uint64_t acc_n = 0, acc_key = 0, acc_rr = 0, acc_rs = 0, acc_mix = 0;
  uint32_t slot = 0, phase = 0;  // bit s of `phase` = parity to wait for on bars[s]
  while (v_cur) {
    ulonglong2 f_lo, f_hi;
    const bool v_far = id + 2 * (uint64_t)stride < total;
    if (v_far) fetch(id + 2 * (uint64_t)stride, f_lo, f_hi);
    if (v_nxt) issue(n_lo, n_hi, slot ^ 1, xs_nxt);

    const uint32_t r_n = (uint32_t)c_hi.x, s_n = (uint32_t)(c_hi.x >> 32), partition = (uint32_t)c_hi.y;
    const uint32_t nb = r_n <= 1 ? 1u : min(1u << (32 - __clz(r_n - 1)), bucket_slots<CAP>());
    const uint32_t hmask = nb - 1;
    for (uint32_t i = tid; i < nb; i += TPB) bucket[i] = 0;
    __syncthreads();
    mbar_wait(&bars[slot], (phase >> slot) & 1);
    phase ^= 1u << slot;

    // ---- build (BuildProbe.cpp:82-86) ----
    const uint64_t *const t = tup0 + slot * CAP;
    for (uint32_t i = tid; i < r_n; i += TPB) {
      const uint32_t idx = KS32 ? (((uint32_t)(t[i] >> 32) >> hash_shift32) & hmask)
                                : ((uint32_t)(t[i] >> a.shift_bits) & hmask);
      next[i] = (uint16_t)atomicExch(&bucket[idx], i + 1);
    }
    __syncthreads();

    // ---- probe (BuildProbe.cpp:97-106) ----
    uint32_t item_n = 0;
    uint64_t item_kpart = 0;
    // shared-window addresses of the table, biased for the 1-based chain indices
    const uint32_t sa_bucket = pin_u32(smem_u32(bucket));
    const uint32_t sa_t = pin_u32(smem_u32(t) - 8);
    const uint32_t sa_next = pin_u32(smem_u32(next) - 2);
    auto probe = [&](uint64_t v) {
      if (KS32) {  // rid = low word, key part = high word
        const uint32_t vhi = (uint32_t)(v >> 32), vlo = (uint32_t)v;
        uint32_t hit = lds_u32(sa_bucket + ((vhi >> hash_shift32) & hmask) * 4);
        while (hit) {
          const uint64_t r = lds_u64(sa_t + hit * 8);
          const uint32_t nx = lds_u16(sa_next + hit * 2);
          if ((uint32_t)(r >> 32) == vhi) {
            const uint32_t rlo = (uint32_t)r;
            item_n += 1;
            item_kpart += vhi;
            acc_rr += rlo;
            acc_rs += vlo;
            acc_mix += mix_triple(((uint64_t)vhi << a.net_bits) | partition, rlo, vlo);
          }
          hit = nx;
        }
      } else {
        const uint64_t kpart = v >> a.key_shift;
        uint32_t hit = bucket[(uint32_t)(v >> a.shift_bits) & hmask];
        while (hit) {
          const uint64_t r = t[hit - 1];
          if ((r >> a.key_shift) == kpart) {
            const uint64_t rid_r = r & rid_mask, rid_s = v & rid_mask;
            item_n += 1;
            item_kpart += kpart;
            acc_rr += rid_r;
            acc_rs += rid_s;
            acc_mix += mix_triple((kpart << a.net_bits) | partition, rid_r, rid_s);
          }
          hit = next[hit - 1];
        }
      }
    };
    if constexpr (!EMIT) {
      const ulonglong2 *v2 = reinterpret_cast<const ulonglong2 *>(a.src[1] + c_lo.y) + tid;
      for (uint32_t base = 0; base < s_n; base += BATCH) {
        if (base > 0) {
#pragma unroll
          for (int u = 0; u < kBp3SV; ++u)
            if (base + 2 * (u * TPB + tid) < s_n) xs_cur[u] = ld_stream_v2(v2 + base / 2 + u * TPB);
        }
        if (base + BATCH <= s_n) {  // whole batch in range: no per-tuple bounds checks
#pragma unroll
          for (int u = 0; u < kBp3SV; ++u) {
            probe(xs_cur[u].x);
            probe(xs_cur[u].y);
          }
        } else {
#pragma unroll
          for (int u = 0; u < kBp3SV; ++u) {
            const uint32_t e = base + 2 * (u * TPB + tid);
            if (e < s_n) probe(xs_cur[u].x);
            if (e + 1 < s_n) probe(xs_cur[u].y);
          }
        }
      }
    } else {
      // Materialising probe.  Per batch: (1) probe as above, remembering per probe tuple how
      // many matches it has and its first one; (2) one output range per CTA is reserved with a
      // single atomic and cut up by a scan -- inside a warp in tuple-major order, so that the
      // lanes of one store instruction write neighbouring 16-byte pairs; (3) the pairs are
      // written; only tuples with more than one match walk their chain a second time.
      constexpr int NQ = 2 * kBp3SV;
      const uint32_t lane = tid & 31, warp = tid >> 5;
      const ulonglong2 *v2 = reinterpret_cast<const ulonglong2 *>(a.src[1] + c_lo.y) + tid;
      for (uint32_t base = 0; base < s_n; base += BATCH) {
        if (base > 0) {
#pragma unroll
          for (int u = 0; u < kBp3SV; ++u)
            if (base + 2 * (u * TPB + tid) < s_n) xs_cur[u] = ld_stream_v2(v2 + base / 2 + u * TPB);
        }
        uint32_t m[NQ], first[NQ];
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
          const uint32_t e = base + 2 * ((q >> 1) * TPB + tid) + (q & 1);
          const uint64_t v = (q & 1) ? xs_cur[q >> 1].y : xs_cur[q >> 1].x;
          m[q] = 0;
          first[q] = 0;
          if (e < s_n) {
            if (KS32) {  // rid = low word, key part = high word
              const uint32_t vhi = (uint32_t)(v >> 32), vlo = (uint32_t)v;
              uint32_t hit = bucket[(vhi >> hash_shift32) & hmask];
              while (hit) {
                const uint64_t r = t[hit - 1];
                if ((uint32_t)(r >> 32) == vhi) {
                  const uint32_t rlo = (uint32_t)r;
                  if (m[q] == 0) first[q] = hit;
                  m[q] += 1;
                  item_kpart += vhi;
                  acc_rr += rlo;
                  acc_rs += vlo;
                  acc_mix += mix_triple(((uint64_t)vhi << a.net_bits) | partition, rlo, vlo);
                }
                hit = next[hit - 1];
              }
            } else {
              const uint64_t kpart = v >> a.key_shift;
              uint32_t hit = bucket[(uint32_t)(v >> a.shift_bits) & hmask];
              while (hit) {
                const uint64_t r = t[hit - 1];
                if ((r >> a.key_shift) == kpart) {
                  const uint64_t rid_r = r & rid_mask, rid_s = v & rid_mask;
                  if (m[q] == 0) first[q] = hit;
                  m[q] += 1;
                  item_kpart += kpart;
                  acc_rr += rid_r;
                  acc_rs += rid_s;
                  acc_mix += mix_triple((kpart << a.net_bits) | partition, rid_r, rid_s);
                }
                hit = next[hit - 1];
              }
            }
            item_n += m[q];
          }
        }
        uint32_t off[NQ], wtot = 0;
        uint32_t mmax = 0;
#pragma unroll
        for (int q = 0; q < NQ; ++q) mmax = max(mmax, m[q]);
        if (!__any_sync(0xFFFFFFFFu, mmax > 1)) {  // unique build keys: one vote per tuple slot
#pragma unroll
          for (int q = 0; q < NQ; ++q) {
            const uint32_t b = __ballot_sync(0xFFFFFFFFu, m[q] != 0);
            off[q] = wtot + __popc(b & ((1u << lane) - 1));
            wtot += __popc(b);
          }
        } else {
#pragma unroll
          for (int q = 0; q < NQ; ++q) {
            const uint32_t incl = warp_inclusive_scan(m[q]);
            off[q] = wtot + incl - m[q];
            wtot += __shfl_sync(0xFFFFFFFFu, incl, 31);
          }
        }
        if (lane == 0) s_wtot[warp] = wtot;
        __syncthreads();
        if (warp == 0) {
          const uint32_t x = lane < TPB / 32 ? s_wtot[lane] : 0;
          const uint32_t incl = warp_inclusive_scan(x);
          const uint32_t cta_total = __shfl_sync(0xFFFFFFFFu, incl, 31);
          unsigned long long gb = 0;
          if (lane == 0 && cta_total) gb = atomicAdd(a.pair_cursor, (unsigned long long)cta_total);
          gb = __shfl_sync(0xFFFFFFFFu, gb, 0);
          s_wbase[lane] = gb + incl - x;
        }
        __syncthreads();
        const uint64_t wb = s_wbase[warp];
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
          if (m[q] == 0) continue;
          const uint64_t v = (q & 1) ? xs_cur[q >> 1].y : xs_cur[q >> 1].x;
          const uint64_t rid_s = v & rid_mask;
          uint64_t pos = wb + off[q];
          if (m[q] == 1) {
            if (pos < a.pairs_cap) a.pairs[pos] = make_ulonglong2(t[first[q] - 1] & rid_mask, rid_s);
          } else {
            const uint64_t kpart = v >> a.key_shift;
            for (uint32_t hit = first[q]; hit; hit = next[hit - 1]) {
              const uint64_t r = t[hit - 1];
              if ((r >> a.key_shift) == kpart) {
                if (pos < a.pairs_cap) a.pairs[pos] = make_ulonglong2(r & rid_mask, rid_s);
                ++pos;
              }
            }
          }
        }
      }
    }
    acc_n += item_n;
    acc_key += (item_kpart << a.net_bits) + (uint64_t)item_n * partition;

    fence_proxy_async_smem();
    __syncthreads();
    c_lo = n_lo;
    c_hi = n_hi;
#pragma unroll
    for (int u = 0; u < kBp3SV; ++u) xs_cur[u] = xs_nxt[u];
    v_cur = v_nxt;
    n_lo = f_lo;
    n_hi = f_hi;
    v_nxt = v_far;
    id += stride;
    slot ^= 1;
  }

  uint64_t v[5] = {acc_n, acc_key, acc_rr, acc_rs, acc_mix};
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    uint64_t w = warp_sum(v[k]);
    if ((tid & 31) == 0) s_red[k][tid >> 5] = w;
  }
  __syncthreads();
  if (tid < 5) {
    unsigned long long sum = 0;
    for (int w = 0; w < TPB / 32; ++w) sum += s_red[tid][w];
    if (sum) atomicAdd(&a.result[tid], sum);
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

// ---------------------------------------------------------------------------------
// Arrival flags: the barrier between the network pass and the local pass
// (operators/HashJoin.cpp:119-121, MPI_Barrier after MPI_Win_flush) without a collective.
// After its scatter kernel of relation k a rank writes the join's epoch into slot [k][me] of
// every peer's flag array (release at system scope: the full-line stores of the scatter kernel,
// which completed earlier in the same stream, are visible before the flag is); before its local
// pass of relation k a rank waits until every peer's slot shows the epoch.  One rank per GPU
// only: ranks emulated on one device must not wait for kernels that are queued behind them.
// ---------------------------------------------------------------------------------
struct FlagPeers {
  uint64_t addr[64];  // byte address of every rank's flag array u32[2][64]; 0 = no such peer
};

__global__ void k_signal_peers(FlagPeers peers, uint32_t nranks, uint32_t me, uint32_t which, uint32_t epoch) {
  const uint32_t r = threadIdx.x;
  if (r >= nranks || r == me || peers.addr[r] == 0) return;
  uint32_t *flag = reinterpret_cast<uint32_t *>(peers.addr[r]) + which * 64 + me;
  __threadfence_system();
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(epoch) : "memory");
}

__global__ void k_wait_peers(const uint32_t *flags, uint32_t nranks, uint32_t me, uint32_t which, uint32_t epoch,
                             unsigned long long timeout_ns, int *status) {
  const uint32_t r = threadIdx.x;
  if (r >= nranks || r == me) return;
  const uint32_t *flag = flags + which * 64 + r;
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  while (true) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
    if ((int32_t)(v - epoch) >= 0) break;
    unsigned long long t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (t1 - t0 > timeout_ns) {
      atomicMax(status, (int)kDevPeerTimeout);
      break;
    }
    __nanosleep(200);
  }
}

// vals[i * stride] <- perm(vals[i * stride]): the generator's bijection applied to values that are
// already there (rank -> key for a skewed probe side, oracle orc_fill_zipf)
__global__ void k_apply_perm(uint64_t *vals, uint64_t n, uint32_t stride, PermParams pp) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    vals[i * stride] = perm_apply(vals[i * stride], pp);
}

}  // namespace rhj
