// rhj_device.cuh -- device-side building blocks shared by the join kernels (sm_100a).
//
// Nothing here is a port of the reference's CUDA (operators/gpu/*.cu is Pascal-era,
// mostly uncompiled); these are the primitives the four stages
// are assembled from: streaming 8/16-byte loads, warp/block scans, the packed
// tuple format of tasks/NetworkPartitioning.cpp:128-129 and the checksum term.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace rhj {

constexpr int kWarp = 32;

// ---- streaming memory access -------------------------------------------------
// Inputs of every pass are read exactly once: bypass L1 allocation so the
// shared-memory carve-out is not fighting a useless L1 working set.
__device__ __forceinline__ ulonglong2 ld_stream_v2(const ulonglong2 *p) {
  ulonglong2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];" : "=l"(r.x), "=l"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ uint64_t ld_stream_u64(const uint64_t *p) {
  uint64_t r;
  asm volatile("ld.global.nc.L1::no_allocate.u64 %0, [%1];" : "=l"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ ulonglong2 ld_stream(const ulonglong2 *p) { return ld_stream_v2(p); }
__device__ __forceinline__ uint64_t ld_stream(const uint64_t *p) { return ld_stream_u64(p); }
// Scatter output is write-once, never re-read by this kernel; may target a peer GPU.
__device__ __forceinline__ void st_stream_u64(uint64_t *p, uint64_t v) {
  asm volatile("st.global.L1::no_allocate.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ---- tuple formats --------------------------------------------------------------
// data/Tuple.h:15-22 {u64 key; u64 rid}  (tuple_bytes == 16), or the packed 8-byte
// input of BASELINE config 5: key in bits [32,64), rid in bits [0,32).
template <int TUPLE_BYTES>
struct InputTuple;

template <>
struct InputTuple<16> {
  static __device__ __forceinline__ void load(const void *base, uint64_t i, uint64_t &key, uint64_t &rid) {
    ulonglong2 t = ld_stream_v2(reinterpret_cast<const ulonglong2 *>(base) + i);
    key = t.x;
    rid = t.y;
  }
  static __device__ __forceinline__ uint64_t load_key(const void *base, uint64_t i) {
    return ld_stream_u64(reinterpret_cast<const uint64_t *>(base) + 2 * i);
  }
};

template <>
struct InputTuple<8> {
  static __device__ __forceinline__ void load(const void *base, uint64_t i, uint64_t &key, uint64_t &rid) {
    uint64_t v = ld_stream_u64(reinterpret_cast<const uint64_t *>(base) + i);
    key = v >> 32;
    rid = v & 0xFFFFFFFFull;
  }
  static __device__ __forceinline__ uint64_t load_key(const void *base, uint64_t i) {
    return ld_stream_u64(reinterpret_cast<const uint64_t *>(base) + i) >> 32;
  }
};

// tasks/NetworkPartitioning.cpp:128-129.  `+`, not `|`: an oversized rid carries
// into the key field exactly as it does in the reference.
__device__ __forceinline__ uint64_t compress_tuple(uint64_t key, uint64_t rid, uint32_t net_bits, uint32_t key_shift) {
  return rid + ((key >> net_bits) << key_shift);
}

// Checksum term of one matching pair; bit-identical to oracle/hpcjoin_oracle.c orc_mix().
// Two 32x32->64 multiply-adds (IMAD.WIDE): the build/probe kernel is issue-bound, so the
// term is kept to a handful of instructions while still tying key, rid_R and rid_S together.
__device__ __forceinline__ uint64_t mix_triple(uint64_t key, uint64_t rid_r, uint64_t rid_s) {
  const uint32_t k32 = (uint32_t)key ^ (uint32_t)(key >> 32);
  const uint32_t rr = (uint32_t)rid_r ^ (uint32_t)(rid_r >> 32);
  const uint32_t rs = (uint32_t)rid_s ^ (uint32_t)(rid_s >> 32);
  const uint64_t a = (uint64_t)(rr ^ 0x9E3779B9u) * (uint64_t)(rs ^ 0x85EBCA6Bu);
  return a + (uint64_t)(k32 ^ 0xC2B2AE35u) * (uint64_t)(uint32_t)(rr + rs + 0x27D4EB2Fu);
}

// ---- scans -------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ T warp_inclusive_scan(T v) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    T o = __shfl_up_sync(0xFFFFFFFFu, v, d);
    if (lane >= d) v += o;
  }
  return v;
}

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, d);
  return v;
}

// Exclusive scan of one value per thread across the block (blockDim.x a multiple of
// 32, <= 1024).  `tmp` has 33 entries.  Returns the exclusive prefix; *total gets
// the block sum.  Contains two __syncthreads(); safe to call back to back.
template <typename T>
__device__ __forceinline__ T block_exclusive_scan(T v, T *tmp, T *total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  T incl = warp_inclusive_scan(v);
  if (lane == 31) tmp[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    T w = lane < nwarps ? tmp[lane] : T(0);
    T wi = warp_inclusive_scan(w);
    tmp[lane] = wi - w;
    if (lane == 31) tmp[32] = wi;
  }
  __syncthreads();
  T excl = tmp[warp] + incl - v;
  *total = tmp[32];
  __syncthreads();
  return excl;
}

__device__ __forceinline__ uint64_t round_up8(uint64_t v) { return (v + 7) & ~uint64_t(7); }

// Largest i in [0, n) with a[i] <= x, for an ascending array with a[0] <= x.
template <typename T>
__device__ __forceinline__ uint32_t upper_bound_idx(const T *a, uint32_t n, T x) {
  uint32_t lo = 0, hi = n;
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (a[mid] <= x) lo = mid; else hi = mid;
  }
  return lo;
}

// Same result as upper_bound_idx, computed by a whole (converged) warp with a 32-ary
// search: two dependent loads for 256 entries instead of eight.
template <typename T>
__device__ __forceinline__ uint32_t warp_upper_bound(const T *a, uint32_t n, T x) {
  const uint32_t lane = threadIdx.x & 31;
  uint32_t lo = 0, len = n;
  while (len > 1) {
    const uint32_t step = (len + 31) >> 5;
    const uint32_t idx = lo + lane * step;
    const bool le = (lane * step < len) && (a[idx] <= x);
    const uint32_t k = __popc(__ballot_sync(0xFFFFFFFFu, le)) - 1;  // lane 0 is always true
    lo += k * step;
    const uint32_t rest = len - k * step;
    len = rest < step ? rest : step;
  }
  return lo;
}

}  // namespace rhj

// ---- mbarrier + bulk async copy (TMA, non-tensor form) ---------------------------------
// SASS: UBLKCP (cp.async.bulk), SYNCS.* (mbarrier).  A bulk copy moves a contiguous,
// 16-byte aligned range global -> shared without passing through registers and signals
// completion on an mbarrier, so the load latency is off the consumers' critical path.
namespace rhj {

// Shared-memory loads through an explicit 32-bit shared-window address.  `pin_u32` makes the
// address opaque to the compiler, which otherwise recomputes the window base (S2UR CgaCtaId,
// ULEA, ...) in front of every access of a hot loop instead of spending a register on it.
__device__ __forceinline__ uint32_t pin_u32(uint32_t v) {
  uint32_t r;
  asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(v));
  return r;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint32_t lds_u16(uint32_t addr) {
  uint16_t v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint64_t lds_u64(uint32_t addr) {
  uint64_t v;
  asm volatile("ld.shared.u64 %0, [%1];" : "=l"(v) : "r"(addr));
  return v;
}

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// one arrival that also announces `bytes` of asynchronous transfers to come
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
// global -> shared bulk copy; dst/src 16-byte aligned, bytes a multiple of 16
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// order this thread's earlier generic-proxy shared accesses before later async-proxy ones
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

}  // namespace rhj

namespace rhj {
// Ask L2 to fetch a contiguous range ahead of time (no registers, no shared memory, no
// completion to wait for).  Lets the real loads, issued a tile or two later, hit in the
// 126 MB L2 instead of paying the loaded HBM latency.  addr 16-byte aligned, bytes % 16 == 0.
__device__ __forceinline__ void prefetch_l2_bulk(const void *addr, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(addr), "r"(bytes) : "memory");
}
// [p, p+bytes) shrunk to 16-byte boundaries, then prefetched
__device__ __forceinline__ void prefetch_l2_range(const void *p, uint64_t bytes) {
  const uintptr_t a0 = (reinterpret_cast<uintptr_t>(p) + 15) & ~uintptr_t(15);
  const uintptr_t a1 = (reinterpret_cast<uintptr_t>(p) + bytes) & ~uintptr_t(15);
  if (a1 > a0) prefetch_l2_bulk(reinterpret_cast<const void *>(a0), (uint32_t)(a1 - a0));
}
}  // namespace rhj
