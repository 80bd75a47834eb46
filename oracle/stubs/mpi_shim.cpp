/*
 * mpi_shim.cpp -- fork()+shared-memory backing for oracle/stubs/mpi.h.
 * TEST INFRASTRUCTURE: lets the unmodified reference run as N single-threaded
 * "ranks" (processes, because the reference keeps Pool / Measurements /
 * RESULT_COUNTER / TASK_QUEUE in statics) on a box without MPI.
 *
 * One MAP_SHARED|MAP_ANONYMOUS region is mapped BEFORE fork, so every pointer
 * into it is valid in every rank: MPI_Alloc_mem carves from it, MPI_Put is a
 * memcpy to the target's registered window base.
 */
#include "mpi.h"

#include <atomic>
#include <sched.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <unistd.h>

namespace {

constexpr int kMaxRanks = 64;
constexpr int kMaxWins = 64;
constexpr size_t kScratchPerRank = 1 << 16;
constexpr size_t kChunk = 1 << 18;
constexpr size_t kResultPerRank = 4096;

struct Channel {
  std::atomic<uint64_t> produced;
  std::atomic<uint64_t> consumed;
};

struct Control {
  int nranks;
  std::atomic<int> bar_count;
  std::atomic<int> bar_sense;
  void *win_base[kMaxWins][kMaxRanks];
  size_t heap_bytes_per_rank;
  std::atomic<size_t> heap_used[kMaxRanks];
  Channel chan[kMaxRanks][kMaxRanks];
};

Control *g_ctl = nullptr;
char *g_scratch = nullptr; /* nranks * kScratchPerRank */
char *g_chunks = nullptr;  /* nranks*nranks * kChunk */
char *g_results = nullptr; /* nranks * kResultPerRank */
char *g_heap = nullptr;    /* nranks * heap_bytes_per_rank */
int g_rank = 0;
int g_local_sense = 0;
int g_next_win = 0;

inline void relax(unsigned &spins) {
  if (++spins > 2000) {
    sched_yield();
  } else {
#if defined(__x86_64__)
    __builtin_ia32_pause();
#endif
  }
}

void barrier() {
  Control *c = g_ctl;
  if (c->nranks == 1) return;
  g_local_sense ^= 1;
  if (c->bar_count.fetch_add(1, std::memory_order_acq_rel) == c->nranks - 1) {
    c->bar_count.store(0, std::memory_order_relaxed);
    c->bar_sense.store(g_local_sense, std::memory_order_release);
  } else {
    unsigned spins = 0;
    while (c->bar_sense.load(std::memory_order_acquire) != g_local_sense) relax(spins);
  }
}

size_t type_size(MPI_Datatype t) { return t == MPI_UINT64_T ? 8 : 1; }

}  // namespace

extern "C" {

int shim_setup(int nranks, size_t shared_bytes) {
  if (nranks < 1 || nranks > kMaxRanks) return -1;
  size_t ctl_bytes = (sizeof(Control) + 4095) & ~size_t(4095);
  size_t scratch = size_t(nranks) * kScratchPerRank;
  size_t chunks = size_t(nranks) * nranks * kChunk;
  size_t results = size_t(nranks) * kResultPerRank;
  size_t heap_per = ((shared_bytes / nranks) + 4095) & ~size_t(4095);
  size_t total = ctl_bytes + scratch + chunks + results + heap_per * nranks;
  void *m = mmap(nullptr, total, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
  if (m == MAP_FAILED) {
    perror("shim mmap");
    return -1;
  }
  char *p = static_cast<char *>(m);
  g_ctl = new (p) Control();
  g_ctl->nranks = nranks;
  g_ctl->heap_bytes_per_rank = heap_per;
  p += ctl_bytes;
  g_scratch = p; p += scratch;
  g_chunks = p; p += chunks;
  g_results = p; p += results;
  g_heap = p;
  return 0;
}

void shim_set_rank(int rank) { g_rank = rank; }
void *shim_result_area(void) { return g_results; }

int MPI_Init(int *, char ***) { return MPI_SUCCESS; }
int MPI_Finalize(void) { return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm, int *size) { *size = g_ctl->nranks; return MPI_SUCCESS; }
int MPI_Comm_rank(MPI_Comm, int *rank) { *rank = g_rank; return MPI_SUCCESS; }
int MPI_Barrier(MPI_Comm) { barrier(); return MPI_SUCCESS; }

int MPI_Allreduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op, MPI_Comm) {
  size_t bytes = size_t(count) * type_size(type);
  if (type != MPI_UINT64_T || bytes > kScratchPerRank) abort();
  memcpy(g_scratch + size_t(g_rank) * kScratchPerRank, sendbuf, bytes);
  barrier();
  uint64_t *out = static_cast<uint64_t *>(recvbuf);
  for (int i = 0; i < count; ++i) {
    uint64_t s = 0;
    for (int r = 0; r < g_ctl->nranks; ++r)
      s += reinterpret_cast<uint64_t *>(g_scratch + size_t(r) * kScratchPerRank)[i];
    out[i] = s;
  }
  barrier();
  return MPI_SUCCESS;
}

/* rank 0's recvbuf is left untouched, as MPI specifies (the reference relies on calloc) */
int MPI_Exscan(const void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op, MPI_Comm) {
  size_t bytes = size_t(count) * type_size(type);
  if (type != MPI_UINT64_T || bytes > kScratchPerRank) abort();
  memcpy(g_scratch + size_t(g_rank) * kScratchPerRank, sendbuf, bytes);
  barrier();
  if (g_rank > 0) {
    uint64_t *out = static_cast<uint64_t *>(recvbuf);
    for (int i = 0; i < count; ++i) {
      uint64_t s = 0;
      for (int r = 0; r < g_rank; ++r)
        s += reinterpret_cast<uint64_t *>(g_scratch + size_t(r) * kScratchPerRank)[i];
      out[i] = s;
    }
  }
  barrier();
  return MPI_SUCCESS;
}

int MPI_Bcast(void *buf, int count, MPI_Datatype type, int root, MPI_Comm) {
  size_t bytes = size_t(count) * type_size(type);
  if (bytes > kScratchPerRank) abort();
  if (g_rank == root) memcpy(g_scratch + size_t(root) * kScratchPerRank, buf, bytes);
  barrier();
  if (g_rank != root) memcpy(buf, g_scratch + size_t(root) * kScratchPerRank, bytes);
  barrier();
  return MPI_SUCCESS;
}

int MPI_Send(const void *buf, int count, MPI_Datatype type, int dest, int, MPI_Comm) {
  size_t bytes = size_t(count) * type_size(type);
  Channel &ch = g_ctl->chan[g_rank][dest];
  char *slot = g_chunks + (size_t(g_rank) * g_ctl->nranks + dest) * kChunk;
  const char *src = static_cast<const char *>(buf);
  for (size_t off = 0; off < bytes || (bytes == 0 && off == 0); off += kChunk) {
    size_t n = bytes - off < kChunk ? bytes - off : kChunk;
    unsigned spins = 0;
    while (ch.produced.load(std::memory_order_acquire) != ch.consumed.load(std::memory_order_acquire)) relax(spins);
    memcpy(slot, src + off, n);
    ch.produced.fetch_add(1, std::memory_order_release);
    if (bytes == 0) break;
  }
  return MPI_SUCCESS;
}

int MPI_Recv(void *buf, int count, MPI_Datatype type, int source, int, MPI_Comm, MPI_Status *) {
  size_t bytes = size_t(count) * type_size(type);
  Channel &ch = g_ctl->chan[source][g_rank];
  char *slot = g_chunks + (size_t(source) * g_ctl->nranks + g_rank) * kChunk;
  char *dst = static_cast<char *>(buf);
  for (size_t off = 0; off < bytes || (bytes == 0 && off == 0); off += kChunk) {
    size_t n = bytes - off < kChunk ? bytes - off : kChunk;
    unsigned spins = 0;
    while (ch.produced.load(std::memory_order_acquire) == ch.consumed.load(std::memory_order_acquire)) relax(spins);
    memcpy(dst + off, slot, n);
    ch.consumed.fetch_add(1, std::memory_order_release);
    if (bytes == 0) break;
  }
  return MPI_SUCCESS;
}

int MPI_Alloc_mem(MPI_Aint size, MPI_Info, void *baseptr) {
  size_t need = (size_t(size) + 63) & ~size_t(63);
  size_t off = g_ctl->heap_used[g_rank].fetch_add(need);
  if (off + need > g_ctl->heap_bytes_per_rank) {
    fprintf(stderr, "mpi shim: rank %d shared heap exhausted (%zu + %zu > %zu)\n", g_rank, off, need,
            g_ctl->heap_bytes_per_rank);
    abort();
  }
  *static_cast<void **>(baseptr) = g_heap + size_t(g_rank) * g_ctl->heap_bytes_per_rank + off;
  return MPI_SUCCESS;
}

int MPI_Free_mem(void *) { return MPI_SUCCESS; } /* bump heap: released at exit */

int MPI_Win_create(void *base, MPI_Aint, int, MPI_Info, MPI_Comm, MPI_Win *win) {
  int id = g_next_win++; /* collective: every rank creates windows in the same order */
  if (id >= kMaxWins) abort();
  g_ctl->win_base[id][g_rank] = base;
  barrier();
  *win = id;
  return MPI_SUCCESS;
}

int MPI_Win_free(MPI_Win *) { barrier(); return MPI_SUCCESS; }
int MPI_Win_lock_all(int, MPI_Win) { return MPI_SUCCESS; }
int MPI_Win_unlock_all(MPI_Win) { return MPI_SUCCESS; }

int MPI_Put(const void *origin, int origin_count, MPI_Datatype origin_type, int target_rank,
            MPI_Aint target_disp, int, MPI_Datatype, MPI_Win win) {
  char *dst = static_cast<char *>(g_ctl->win_base[win][target_rank]) + target_disp;
  memcpy(dst, origin, size_t(origin_count) * type_size(origin_type));
  return MPI_SUCCESS;
}

int MPI_Win_flush_local(int, MPI_Win) { return MPI_SUCCESS; }
int MPI_Win_flush_local_all(MPI_Win) {
  std::atomic_thread_fence(std::memory_order_seq_cst);
  return MPI_SUCCESS;
}

}  // extern "C"
