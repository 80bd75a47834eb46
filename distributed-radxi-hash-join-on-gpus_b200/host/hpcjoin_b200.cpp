// hpcjoin_b200.cpp -- see hpcjoin_b200.hpp.  Host-side sequencing only.
#include "hpcjoin_b200.hpp"

#include <unordered_map>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/time.h>
#include <sys/wait.h>
#include <sched.h>
#include <unistd.h>

#include <atomic>
#include <new>

namespace hpcjoin_b200 {

void check(int rc) {
  if (rc != 0) throw Error{rc, rhj_last_error()};
}

static double now_us() {
  timeval tv;
  gettimeofday(&tv, nullptr);
  return tv.tv_sec * 1e6 + tv.tv_usec;
}

// ---- core ---------------------------------------------------------------------------
namespace core {
uint32_t Configuration::RESULT_AGGREGATION_NODE = 0;
bool Configuration::ENABLE_TWO_LEVEL_PARTITIONING = false;  // shipped value, core/Configuration.h:28
uint64_t Configuration::NETWORK_PARTITIONING_FANOUT = 5;    // :30
uint64_t Configuration::LOCAL_PARTITIONING_FANOUT = 5;      // :31
double Configuration::ALLOCATION_FACTOR = 1.1;              // :35
uint32_t Configuration::PAYLOAD_BITS = 27;                  // :37

rhj_config Configuration::toConfig(uint32_t flags) {
  rhj_config c;
  memset(&c, 0, sizeof(c));
  c.network_fanout_bits = (uint32_t)NETWORK_PARTITIONING_FANOUT;
  c.local_fanout_bits = (uint32_t)LOCAL_PARTITIONING_FANOUT;
  c.payload_bits = PAYLOAD_BITS;
  c.enable_two_level = ENABLE_TWO_LEVEL_PARTITIONING ? 1 : 0;
  c.tuple_bytes = sizeof(data::Tuple);
  c.allocation_factor = ALLOCATION_FACTOR;
  c.flags = flags;
  return c;
}
}  // namespace core

// ---- data ---------------------------------------------------------------------------
namespace data {
void SingleRank::allgatherHost(const void *send, void *recv, uint64_t bytes) { memcpy(recv, send, bytes); }

// ---- ProcessComm: ranks = forked processes, exchanges through one MAP_SHARED block ----
namespace {
constexpr uint64_t kGatherSlot = 1ull << 20;  // bytes per rank in the all-gather area
constexpr uint64_t kMailbox = 4ull << 20;     // bytes per ordered pair of ranks in exchange()
struct ControlBlock {
  std::atomic<uint32_t> arrived;
  std::atomic<uint32_t> sense;
  std::atomic<uint32_t> failed;
};
struct Mail {
  std::atomic<uint32_t> full;
  uint32_t pad[15];
};
inline ControlBlock *control(void *shared) { return static_cast<ControlBlock *>(shared); }
inline Mail *mail(void *shared, uint32_t n, uint32_t from, uint32_t to) {
  return reinterpret_cast<Mail *>(static_cast<char *>(shared) + 4096) + (uint64_t)from * n + to;
}
inline char *gather_area(void *shared, uint32_t n) { return static_cast<char *>(shared) + 4096 + (uint64_t)n * n * sizeof(Mail); }
inline char *mailbox(void *shared, uint32_t n, uint32_t from, uint32_t to) {
  return gather_area(shared, n) + (uint64_t)n * kGatherSlot + ((uint64_t)from * n + to) * kMailbox;
}
inline uint64_t shared_bytes(uint32_t n) {
  return 4096 + (uint64_t)n * n * sizeof(Mail) + (uint64_t)n * kGatherSlot + (uint64_t)n * n * kMailbox;
}
inline void spin_until(const std::atomic<uint32_t> &flag, uint32_t value, void *shared) {
  while (flag.load(std::memory_order_acquire) != value) {
    if (control(shared)->failed.load(std::memory_order_relaxed)) throw Error{RHJ_ERR_STATE, "another rank failed"};
    sched_yield();
  }
}
}  // namespace

int ProcessComm::run(uint32_t ranks, int (*body)(ProcessComm &, void *), void *arg) {
  if (ranks < 1 || ranks > 64) throw Error{RHJ_ERR_INVALID, "ranks must be in [1,64]"};
  void *shared = mmap(nullptr, shared_bytes(ranks), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
  if (shared == MAP_FAILED) throw Error{RHJ_ERR_INVALID, "mmap of the shared control block failed"};
  new (control(shared)) ControlBlock();
  for (uint32_t i = 0; i < ranks * ranks; ++i) new (mail(shared, ranks, 0, 0) + i) Mail();
  fflush(nullptr);
  std::vector<pid_t> children;
  uint32_t me = 0;
  for (uint32_t r = 1; r < ranks; ++r) {
    pid_t pid = fork();
    if (pid < 0) throw Error{RHJ_ERR_INVALID, "fork failed"};
    if (pid == 0) {
      me = r;
      children.clear();
      break;
    }
    children.push_back(pid);
  }
  int code = 1;
  try {
    ProcessComm comm(me, ranks, shared);
    code = body(comm, arg);
  } catch (const Error &e) {
    fprintf(stderr, "rank %u: rhj error %d: %s\n", me, e.code, e.message.c_str());
    code = 1;
  }
  if (code) control(shared)->failed.store(1);
  fflush(nullptr);
  if (me != 0) _exit(code);
  for (pid_t pid : children) {
    int st = 0;
    waitpid(pid, &st, 0);
    const int c = WIFEXITED(st) ? WEXITSTATUS(st) : 1;
    if (c > code) code = c;
  }
  munmap(shared, shared_bytes(ranks));
  return code;
}

void ProcessComm::barrier(void *) {  // sense-reversing; the caller has synchronised its GPU work
  ControlBlock *cb = control(shared);
  sense ^= 1;
  if (cb->arrived.fetch_add(1, std::memory_order_acq_rel) + 1 == n) {
    cb->arrived.store(0, std::memory_order_relaxed);
    cb->sense.store(sense, std::memory_order_release);
  } else {
    spin_until(cb->sense, sense, shared);
  }
}

void ProcessComm::allgatherHost(const void *send, void *recv, uint64_t bytes) {
  if (bytes > kGatherSlot) throw Error{RHJ_ERR_INVALID, "all-gather block larger than the shared slot"};
  memcpy(gather_area(shared, n) + (uint64_t)me * kGatherSlot, send, bytes);
  barrier(nullptr);
  for (uint32_t r = 0; r < n; ++r)
    memcpy(static_cast<char *>(recv) + (uint64_t)r * bytes, gather_area(shared, n) + (uint64_t)r * kGatherSlot, bytes);
  barrier(nullptr);
}

void ProcessComm::allgatherDevice(const void *send_dev, void *recv_dev, uint64_t bytes, void *) {
  // u64[2][P] per rank: staged through the host (blocking copies are ordered behind the kernels)
  if (bytes > kGatherSlot) throw Error{RHJ_ERR_INVALID, "all-gather block larger than the shared slot"};
  check(rhj_memcpy_d2h(gather_area(shared, n) + (uint64_t)me * kGatherSlot, send_dev, bytes));
  barrier(nullptr);
  for (uint32_t r = 0; r < n; ++r)
    check(rhj_memcpy_h2d(static_cast<char *>(recv_dev) + (uint64_t)r * bytes, gather_area(shared, n) + (uint64_t)r * kGatherSlot, bytes));
  barrier(nullptr);
}

void ProcessComm::exchange(const void *send, void *recv, uint64_t bytes, uint32_t peer) {
  // pairwise swap in mailbox-sized pieces; every ordered pair has its own mailbox and flag, both
  // sides post before they fetch, so the order in which ranks visit their peers cannot deadlock
  Mail *out = mail(shared, n, me, peer), *in = mail(shared, n, peer, me);
  for (uint64_t off = 0; off < bytes || off == 0; off += kMailbox) {
    const uint64_t len = bytes - off < kMailbox ? bytes - off : kMailbox;
    spin_until(out->full, 0, shared);
    memcpy(mailbox(shared, n, me, peer), static_cast<const char *>(send) + off, len);
    out->full.store(1, std::memory_order_release);
    spin_until(in->full, 1, shared);
    memcpy(static_cast<char *>(recv) + off, mailbox(shared, n, peer, me), len);
    in->full.store(0, std::memory_order_release);
    if (bytes == 0) break;
  }
}

Relation::Relation(uint64_t localSize, uint64_t globalSize) : localSize(localSize), globalSize(globalSize), data(nullptr) {
  void *p = nullptr;
  check(rhj_host_alloc(&p, (localSize ? localSize : 1) * sizeof(Tuple)));
  data = static_cast<Tuple *>(p);
  memset(data, 0, localSize * sizeof(Tuple));
}
Relation::~Relation() { rhj_host_free(data); }

// keys start+i, shuffled, then rids assigned in final positions
void Relation::fillUniqueValues(uint64_t startKeyValue, uint64_t startRidValue) {
  for (uint64_t i = 0; i < localSize; ++i) data[i].key = startKeyValue + i;
  randomOrder();
  for (uint64_t i = 0; i < localSize; ++i) data[i].rid = startRidValue + i;
}
void Relation::fillModuloValues(uint64_t startKeyValue, uint64_t startRidValue, uint64_t innerRelationSize) {
  for (uint64_t i = 0; i < localSize; ++i) data[i].key = startKeyValue + (i % innerRelationSize);
  randomOrder();
  for (uint64_t i = 0; i < localSize; ++i) data[i].rid = startRidValue + i;
}
// Fisher-Yates from the top on the key field, j drawn as (rand()/(RAND_MAX+1)) * i in double
// precision: the draw sequence (and so the permutation) equals the reference's for a given srand().
void Relation::randomOrder() {
  if (localSize < 2) return;
  for (uint64_t i = localSize - 1; i > 0; --i) {
    const double u = (double)rand() / ((double)RAND_MAX + 1.0);
    const uint64_t j = (uint64_t)(u * (double)i);
    const uint64_t k = data[i].key;
    data[i].key = data[j].key;
    data[j].key = k;
  }
}
// rank me and rank i swap their sections numbered (me+i) % n, then keys are reshuffled
void Relation::distribute(Comm &comm) {
  const uint32_t n = comm.size(), me = comm.rank();
  if (n > 1) {
    const uint64_t sec = localSize / n, last = localSize - (uint64_t)(n - 1) * sec;
    std::vector<Tuple> incoming(last ? last : 1);
    for (uint32_t i = 0; i < n; ++i) {
      if (i == me) continue;
      const uint32_t section = (me + i) % n;
      const uint64_t start = (uint64_t)section * sec, size = section == n - 1 ? last : sec;
      comm.exchange(data + start, incoming.data(), size * sizeof(Tuple), i);
      memcpy(data + start, incoming.data(), size * sizeof(Tuple));
    }
  }
  randomOrder();
}
}  // namespace data

// ---- memory -------------------------------------------------------------------------
namespace memory {
rhj_ctx *Pool::context = nullptr;
void Pool::allocate(uint64_t size) { check(rhj_pool_allocate(context, size)); }
void *Pool::getMemory(uint64_t size) {
  void *p = nullptr;
  check(rhj_pool_get_memory(context, size, &p));
  return p;
}
void Pool::reset() { check(rhj_pool_reset(context)); }
void Pool::freeAll() { check(rhj_pool_free_all(context)); }
}  // namespace memory

// ---- performance --------------------------------------------------------------------
namespace performance {
rhj_timings Measurements::last = {};
double Measurements::wallJoinUs = 0;
std::string Measurements::directory;
uint32_t Measurements::node = 0;
double Measurements::t0 = 0;

void Measurements::init(uint32_t nodeId, uint32_t numberOfNodes, const std::string &tag) {
  node = nodeId;
  char buf[512];
  snprintf(buf, sizeof(buf), "%s-%u", tag.c_str(), numberOfNodes);
  directory = buf;
  mkdir(directory.c_str(), 0755);
}
void Measurements::startJoin() { t0 = now_us(); }
void Measurements::stopJoin(rhj_ctx *ctx) {
  wallJoinUs = now_us() - t0;
  rhj_get_timings(ctx, &last);
}
// same keys and "key\tvalue\tunit" format as storePhaseData / storeSpecialData / store*Data
void Measurements::storeAllMeasurements() {
  if (directory.empty()) return;
  char path[600];
  snprintf(path, sizeof(path), "%s/%u.perf", directory.c_str(), node);
  FILE *f = fopen(path, "w");
  if (!f) return;
  fprintf(f, "JTOTAL\t%lu\tus\n", (unsigned long)last.jtotal_us);
  fprintf(f, "JHIST\t%lu\tus\n", (unsigned long)last.jhist_us);
  fprintf(f, "JMPI\t%lu\tus\n", (unsigned long)last.jmpi_us);
  fprintf(f, "JPROC\t%lu\tus\n", (unsigned long)last.jproc_us);
  fprintf(f, "LPTASKTIME\t%lu\tus\n", (unsigned long)last.lp_us);
  fprintf(f, "BPTASKTIME\t%lu\tus\n", (unsigned long)last.bp_us);
  // special keys (performance/Measurements.cpp:176-178): the windows are carved when the context is planned and
  // the local pass needs no preparation step here, so only the wait for the network phase is non-zero
  fprintf(f, "SWINALLOC\t%lu\tus\n", 0ul);
  fprintf(f, "SNETCOMPL\t%lu\tus\n", (unsigned long)last.snetcompl_us);
  fprintf(f, "SLOCPREP\t%lu\tus\n", 0ul);
  fprintf(f, "JWALL\t%lu\tus\n", (unsigned long)wallJoinUs);
  fclose(f);
}
void Measurements::printMeasurements(uint32_t numberOfNodes, uint32_t, const std::vector<uint64_t> &tuplesPerNode) {
  printf("[RESULTS] Tuples:\t");
  uint64_t total = 0;
  for (uint32_t n = 0; n < numberOfNodes; ++n) {
    printf("%lu\t", (unsigned long)tuplesPerNode[n]);
    total += tuplesPerNode[n];
  }
  printf("\n[RESULTS] Join:\t%.3f\n", last.jtotal_us / 1000.0);
  printf("[RESULTS] Summary:\t%lu\t%.3f\t%.3f\t%.3f\t%.3f\n", (unsigned long)total, last.jtotal_us / 1000.0,
         last.jhist_us / 1000.0, last.jmpi_us / 1000.0, last.jproc_us / 1000.0);
}
// the ten values a rank sends to the aggregator (performance/Measurements.cpp:548-565, serializeResults):
// matches, total, histogram, network, local phase, window allocation, network completion wait, local
// preparation, local-partitioning task sum, build/probe task sum (times in microseconds)
std::vector<uint64_t> Measurements::serializeResults(uint64_t matches) {
  return {matches, (uint64_t)last.jtotal_us, (uint64_t)last.jhist_us, (uint64_t)last.jmpi_us, (uint64_t)last.jproc_us,
          0, (uint64_t)last.snetcompl_us, 0, (uint64_t)last.lp_us, (uint64_t)last.bp_us};
}
}  // namespace performance

// ---- tasks --------------------------------------------------------------------------
namespace tasks {
HistogramComputation::HistogramComputation(rhj_ctx *ctx, data::Comm *comm, const void *dInner, uint64_t nInner,
                                           const void *dOuter, uint64_t nOuter, void *stream)
    : ctx(ctx), comm(comm), dInner(dInner), dOuter(dOuter), nInner(nInner), nOuter(nOuter), stream(stream) {}

void HistogramComputation::execute() {
  check(rhj_histogram(ctx, dInner, nInner, dOuter, nOuter, stream));  // LocalHistogram x2
  const uint64_t P = core::Configuration::NETWORK_PARTITIONING_COUNT();
  if (comm->size() > 1) {
    // one all-gather of u64[2][P] replaces MPI_Allreduce (GlobalHistogram.cpp:37) + MPI_Exscan (OffsetMap.cpp:78)
    // one gather buffer per context for all joins on it: this task object is built anew by every join(),
    // and a bump allocation per join would exhaust the pool after a few of them
    static std::unordered_map<rhj_ctx *, void *> buffers;
    void *&slot = buffers[ctx];
    if (!slot) slot = memory::Pool::getMemory((uint64_t)comm->size() * 2 * P * 8);
    gathered = slot;
    comm->allgatherDevice(rhj_local_histogram_dev(ctx), gathered, 2 * P * 8, stream);
    check(rhj_compute_offsets(ctx, static_cast<const uint64_t *>(gathered), stream));
  } else {
    check(rhj_compute_offsets(ctx, nullptr, stream));
  }
}
std::vector<uint32_t> HistogramComputation::getAssignment() {
  std::vector<uint32_t> v(core::Configuration::NETWORK_PARTITIONING_COUNT());
  check(rhj_get_assignment(ctx, v.data(), stream));
  return v;
}
#define RHJ_GETTER(NAME, FN, WHICH)                                          \
  std::vector<uint64_t> HistogramComputation::NAME() {                       \
    std::vector<uint64_t> v(core::Configuration::NETWORK_PARTITIONING_COUNT()); \
    check(FN(ctx, WHICH, v.data(), stream));                                 \
    return v;                                                                \
  }
RHJ_GETTER(getInnerRelationLocalHistogram, rhj_get_local_histogram, 0)
RHJ_GETTER(getOuterRelationLocalHistogram, rhj_get_local_histogram, 1)
RHJ_GETTER(getInnerRelationGlobalHistogram, rhj_get_global_histogram, 0)
RHJ_GETTER(getOuterRelationGlobalHistogram, rhj_get_global_histogram, 1)
RHJ_GETTER(getInnerRelationBaseOffsets, rhj_get_base_offsets, 0)
RHJ_GETTER(getOuterRelationBaseOffsets, rhj_get_base_offsets, 1)
RHJ_GETTER(getInnerRelationWriteOffsets, rhj_get_write_offsets, 0)
RHJ_GETTER(getOuterRelationWriteOffsets, rhj_get_write_offsets, 1)
#undef RHJ_GETTER

void LocalPartitioning::execute() {
  check(rhj_local_partition(ctx, stream));
  operators::HashJoin::TASK_QUEUE.push(new BuildProbe(ctx, stream, pairs, capacity));
}
void BuildProbe::execute() {
  if (pairs) check(rhj_build_probe_pairs(ctx, pairs, capacity, stream));
  else check(rhj_build_probe(ctx, stream));
  rhj_result r;
  check(rhj_get_result(ctx, &r, stream));
  operators::HashJoin::RESULT_COUNTER += r.matches;
  operators::HashJoin::RESULT_CHECKSUM[0] += r.sum_key;
  operators::HashJoin::RESULT_CHECKSUM[1] += r.sum_rid_r;
  operators::HashJoin::RESULT_CHECKSUM[2] += r.sum_rid_s;
  operators::HashJoin::RESULT_CHECKSUM[3] += r.sum_mix;
}
}  // namespace tasks

// ---- operators ----------------------------------------------------------------------
namespace operators {
uint64_t HashJoin::RESULT_COUNTER = 0;
uint64_t HashJoin::RESULT_CHECKSUM[4] = {0, 0, 0, 0};
uint64_t HashJoin::RESULT_PAIR_CAPACITY = 0;
std::queue<tasks::Task *> HashJoin::TASK_QUEUE;

HashJoin::HashJoin(uint32_t numberOfNodes, uint32_t nodeId, data::Relation *innerRelation,
                   data::Relation *outerRelation, data::Comm *comm_, int device)
    : numberOfNodes(numberOfNodes), nodeId(nodeId), innerRelation(innerRelation), outerRelation(outerRelation),
      comm(comm_ ? comm_ : &single), ctx(nullptr), dInner(nullptr), dOuter(nullptr), dPairs(nullptr), pairsThisJoin(0),
      stream(nullptr) {
  rhj_config cfg = core::Configuration::toConfig();
  check(rhj_init(&ctx, (int)nodeId, (int)numberOfNodes, device, &cfg));
  memory::Pool::bind(ctx);
  const uint64_t ni = innerRelation->getLocalSize(), no = outerRelation->getLocalSize();
  // main.cpp:86-88 sizes the pool for the relations; here it also holds windows and intermediates
  const uint64_t inputs = ((ni + no) * sizeof(data::Tuple)) + 4096;
  // window head-room: a rank can receive up to allocation_factor x the mean (skew beyond that is an error)
  const uint64_t wi = (uint64_t)(core::Configuration::ALLOCATION_FACTOR * (double)innerRelation->getGlobalSize() / numberOfNodes) + 1024;
  const uint64_t wo = (uint64_t)(core::Configuration::ALLOCATION_FACTOR * (double)outerRelation->getGlobalSize() / numberOfNodes) + 1024;
  memory::Pool::allocate(inputs + rhj_workspace_bytes(&cfg, (int)numberOfNodes, ni, no, wi, wo) +
                         (uint64_t)numberOfNodes * 4 * core::Configuration::NETWORK_PARTITIONING_COUNT() * 8 + 4096 +
                         RESULT_PAIR_CAPACITY * sizeof(rhj_pair) + 4096);
  dInner = memory::Pool::getMemory(ni * sizeof(data::Tuple));
  dOuter = memory::Pool::getMemory(no * sizeof(data::Tuple));
  if (RESULT_PAIR_CAPACITY) dPairs = static_cast<rhj_pair *>(memory::Pool::getMemory(RESULT_PAIR_CAPACITY * sizeof(rhj_pair)));
  check(rhj_memcpy_h2d(dInner, innerRelation->getData(), ni * sizeof(data::Tuple)));
  check(rhj_memcpy_h2d(dOuter, outerRelation->getData(), no * sizeof(data::Tuple)));
  check(rhj_plan(ctx, ni, no, wi, wo));
  if (numberOfNodes > 1) {  // replaces MPI_Win_create (data/Window.cpp:35-45)
    rhj_window_handle mine;
    check(rhj_window_export(ctx, &mine));
    std::vector<rhj_window_handle> all(numberOfNodes);
    comm->allgatherHost(&mine, all.data(), sizeof(mine));
    for (uint32_t r = 0; r < numberOfNodes; ++r)
      if (r != nodeId) check(rhj_window_import(ctx, (int)r, &all[r]));
  }
}

HashJoin::~HashJoin() {
  memory::Pool::bind(nullptr);
  rhj_destroy(ctx);
}

// operators/HashJoin.cpp:45-220: histogram -> (windows exist already) -> network partitioning ->
// barrier -> task queue with the local phase.
void HashJoin::join() {
  comm->barrier(stream);
  performance::Measurements::startJoin();

  tasks::HistogramComputation histogramComputation(ctx, comm, dInner, innerRelation->getLocalSize(), dOuter,
                                                   outerRelation->getLocalSize(), stream);
  histogramComputation.execute();

  tasks::NetworkPartitioning networkPartitioning(ctx, stream);
  networkPartitioning.execute();

  if (numberOfNodes > 1) {  // MPI_Win_flush_local_all (data/Window.cpp:193-201): my stores have left
    check(rhj_exchange_sync(ctx, 0));
    check(rhj_exchange_sync(ctx, 1));
  }
  comm->barrier(stream);  // operators/HashJoin.cpp:119-121: every put has landed

  // two-level off: the pass is a no-op and only queues BuildProbe (HashJoin.cpp:151-156)
  const uint64_t before = RESULT_COUNTER;
  TASK_QUEUE.push(new tasks::LocalPartitioning(ctx, stream, dPairs, dPairs ? RESULT_PAIR_CAPACITY : 0));

  while (!TASK_QUEUE.empty()) {  // operators/HashJoin.cpp:187-204
    tasks::Task *task = TASK_QUEUE.front();
    TASK_QUEUE.pop();
    task->execute();
    delete task;
  }
  performance::Measurements::stopJoin(ctx);
  pairsThisJoin = RESULT_COUNTER - before;
}

std::vector<rhj_pair> HashJoin::getResultPairs() {
  const uint64_t n = dPairs ? (pairsThisJoin < RESULT_PAIR_CAPACITY ? pairsThisJoin : RESULT_PAIR_CAPACITY) : 0;
  std::vector<rhj_pair> out(n);
  if (n) check(rhj_memcpy_d2h(out.data(), dPairs, n * sizeof(rhj_pair)));
  return out;
}
}  // namespace operators

}  // namespace hpcjoin_b200
