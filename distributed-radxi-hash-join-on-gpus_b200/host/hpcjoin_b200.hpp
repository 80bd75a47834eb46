// hpcjoin_b200.hpp -- C++ host side above the C ABI (include/rhj.h), mirroring the
// reference's operator interface so that code written against hpcjoin::operators::HashJoin
// ports by changing a namespace.  Names and argument meaning follow:
//   core/Configuration.h:19-38      -> hpcjoin_b200::core::Configuration (runtime knobs)
//   data/Tuple.h, CompressedTuple.h -> hpcjoin_b200::data::Tuple / CompressedTuple
//   data/Relation.h:18-60           -> hpcjoin_b200::data::Relation (same generators)
//   memory/Pool.h:19-27             -> hpcjoin_b200::memory::Pool (device arena)
//   tasks/Task.h:10-30 + the four tasks
//   operators/HashJoin.h:23-43      -> hpcjoin_b200::operators::HashJoin
//   performance/Measurements.h      -> hpcjoin_b200::performance::Measurements (phase keys)
// All per-tuple work happens in librhj_b200.so; nothing here computes on the CPU except the
// reference's own data generator (rand() Fisher-Yates), which is workload set-up.
#pragma once

#include <stdint.h>

#include <queue>
#include <string>
#include <vector>

#include "../../include/rhj.h"

enum task_type_t { TASK_HISTOGRAM, TASK_NET_PARTITION, TASK_PARTITION, TASK_BUILD_PROBE };

namespace hpcjoin_b200 {

struct Error {
  int code;
  std::string message;
};
void check(int rc);  // throws Error with rhj_last_error()

namespace core {
// core/Configuration.h:19-38; static like the reference, but assignable before the first join.
struct Configuration {
  static uint32_t RESULT_AGGREGATION_NODE;
  static bool ENABLE_TWO_LEVEL_PARTITIONING;
  static uint64_t NETWORK_PARTITIONING_FANOUT;
  static uint64_t LOCAL_PARTITIONING_FANOUT;
  static double ALLOCATION_FACTOR;
  static uint32_t PAYLOAD_BITS;
  static uint64_t NETWORK_PARTITIONING_COUNT() { return 1ull << NETWORK_PARTITIONING_FANOUT; }
  static uint64_t LOCAL_PARTITIONING_COUNT() { return 1ull << LOCAL_PARTITIONING_FANOUT; }
  static rhj_config toConfig(uint32_t flags = 0);
};
}  // namespace core

namespace data {
struct Tuple {  // data/Tuple.h:15-22
  uint64_t key;
  uint64_t rid;
};
struct CompressedTuple {  // data/CompressedTuple.h:14-20
  unsigned long long value;
};

// Communication needed by Relation::distribute and the multi-rank join.  Built in: SingleRank, and
// ProcessComm for the ranks of one node (one forked process per GPU, host-side exchanges through
// shared memory -- what `mpirun -np N` on one node is to the reference).  Other launches use the
// Python launcher (NCCL) or MPI as sketched in INTEGRATION.md.
struct Comm {
  virtual ~Comm() {}
  virtual uint32_t rank() const = 0;
  virtual uint32_t size() const = 0;
  virtual void exchange(const void *send, void *recv, uint64_t bytes, uint32_t peer) = 0;  // pairwise swap
  virtual void allgatherDevice(const void *send_dev, void *recv_dev, uint64_t bytes_per_rank, void *stream) = 0;
  virtual void allgatherHost(const void *send, void *recv, uint64_t bytes_per_rank) = 0;
  virtual void barrier(void *stream) = 0;
};
struct SingleRank : Comm {
  uint32_t rank() const override { return 0; }
  uint32_t size() const override { return 1; }
  void exchange(const void *, void *, uint64_t, uint32_t) override {}
  void allgatherDevice(const void *, void *, uint64_t, void *) override {}
  void allgatherHost(const void *send, void *recv, uint64_t bytes) override;
  void barrier(void *) override {}
};

// N ranks as N processes of this node.  run() maps the shared control block, forks ranks 1..N-1
// (it must be called before anything initialises CUDA in the calling process), runs `body` in
// every rank and returns the largest exit code.  The payload of the join never goes through it:
// only the histogram all-gather (a few KiB), the window handles, barriers, and Relation::distribute.
class ProcessComm : public Comm {
 public:
  static int run(uint32_t ranks, int (*body)(ProcessComm &comm, void *arg), void *arg);
  uint32_t rank() const override { return me; }
  uint32_t size() const override { return n; }
  void exchange(const void *send, void *recv, uint64_t bytes, uint32_t peer) override;
  void allgatherDevice(const void *send_dev, void *recv_dev, uint64_t bytes_per_rank, void *stream) override;
  void allgatherHost(const void *send, void *recv, uint64_t bytes_per_rank) override;
  void barrier(void *stream) override;

 private:
  ProcessComm(uint32_t me, uint32_t n, void *shared) : me(me), n(n), shared(shared), sense(0) {}
  uint32_t me, n;
  void *shared;
  uint32_t sense;
};

class Relation {  // data/Relation.h:18-60
 public:
  Relation(uint64_t localSize, uint64_t globalSize);
  ~Relation();
  uint64_t getLocalSize() const { return localSize; }
  uint64_t getGlobalSize() const { return globalSize; }
  Tuple *getData() { return data; }
  void fillUniqueValues(uint64_t startKeyValue, uint64_t startRidValue);                              // Relation.cpp:63-73
  void fillModuloValues(uint64_t startKeyValue, uint64_t startRidValue, uint64_t innerRelationSize);  // :75-85
  void distribute(Comm &comm);                                                                        // :99-141
 protected:
  void randomOrder();  // :87-97
  uint64_t localSize, globalSize;
  Tuple *data;  // pinned host memory (what MPI_Alloc_mem / posix_memalign is to the reference)
};
}  // namespace data

namespace memory {
// memory/Pool.h:19-27 on device memory.  Bound to the context of the HashJoin that is alive.
class Pool {
 public:
  static void bind(rhj_ctx *ctx) { context = ctx; }
  static void allocate(uint64_t size);
  static void *getMemory(uint64_t size);
  static void reset();
  static void freeAll();
 private:
  static rhj_ctx *context;
};
}  // namespace memory

namespace performance {
// The phase-level hooks of performance/Measurements.h:27-41,66-76 and the keys of
// storePhaseData/storeSpecialData (Measurements.cpp:136-142, 175-179); device time comes
// from the library's cudaEvents (rhj_get_timings), not gettimeofday.
class Measurements {
 public:
  static void init(uint32_t nodeId, uint32_t numberOfNodes, const std::string &tag);
  static void startJoin();
  static void stopJoin(rhj_ctx *ctx);
  static void storeAllMeasurements();
  static void printMeasurements(uint32_t numberOfNodes, uint32_t nodeId, const std::vector<uint64_t> &tuplesPerNode);
  static std::vector<uint64_t> serializeResults(uint64_t matches);  // performance/Measurements.cpp:548-565
  static rhj_timings last;
  static double wallJoinUs;
 private:
  static std::string directory;
  static uint32_t node;
  static double t0;
};
}  // namespace performance

namespace tasks {
class Task {  // tasks/Task.h:20-30
 public:
  virtual ~Task() {}
  virtual void execute() = 0;
  virtual task_type_t getType() = 0;
};
}  // namespace tasks

namespace operators {
class HashJoin {  // operators/HashJoin.h:23-43
 public:
  HashJoin(uint32_t numberOfNodes, uint32_t nodeId, data::Relation *innerRelation, data::Relation *outerRelation,
           data::Comm *comm = nullptr, int device = 0);
  ~HashJoin();
  void join();
  rhj_ctx *context() { return ctx; }
  // Materialised output (no counterpart in the reference, which only counts): set
  // RESULT_PAIR_CAPACITY before constructing the operator and join() also writes up to that many
  // (rid_R, rid_S) pairs of this rank to device memory; getResultPairs copies them to the host.
  static uint64_t RESULT_PAIR_CAPACITY;
  rhj_pair *resultPairsDevice() { return dPairs; }
  std::vector<rhj_pair> getResultPairs();

  static uint64_t RESULT_COUNTER;                       // operators/HashJoin.cpp:28
  static uint64_t RESULT_CHECKSUM[4];                   // sum key, sum rid_R, sum rid_S, sum mix (this project's addition)
  static std::queue<hpcjoin_b200::tasks::Task *> TASK_QUEUE;  // operators/HashJoin.cpp:29

 protected:
  uint32_t numberOfNodes, nodeId;
  data::Relation *innerRelation, *outerRelation;
  data::Comm *comm;
  data::SingleRank single;
  rhj_ctx *ctx;
  void *dInner, *dOuter;  // device copies of the relations (pool memory)
  rhj_pair *dPairs;       // pool memory, RESULT_PAIR_CAPACITY entries (nullptr: count only)
  uint64_t pairsThisJoin;
  void *stream;
};
}  // namespace operators

namespace tasks {
// One Task per stage, each a single call through the C ABI (tasks/*.h in the reference).
class HistogramComputation : public Task {
 public:
  HistogramComputation(rhj_ctx *ctx, data::Comm *comm, const void *dInner, uint64_t nInner, const void *dOuter,
                       uint64_t nOuter, void *stream);
  void execute() override;
  task_type_t getType() override { return TASK_HISTOGRAM; }
  // tasks/HistogramComputation.cpp:78-130 (host copies)
  std::vector<uint32_t> getAssignment();
  std::vector<uint64_t> getInnerRelationLocalHistogram(), getOuterRelationLocalHistogram();
  std::vector<uint64_t> getInnerRelationGlobalHistogram(), getOuterRelationGlobalHistogram();
  std::vector<uint64_t> getInnerRelationBaseOffsets(), getOuterRelationBaseOffsets();
  std::vector<uint64_t> getInnerRelationWriteOffsets(), getOuterRelationWriteOffsets();
 private:
  rhj_ctx *ctx;
  data::Comm *comm;
  const void *dInner, *dOuter;
  uint64_t nInner, nOuter;
  void *stream;
  void *gathered = nullptr;
};
class NetworkPartitioning : public Task {
 public:
  NetworkPartitioning(rhj_ctx *ctx, void *stream) : ctx(ctx), stream(stream) {}
  void execute() override { check(rhj_network_partition(ctx, stream)); }
  task_type_t getType() override { return TASK_NET_PARTITION; }
 private:
  rhj_ctx *ctx;
  void *stream;
};
class LocalPartitioning : public Task {
 public:
  LocalPartitioning(rhj_ctx *ctx, void *stream, rhj_pair *pairs = nullptr, uint64_t capacity = 0)
      : ctx(ctx), stream(stream), pairs(pairs), capacity(capacity) {}
  void execute() override;  // also queues the BuildProbe task, as tasks/LocalPartitioning.cpp:116-124 does
  task_type_t getType() override { return TASK_PARTITION; }
 private:
  rhj_ctx *ctx;
  void *stream;
  rhj_pair *pairs;
  uint64_t capacity;
};
class BuildProbe : public Task {
 public:
  BuildProbe(rhj_ctx *ctx, void *stream, rhj_pair *pairs = nullptr, uint64_t capacity = 0)
      : ctx(ctx), stream(stream), pairs(pairs), capacity(capacity) {}
  void execute() override;  // adds this rank's matches to HashJoin::RESULT_COUNTER (tasks/BuildProbe.cpp:115)
  task_type_t getType() override { return TASK_BUILD_PROBE; }
 private:
  rhj_ctx *ctx;
  void *stream;
  rhj_pair *pairs;
  uint64_t capacity;
};
}  // namespace tasks

}  // namespace hpcjoin_b200
