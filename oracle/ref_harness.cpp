/*
 * ref_harness.cpp -- driver that runs the UNMODIFIED reference classes
 * (compiled in place from /root/reference by build_ref.sh) as N forked ranks
 * over oracle/stubs/mpi_shim.cpp.  TEST INFRASTRUCTURE ONLY.
 *
 * It plays the part of the reference's main.cpp:28-149 (whose sizes are
 * literals) with sizes / generator / rank count on the command line, and can
 * dump every intermediate the reference's getters expose, which is what pins
 * oracle/hpcjoin_oracle.c and the CUDA path.
 *
 *   --mode join    HashJoin::join() exactly as main.cpp:113-121 drives it
 *   --mode stages  HistogramComputation -> Window -> NetworkPartitioning ->
 *                  (LocalPartitioning) -> BuildProbe, dumping arrays to --dir
 *
 * Output: one line "REFJSON {...}" on stdout (counts and join microseconds per rank).
 */
#include <mpi.h>

#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <sys/time.h>
#include <sys/wait.h>
#include <unistd.h>
#include <vector>

#include "core/Configuration.h"
#include "data/CompressedTuple.h"
#include "data/Relation.h"
#include "data/Tuple.h"
#include "data/Window.h"
#include "memory/Pool.h"
#include "operators/HashJoin.h"
#include "tasks/BuildProbe.h"
#include "tasks/HistogramComputation.h"
#include "tasks/LocalPartitioning.h"
#include "tasks/NetworkPartitioning.h"

using hpcjoin::core::Configuration;
using hpcjoin::data::CompressedTuple;
using hpcjoin::data::Relation;
using hpcjoin::data::Tuple;
using hpcjoin::data::Window;

namespace {

struct Options {
  int ranks = 1;
  uint64_t inner = 1 << 20, outer = 1 << 20; /* tuples PER RANK */
  std::string gen = "unique", mode = "join", dir = ".";
  bool dump = false;
  unsigned seed = 1234;
};

struct RankResult {
  uint64_t matches;
  uint64_t join_us;
  uint64_t ok;
};

uint64_t now_us() {
  timeval tv;
  gettimeofday(&tv, nullptr);
  return uint64_t(tv.tv_sec) * 1000000ull + tv.tv_usec;
}

void write_file(const Options &o, int rank, const char *name, const void *data, size_t bytes) {
  char path[2048];
  snprintf(path, sizeof(path), "%s/rank%d_%s.bin", o.dir.c_str(), rank, name);
  FILE *f = fopen(path, "wb");
  if (!f) { perror(path); exit(2); }
  if (bytes && fwrite(data, 1, bytes, f) != bytes) { perror("fwrite"); exit(2); }
  fclose(f);
}

void read_file(const Options &o, int rank, const char *name, void *data, size_t bytes) {
  char path[2048];
  snprintf(path, sizeof(path), "%s/rank%d_%s.bin", o.dir.c_str(), rank, name);
  FILE *f = fopen(path, "rb");
  if (!f) { perror(path); exit(2); }
  if (bytes && fread(data, 1, bytes, f) != bytes) { fprintf(stderr, "short read %s\n", path); exit(2); }
  fclose(f);
}

/* exposes LocalPartitioning's protected static helpers (tasks/LocalPartitioning.h) */
struct LocalPartitioningProbe : hpcjoin::tasks::LocalPartitioning {
  using LocalPartitioning::computeHistogram;
  using LocalPartitioning::computePrefixSum;
  using LocalPartitioning::partitionData;
};

void dump_local_partitioning(const Options &o, int rank, const char *tag, CompressedTuple *part, uint64_t size,
                             uint32_t p) {
  const uint64_t L = Configuration::LOCAL_PARTITIONING_COUNT;
  uint64_t *hist = LocalPartitioningProbe::computeHistogram(part, size);
  uint64_t *prefix = LocalPartitioningProbe::computePrefixSum(hist);
  uint64_t out_n = size + L * 8;
  auto *out = static_cast<CompressedTuple *>(aligned_alloc(64, ((out_n * 8 + 63) / 64) * 64));
  memset(out, 0xFF, out_n * 8);
  LocalPartitioningProbe::partitionData(part, size, out, prefix, hist);
  char name[128];
  snprintf(name, sizeof(name), "%s_p%u_lhist", tag, p);
  write_file(o, rank, name, hist, L * 8);
  snprintf(name, sizeof(name), "%s_p%u_lprefix", tag, p);
  write_file(o, rank, name, prefix, L * 8);
  snprintf(name, sizeof(name), "%s_p%u_lout", tag, p);
  write_file(o, rank, name, out, out_n * 8);
  free(hist);
  free(prefix);
  free(out);
}

void drain_queue() {
  auto &q = hpcjoin::operators::HashJoin::TASK_QUEUE;
  while (!q.empty()) {
    hpcjoin::tasks::Task *t = q.front();
    q.pop();
    t->execute();
    delete t;
  }
}

int run_rank(const Options &o, int rank) {
  shim_set_rank(rank);
  MPI_Init(nullptr, nullptr);
  const uint32_t n = o.ranks;
  const uint64_t gi = o.inner * n, go = o.outer * n;

  /* main.cpp:86-90 */
  hpcjoin::memory::Pool::allocate(static_cast<uint64_t>(Configuration::ALLOCATION_FACTOR * (o.inner + o.outer) * sizeof(Tuple)));
  auto *inner = new Relation(o.inner, gi);
  auto *outer = new Relation(o.outer, go);

  if (o.gen == "file") {
    read_file(o, rank, "inner", inner->getData(), o.inner * sizeof(Tuple));
    read_file(o, rank, "outer", outer->getData(), o.outer * sizeof(Tuple));
  } else {
    srand(o.seed + rank); /* main.cpp:94 */
    inner->fillUniqueValues(rank * (gi / n), rank * (gi / n)); /* main.cpp:95-96 */
    if (o.gen == "unique") {
      outer->fillUniqueValues((n - rank - 1) * (go / n), rank * (go / n)); /* main.cpp:97-98 */
    } else { /* main.cpp:99 (commented-out variant) */
      outer->fillModuloValues((n - rank - 1) * (gi / n), rank * (go / n), inner->getLocalSize());
    }
    if (n > 1) { /* main.cpp:101-104 */
      inner->distribute(rank, n);
      outer->distribute(rank, n);
    }
  }
  if (o.dump) {
    write_file(o, rank, "inner", inner->getData(), o.inner * sizeof(Tuple));
    write_file(o, rank, "outer", outer->getData(), o.outer * sizeof(Tuple));
  }

  auto *res = reinterpret_cast<RankResult *>(static_cast<char *>(shim_result_area()) + size_t(rank) * 4096);
  if (o.mode == "join") {
    auto *hj = new hpcjoin::operators::HashJoin(n, rank, inner, outer);
    MPI_Barrier(MPI_COMM_WORLD);
    uint64_t t0 = now_us();
    hj->join();
    uint64_t t1 = now_us();
    res->matches = hpcjoin::operators::HashJoin::RESULT_COUNTER;
    res->join_us = t1 - t0;
  } else {
    const uint64_t P = Configuration::NETWORK_PARTITIONING_COUNT;
    MPI_Barrier(MPI_COMM_WORLD);
    uint64_t t0 = now_us();
    auto *hc = new hpcjoin::tasks::HistogramComputation(n, rank, inner, outer);
    hc->execute();
    if (o.dump) {
      write_file(o, rank, "assignment", hc->getAssignment(), P * 4);
      write_file(o, rank, "inner_local_hist", hc->getInnerRelationLocalHistogram(), P * 8);
      write_file(o, rank, "outer_local_hist", hc->getOuterRelationLocalHistogram(), P * 8);
      write_file(o, rank, "inner_global_hist", hc->getInnerRelationGlobalHistogram(), P * 8);
      write_file(o, rank, "outer_global_hist", hc->getOuterRelationGlobalHistogram(), P * 8);
      write_file(o, rank, "inner_base", hc->getInnerRelationBaseOffsets(), P * 8);
      write_file(o, rank, "outer_base", hc->getOuterRelationBaseOffsets(), P * 8);
      write_file(o, rank, "inner_write", hc->getInnerRelationWriteOffsets(), P * 8);
      write_file(o, rank, "outer_write", hc->getOuterRelationWriteOffsets(), P * 8);
    }
    auto *iw = new Window(n, rank, hc->getAssignment(), hc->getInnerRelationLocalHistogram(),
                          hc->getInnerRelationGlobalHistogram(), hc->getInnerRelationBaseOffsets(),
                          hc->getInnerRelationWriteOffsets());
    auto *ow = new Window(n, rank, hc->getAssignment(), hc->getOuterRelationLocalHistogram(),
                          hc->getOuterRelationGlobalHistogram(), hc->getOuterRelationBaseOffsets(),
                          hc->getOuterRelationWriteOffsets());
    auto *np = new hpcjoin::tasks::NetworkPartitioning(rank, inner, outer, iw, ow);
    np->execute();
    MPI_Barrier(MPI_COMM_WORLD);
    uint32_t *assign = hc->getAssignment();
    if (o.dump && uint64_t(rank) < P) {
      /* partition `rank` is the first one this rank owns, base offset 0 (OffsetMap.cpp:59-73) */
      write_file(o, rank, "inner_window", iw->getPartition(rank), iw->computeLocalWindowSize() * 8);
      write_file(o, rank, "outer_window", ow->getPartition(rank), ow->computeLocalWindowSize() * 8);
    }
    if (Configuration::ENABLE_TWO_LEVEL_PARTITIONING && o.dump) {
      for (uint32_t p = 0; p < P; ++p)
        if (assign[p] == uint32_t(rank)) {
          dump_local_partitioning(o, rank, "inner", iw->getPartition(p), iw->getPartitionSize(p), p);
          dump_local_partitioning(o, rank, "outer", ow->getPartition(p), ow->getPartitionSize(p), p);
        }
    }
    /* operators/HashJoin.cpp:131-166 + :187-204 */
    if (Configuration::ENABLE_TWO_LEVEL_PARTITIONING) hpcjoin::memory::Pool::reset();
    for (uint32_t p = 0; p < P; ++p) {
      if (assign[p] != uint32_t(rank)) continue;
      if (Configuration::ENABLE_TWO_LEVEL_PARTITIONING) {
        hpcjoin::operators::HashJoin::TASK_QUEUE.push(new hpcjoin::tasks::LocalPartitioning(
            iw->getPartitionSize(p), iw->getPartition(p), ow->getPartitionSize(p), ow->getPartition(p)));
      } else {
        hpcjoin::operators::HashJoin::TASK_QUEUE.push(new hpcjoin::tasks::BuildProbe(
            iw->getPartitionSize(p), iw->getPartition(p), ow->getPartitionSize(p), ow->getPartition(p)));
      }
    }
    drain_queue();
    uint64_t t1 = now_us();
    res->matches = hpcjoin::operators::HashJoin::RESULT_COUNTER;
    res->join_us = t1 - t0;
  }
  res->ok = 1;
  MPI_Barrier(MPI_COMM_WORLD);
  fflush(stdout);
  return 0;
}

}  // namespace

int main(int argc, char **argv) {
  Options o;
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    auto next = [&]() -> const char * {
      if (i + 1 >= argc) { fprintf(stderr, "missing value for %s\n", a.c_str()); exit(2); }
      return argv[++i];
    };
    if (a == "--ranks") o.ranks = atoi(next());
    else if (a == "--inner") o.inner = strtoull(next(), nullptr, 10);
    else if (a == "--outer") o.outer = strtoull(next(), nullptr, 10);
    else if (a == "--gen") o.gen = next();
    else if (a == "--mode") o.mode = next();
    else if (a == "--dir") o.dir = next();
    else if (a == "--seed") o.seed = atoi(next());
    else if (a == "--dump") o.dump = true;
    else if (a == "--config") {
      printf("REFJSON {\"net_bits\": %lu, \"loc_bits\": %lu, \"payload_bits\": %u, \"two_level\": %d}\n",
             (unsigned long)Configuration::NETWORK_PARTITIONING_FANOUT,
             (unsigned long)Configuration::LOCAL_PARTITIONING_FANOUT, Configuration::PAYLOAD_BITS,
             int(Configuration::ENABLE_TWO_LEVEL_PARTITIONING));
      return 0;
    } else { fprintf(stderr, "unknown argument %s\n", a.c_str()); return 2; }
  }
  /* shared heap: both windows of every rank (8 B per tuple), with slack for imbalance */
  size_t shared = size_t(o.ranks) * ((o.inner + o.outer) * 8 * 2 + (64u << 20));
  if (shim_setup(o.ranks, shared) != 0) return 3;

  std::vector<pid_t> kids;
  for (int r = 0; r < o.ranks; ++r) {
    pid_t pid = fork();
    if (pid < 0) { perror("fork"); return 3; }
    if (pid == 0) _exit(run_rank(o, r));
    kids.push_back(pid);
  }
  int bad = 0;
  for (pid_t k : kids) {
    int st = 0;
    waitpid(k, &st, 0);
    if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) ++bad;
  }
  auto *base = static_cast<char *>(shim_result_area());
  uint64_t total = 0, max_us = 0;
  std::string counts = "[", times = "[";
  for (int r = 0; r < o.ranks; ++r) {
    auto *res = reinterpret_cast<RankResult *>(base + size_t(r) * 4096);
    if (!res->ok) ++bad;
    total += res->matches;
    if (res->join_us > max_us) max_us = res->join_us;
    counts += std::to_string(res->matches) + (r + 1 < o.ranks ? ", " : "");
    times += std::to_string(res->join_us) + (r + 1 < o.ranks ? ", " : "");
  }
  counts += "]";
  times += "]";
  printf("REFJSON {\"ranks\": %d, \"inner_per_rank\": %lu, \"outer_per_rank\": %lu, \"mode\": \"%s\", "
         "\"gen\": \"%s\", \"matches\": %s, \"total_matches\": %lu, \"join_us\": %s, \"join_us_max\": %lu, "
         "\"failed_ranks\": %d}\n",
         o.ranks, (unsigned long)o.inner, (unsigned long)o.outer, o.mode.c_str(), o.gen.c_str(), counts.c_str(),
         (unsigned long)total, times.c_str(), (unsigned long)max_us, bad);
  return bad ? 4 : 0;
}
