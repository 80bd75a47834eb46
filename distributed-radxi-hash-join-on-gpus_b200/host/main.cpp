// main.cpp -- driver in the shape of the reference's main.cpp:28-149 (whose sizes are literals):
// generate the relations with the reference's generator, distribute them, join on the GPU(s), report.
//   rhj_main [tuples_per_relation_per_rank] [net_bits loc_bits payload_bits two_level] [pair_capacity] [ranks]
// ranks > 1: one forked process per GPU of this node (rank r uses device r), as `mpirun -np ranks`.
#include <stdio.h>
#include <stdlib.h>

#include "hpcjoin_b200.hpp"

using namespace hpcjoin_b200;

struct Args {
  uint64_t n;
};

static int rank_main(data::ProcessComm &comm, void *argp) {
  const uint64_t n = static_cast<Args *>(argp)->n;
  const uint32_t numberOfNodes = comm.size(), nodeId = comm.rank();
  performance::Measurements::init(nodeId, numberOfNodes, "experiment");
  const uint64_t global = n * numberOfNodes;
  data::Relation inner(n, global), outer(n, global);
  srand(1234 + nodeId);                                                  // main.cpp:94
  inner.fillUniqueValues(nodeId * n, nodeId * n);                        // main.cpp:95-96
  outer.fillUniqueValues((numberOfNodes - nodeId - 1) * n, nodeId * n);  // main.cpp:97-98
  if (numberOfNodes > 1) {                                               // main.cpp:101-104
    inner.distribute(comm);
    outer.distribute(comm);
  }
  operators::HashJoin hashJoin(numberOfNodes, nodeId, &inner, &outer, &comm, (int)nodeId);
  hashJoin.join();
  // gather every rank's count and checksum (Measurements::storeAllMeasurements + the Send/Recv
  // of performance/Measurements.cpp:548-565 in the reference)
  uint64_t mine[5] = {operators::HashJoin::RESULT_COUNTER, operators::HashJoin::RESULT_CHECKSUM[0],
                      operators::HashJoin::RESULT_CHECKSUM[1], operators::HashJoin::RESULT_CHECKSUM[2],
                      operators::HashJoin::RESULT_CHECKSUM[3]};
  std::vector<uint64_t> all(5 * numberOfNodes);
  comm.allgatherHost(mine, all.data(), sizeof(mine));
  // the reference's per-rank 10-tuple (matches, total, hist, network, local, winalloc, netwait, locprep, LP sum, BP sum)
  const std::vector<uint64_t> ten = performance::Measurements::serializeResults(operators::HashJoin::RESULT_COUNTER);
  std::vector<uint64_t> tens(10 * numberOfNodes);
  comm.allgatherHost(ten.data(), tens.data(), 10 * sizeof(uint64_t));
  if (nodeId == 0) {
    std::vector<uint64_t> counts(numberOfNodes);
    for (uint32_t r = 0; r < numberOfNodes; ++r) counts[r] = all[5 * r];
    performance::Measurements::printMeasurements(numberOfNodes, nodeId, counts);
    for (uint32_t r = 0; r < numberOfNodes; ++r) {
      printf("[RESULTS] Rank %u:", r);
      for (int i = 0; i < 10; ++i) printf("\t%lu", (unsigned long)tens[10 * r + i]);
      printf("\n");
    }
    for (uint32_t r = 0; r < numberOfNodes; ++r)
      printf("[RESULTS] Checksum%s:\t%lu\t%lu\t%lu\t%lu\n", numberOfNodes > 1 ? (" rank " + std::to_string(r)).c_str() : "",
             (unsigned long)all[5 * r + 1], (unsigned long)all[5 * r + 2], (unsigned long)all[5 * r + 3],
             (unsigned long)all[5 * r + 4]);
  }
  if (operators::HashJoin::RESULT_PAIR_CAPACITY && nodeId == 0 && numberOfNodes == 1) {
    // materialised output: fold it so a test can compare
    uint64_t sr = 0, ss = 0, bad = 0;
    const std::vector<rhj_pair> pairs = hashJoin.getResultPairs();
    for (const rhj_pair &p : pairs) {
      sr += p.rid_r;
      ss += p.rid_s;
      // keys of both sides at these rids must be equal (rid = position, Relation.cpp:63-73)
      bad += inner.getData()[p.rid_r].key != outer.getData()[p.rid_s].key;
    }
    printf("[RESULTS] Pairs:\t%lu\t%lu\t%lu\t%lu\n", (unsigned long)pairs.size(), (unsigned long)sr, (unsigned long)ss,
           (unsigned long)bad);
  }
  performance::Measurements::storeAllMeasurements();
  return 0;
}

int main(int argc, char **argv) {
  Args args{argc > 1 ? strtoull(argv[1], nullptr, 10) : 20000000ull};  // main.cpp:70-71
  if (argc > 5) {
    core::Configuration::NETWORK_PARTITIONING_FANOUT = atoi(argv[2]);
    core::Configuration::LOCAL_PARTITIONING_FANOUT = atoi(argv[3]);
    core::Configuration::PAYLOAD_BITS = atoi(argv[4]);
    core::Configuration::ENABLE_TWO_LEVEL_PARTITIONING = atoi(argv[5]) != 0;
  }
  if (argc > 6) operators::HashJoin::RESULT_PAIR_CAPACITY = strtoull(argv[6], nullptr, 10);
  const uint32_t ranks = argc > 7 ? (uint32_t)atoi(argv[7]) : 1;
  try {
    return data::ProcessComm::run(ranks, rank_main, &args);
  } catch (const Error &e) {
    fprintf(stderr, "rhj error %d: %s\n", e.code, e.message.c_str());
    return 1;
  }
}
