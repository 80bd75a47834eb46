// main.cpp -- single-rank driver in the shape of the reference's main.cpp:28-149 (whose sizes
// are literals): generate the relations with the reference's generator, join on the GPU, report.
//   rhj_main [tuples_per_relation] [net_bits loc_bits payload_bits two_level] [pair_capacity]
#include <stdio.h>
#include <stdlib.h>

#include "hpcjoin_b200.hpp"

using namespace hpcjoin_b200;

int main(int argc, char **argv) {
  const uint64_t n = argc > 1 ? strtoull(argv[1], nullptr, 10) : 20000000ull;  // main.cpp:70-71
  if (argc > 5) {
    core::Configuration::NETWORK_PARTITIONING_FANOUT = atoi(argv[2]);
    core::Configuration::LOCAL_PARTITIONING_FANOUT = atoi(argv[3]);
    core::Configuration::PAYLOAD_BITS = atoi(argv[4]);
    core::Configuration::ENABLE_TWO_LEVEL_PARTITIONING = atoi(argv[5]) != 0;
  }
  if (argc > 6) operators::HashJoin::RESULT_PAIR_CAPACITY = strtoull(argv[6], nullptr, 10);
  try {
    const uint32_t numberOfNodes = 1, nodeId = 0;
    performance::Measurements::init(nodeId, numberOfNodes, "experiment");
    data::Relation inner(n, n), outer(n, n);
    srand(1234 + nodeId);                                      // main.cpp:94
    inner.fillUniqueValues(nodeId * n, nodeId * n);            // main.cpp:95-96
    outer.fillUniqueValues((numberOfNodes - nodeId - 1) * n, nodeId * n);  // main.cpp:97-98
    operators::HashJoin hashJoin(numberOfNodes, nodeId, &inner, &outer);
    hashJoin.join();
    performance::Measurements::printMeasurements(numberOfNodes, nodeId, {operators::HashJoin::RESULT_COUNTER});
    printf("[RESULTS] Checksum:\t%lu\t%lu\t%lu\t%lu\n", (unsigned long)operators::HashJoin::RESULT_CHECKSUM[0],
           (unsigned long)operators::HashJoin::RESULT_CHECKSUM[1], (unsigned long)operators::HashJoin::RESULT_CHECKSUM[2],
           (unsigned long)operators::HashJoin::RESULT_CHECKSUM[3]);
    if (operators::HashJoin::RESULT_PAIR_CAPACITY) {  // materialised output: fold it so a test can compare
      uint64_t sr = 0, ss = 0, bad = 0;
      const std::vector<rhj_pair> pairs = hashJoin.getResultPairs();
      for (const rhj_pair &p : pairs) {
        sr += p.rid_r;
        ss += p.rid_s;
        // keys of both sides at these rids must be equal (rid = position, Relation.cpp:63-73)
        bad += inner.getData()[p.rid_r - nodeId * n].key != outer.getData()[p.rid_s - nodeId * n].key;
      }
      printf("[RESULTS] Pairs:\t%lu\t%lu\t%lu\t%lu\n", (unsigned long)pairs.size(), (unsigned long)sr,
             (unsigned long)ss, (unsigned long)bad);
    }
    performance::Measurements::storeAllMeasurements();
  } catch (const Error &e) {
    fprintf(stderr, "rhj error %d: %s\n", e.code, e.message.c_str());
    return 1;
  }
  return 0;
}
