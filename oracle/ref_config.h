/*
 * ref_config.h -- harness-owned stand-in for the reference's
 * core/Configuration.h (its knobs are compile-time literals, :19-38).
 * build_ref.sh force-includes this file (-include) for every reference source;
 * defining the reference header's include guard makes the original a no-op, so
 * the reference .cpp files are compiled unmodified but with the knobs chosen by
 * -DREF_NET_BITS / -DREF_LOC_BITS / -DREF_PAYLOAD_BITS / -DREF_TWO_LEVEL.
 * TEST INFRASTRUCTURE ONLY.
 */
#ifndef HPCJOIN_CORE_CONFIGURATION_H_
#define HPCJOIN_CORE_CONFIGURATION_H_

#include <stdint.h>

#ifndef REF_NET_BITS
#define REF_NET_BITS 5
#endif
#ifndef REF_LOC_BITS
#define REF_LOC_BITS 5
#endif
#ifndef REF_PAYLOAD_BITS
#define REF_PAYLOAD_BITS 27
#endif
#ifndef REF_TWO_LEVEL
#define REF_TWO_LEVEL 0
#endif

namespace hpcjoin {
namespace core {

class Configuration {
 public:
  static const uint32_t RESULT_AGGREGATION_NODE = 0;
  static const uint32_t CACHELINE_SIZE_BYTES = 64;
  static const uint32_t CACHELINES_PER_MEMORY_BUFFER = 1024;
  static const uint32_t MEMORY_BUFFERS_PER_PARTITION = 2;
  static const uint64_t MEMORY_BUFFER_SIZE_BYTES = CACHELINES_PER_MEMORY_BUFFER * CACHELINE_SIZE_BYTES;
  static const uint64_t MEMORY_PARTITION_SIZE_BYTES = MEMORY_BUFFERS_PER_PARTITION * MEMORY_BUFFER_SIZE_BYTES;
  static const bool ENABLE_TWO_LEVEL_PARTITIONING = (REF_TWO_LEVEL != 0);
  static const uint64_t NETWORK_PARTITIONING_FANOUT = REF_NET_BITS;
  static const uint64_t LOCAL_PARTITIONING_FANOUT = REF_LOC_BITS;
  static const uint64_t NETWORK_PARTITIONING_COUNT = (1 << NETWORK_PARTITIONING_FANOUT);
  static const uint64_t LOCAL_PARTITIONING_COUNT = (1 << LOCAL_PARTITIONING_FANOUT);
  static constexpr double ALLOCATION_FACTOR = 1.1;
  static const uint32_t PAYLOAD_BITS = REF_PAYLOAD_BITS;
};

}  // namespace core
}  // namespace hpcjoin

#endif
