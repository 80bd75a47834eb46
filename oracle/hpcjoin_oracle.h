/*
 * hpcjoin_oracle.h -- CPU restatement of the reference's radix hash join path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (the package directory,
 * include/) may include, link or call this.  Allowed users: tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
 *
 * Parity status: the reference ships no tests or golden vectors,
 * so this restatement is pinned against the reference itself: oracle/_ref is the
 * UNMODIFIED reference compiled from /root/reference by oracle/build_ref.sh, and
 * tests/test_oracle_vs_ref.py + tests/golden/ compare every stage of this file
 * with it (generator output, histograms, offsets, window contents, local
 * partitions, match counts).
 *
 * All citations are file:line in /root/reference.
 */
#ifndef HPCJOIN_ORACLE_H_
#define HPCJOIN_ORACLE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* data/Tuple.h:15-22 */
typedef struct {
  uint64_t key;
  uint64_t rid;
} orc_tuple;

/* core/Configuration.h:28-37 -- the knobs that change results */
typedef struct {
  uint32_t net_bits;     /* NETWORK_PARTITIONING_FANOUT */
  uint32_t loc_bits;     /* LOCAL_PARTITIONING_FANOUT */
  uint32_t payload_bits; /* PAYLOAD_BITS */
  uint32_t two_level;    /* ENABLE_TWO_LEVEL_PARTITIONING */
  uint32_t balanced;     /* 0: AssignmentMap.cpp's p % n; 1: size-balanced assignment (SURVEY.md 8(f3),
                            no counterpart in the reference's built code) */
} orc_config;

/* Result of a join: the reference only counts (tasks/BuildProbe.cpp:97-115);
 * the sums are this project's order-independent checksum (all mod 2^64). */
typedef struct {
  uint64_t matches;
  uint64_t sum_key;
  uint64_t sum_rid_r;
  uint64_t sum_rid_s;
  uint64_t sum_mix;
} orc_result;

/* ---- generators ------------------------------------------------------- */

/* One glibc rand() stream (TYPE_3, as srand()/rand() use), so several ranks
 * can be simulated in one process.  main.cpp:94 seeds 1234+rank. */
typedef struct orc_rng orc_rng;
orc_rng *orc_rng_new(unsigned int seed);
void orc_rng_free(orc_rng *rng);
int32_t orc_rng_next(orc_rng *rng);

void orc_fill_unique(orc_rng *rng, orc_tuple *t, uint64_t n, uint64_t start_key, uint64_t start_rid);
void orc_fill_modulo(orc_rng *rng, orc_tuple *t, uint64_t n, uint64_t start_key, uint64_t start_rid,
                     uint64_t inner_size);
void orc_random_order(orc_rng *rng, orc_tuple *t, uint64_t n);
/* Relation::distribute for all ranks at once (equal local sizes). */
int orc_distribute(orc_rng **rngs, orc_tuple **rels, uint64_t local_size, uint32_t nranks);
/* the same for the harness generator: section swaps, then key'[i] = key[orc_perm(i, local_size, seed + rank)] */
int orc_distribute_perm(orc_tuple **rels, uint64_t local_size, uint32_t nranks, uint64_t seed);

/* Harness-owned counter-based generator for sizes where rand() Fisher-Yates
 * is too slow (SURVEY.md §7 hard part 6): key = perm_seed(first+i) over
 * [0, domain), optionally reduced modulo `modulo` (0 = off); rid = start_rid+i. */
uint64_t orc_perm(uint64_t x, uint64_t domain, uint64_t seed);
void orc_fill_perm(orc_tuple *t, uint64_t n, uint64_t first, uint64_t domain, uint64_t seed,
                   uint64_t modulo, uint64_t start_key, uint64_t start_rid);
/* Zipf(theta) keys over [0,domain) through perm (hot keys spread over partitions). */
void orc_fill_zipf(orc_tuple *t, uint64_t n, uint64_t first, uint64_t domain, double theta, uint64_t seed,
                   uint64_t start_rid);

/* ---- histogram phase --------------------------------------------------- */
void orc_local_histogram(const orc_tuple *t, uint64_t n, const orc_config *cfg, uint64_t *hist);
void orc_global_histogram(const uint64_t *locals, uint32_t nranks, uint32_t parts, uint64_t *global);
void orc_assignment(uint32_t nranks, uint32_t parts, uint32_t *assignment);
/* Size-balanced assignment: partitions by descending weight (global inner + outer count, ties by
 * ascending id) go to the least-loaded rank (ties: lowest rank) that still has a free slot, every
 * rank keeping exactly the number of partitions the round-robin map gives it. */
void orc_assignment_balanced(const uint64_t *global_inner, const uint64_t *global_outer, uint32_t nranks,
                             uint32_t parts, uint32_t *assignment);
void orc_offsets(const uint64_t *locals, const uint64_t *global, const uint32_t *assignment,
                 uint32_t nranks, uint32_t rank, uint32_t parts, uint64_t *base, uint64_t *rel,
                 uint64_t *abs);
uint64_t orc_window_size(const uint64_t *global, const uint32_t *assignment, uint32_t parts, uint32_t rank);

/* ---- network partitioning --------------------------------------------- */
uint64_t orc_compress(uint64_t key, uint64_t rid, const orc_config *cfg);
void orc_network_partition(const orc_tuple *t, uint64_t n, const orc_config *cfg,
                           const uint32_t *assignment, const uint64_t *abs_offsets,
                           uint64_t **windows);

/* ---- local partitioning ------------------------------------------------ */
void orc_local_part_histogram(const uint64_t *part, uint64_t n, const orc_config *cfg, uint64_t *hist);
void orc_local_part_prefix(const uint64_t *hist, uint32_t subparts, uint64_t *prefix);
void orc_local_part_scatter(const uint64_t *part, uint64_t n, const orc_config *cfg,
                            const uint64_t *prefix, uint64_t *out);

/* ---- build / probe ------------------------------------------------------ */
uint64_t orc_mix(uint64_t key, uint64_t rid_r, uint64_t rid_s);
void orc_build_probe(const uint64_t *r, uint64_t nr, const uint64_t *s, uint64_t ns,
                     const orc_config *cfg, uint32_t partition, orc_result *acc);

/* Materialised output: the reference's probe loop with `++matches`
 * (tasks/BuildProbe.cpp:101-104) additionally appending (rid_R, rid_S).  `pairs` holds
 * 2*cap u64; `n` counts every match, also those beyond `cap` (which are not stored). */
typedef struct {
  uint64_t *pairs;
  uint64_t cap;
  uint64_t n;
} orc_pair_sink;
void orc_build_probe_pairs(const uint64_t *r, uint64_t nr, const uint64_t *s, uint64_t ns,
                           const orc_config *cfg, uint32_t partition, orc_result *acc, orc_pair_sink *sink);

/* ---- the whole operator, all ranks simulated in-process ------------------ */
int orc_join(const orc_tuple *const *inner, const uint64_t *inner_n, const orc_tuple *const *outer,
             const uint64_t *outer_n, uint32_t nranks, const orc_config *cfg, orc_result *per_rank);
/* orc_join with one pair sink per rank (NULL: count only). */
int orc_join_pairs(const orc_tuple *const *inner, const uint64_t *inner_n, const orc_tuple *const *outer,
                   const uint64_t *outer_n, uint32_t nranks, const orc_config *cfg, orc_result *per_rank,
                   orc_pair_sink *sinks);
/* Same pipeline starting from already-compressed 8-byte tuples (BASELINE config 5):
 * `value` must be packed as orc_compress() would, partition = low net_bits of `key`
 * supplied beside it is impossible, so input is (partition-less) raw: key in the
 * high bits, see orc_pack8/orc_unpack8. */
uint64_t orc_pack8(uint64_t key, uint64_t rid, uint32_t rid_bits);

/* Independent checker for small inputs: sort-merge equi-join on the raw tuples. */
void orc_bruteforce_join(const orc_tuple *r, uint64_t nr, const orc_tuple *s, uint64_t ns, orc_result *out);

#ifdef __cplusplus
}
#endif
#endif
