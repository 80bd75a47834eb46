/*
 * hpcjoin_oracle.c -- CPU restatement of the reference's radix hash join path.
 * TEST INFRASTRUCTURE ONLY (see hpcjoin_oracle.h for who may use it).
 *
 * Each function names the reference lines it restates.  Where the reference's
 * mechanics (64 KiB staging buffers, AVX streaming stores, MPI_Put batching) do
 * not influence the bytes that end up in memory, only the resulting layout is
 * restated, and the comment says so.
 */
#define _GNU_SOURCE
#include "hpcjoin_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------ */
/* rand() streams                                                            */
/* ------------------------------------------------------------------------ */

struct orc_rng {
  struct random_data data;
  char state[128]; /* glibc's rand() default table: TYPE_3, 128 bytes */
};

orc_rng *orc_rng_new(unsigned int seed) {
  orc_rng *r = (orc_rng *)calloc(1, sizeof(orc_rng));
  if (!r) return NULL;
  initstate_r(seed, r->state, sizeof(r->state), &r->data);
  srandom_r(seed, &r->data);
  return r;
}

void orc_rng_free(orc_rng *rng) { free(rng); }

int32_t orc_rng_next(orc_rng *rng) {
  int32_t v;
  random_r(&rng->data, &v);
  return v;
}

/* data/Relation.cpp:19  RAND_RANGE(N) = ((double) rand() / ((double) RAND_MAX + 1)) * N */
static uint64_t rand_range(orc_rng *rng, uint64_t n) {
  double u = (double)orc_rng_next(rng) / ((double)RAND_MAX + 1);
  return (uint64_t)(u * (double)n);
}

/* data/Relation.cpp:87-97: Fisher-Yates from the top, on the key field only */
void orc_random_order(orc_rng *rng, orc_tuple *t, uint64_t n) {
  if (n < 2) return;
  for (uint64_t i = n - 1; i > 0; --i) {
    uint64_t j = rand_range(rng, i);
    uint64_t k = t[i].key;
    t[i].key = t[j].key;
    t[j].key = k;
  }
}

/* data/Relation.cpp:63-73 */
void orc_fill_unique(orc_rng *rng, orc_tuple *t, uint64_t n, uint64_t start_key, uint64_t start_rid) {
  for (uint64_t i = 0; i < n; ++i) t[i].key = start_key + i;
  orc_random_order(rng, t, n);
  for (uint64_t i = 0; i < n; ++i) t[i].rid = start_rid + i;
}

/* data/Relation.cpp:75-85 */
void orc_fill_modulo(orc_rng *rng, orc_tuple *t, uint64_t n, uint64_t start_key, uint64_t start_rid,
                     uint64_t inner_size) {
  for (uint64_t i = 0; i < n; ++i) t[i].key = start_key + (i % inner_size);
  orc_random_order(rng, t, n);
  for (uint64_t i = 0; i < n; ++i) t[i].rid = start_rid + i;
}

/* data/Relation.cpp:99-141.  Rank `me` and rank `i` swap their sections numbered
 * (me+i) % n (both sides compute the same section index), section (2*me)%n stays;
 * afterwards every rank reshuffles its keys (:139). Whole tuples (key AND rid) move. */
/* Relation::distribute (data/Relation.cpp:99-141) for the harness's counter-based generator: the same
 * pairwise section swaps (rank a and rank b exchange section (a + b) % n), then -- where the reference
 * shuffles the key field with rand() (randomOrder, :87-97: keys move, rids stay) -- the keys of rank r
 * are permuted by the generator's bijection over positions: key'[i] = key[orc_perm(i, local_size, seed + r)].
 * The device-side distribute of the launchers must reproduce this bit for bit. */
int orc_distribute_perm(orc_tuple **rels, uint64_t local_size, uint32_t nranks, uint64_t seed) {
  uint64_t sec = local_size / nranks;
  uint64_t last = local_size - (uint64_t)(nranks - 1) * sec;
  orc_tuple *tmp = (orc_tuple *)malloc((last ? last : 1) * sizeof(orc_tuple));
  uint64_t *keys = (uint64_t *)malloc((local_size ? local_size : 1) * sizeof(uint64_t));
  if (!tmp || !keys) { free(tmp); free(keys); return -1; }
  for (uint32_t a = 0; a < nranks; ++a) {
    for (uint32_t b = a + 1; b < nranks; ++b) {
      uint32_t section = (a + b) % nranks;
      uint64_t start = (uint64_t)section * sec;
      uint64_t size = (section == nranks - 1) ? last : sec;
      memcpy(tmp, rels[a] + start, size * sizeof(orc_tuple));
      memcpy(rels[a] + start, rels[b] + start, size * sizeof(orc_tuple));
      memcpy(rels[b] + start, tmp, size * sizeof(orc_tuple));
    }
  }
  for (uint32_t r = 0; r < nranks; ++r) {
    for (uint64_t i = 0; i < local_size; ++i) keys[i] = rels[r][i].key;
    for (uint64_t i = 0; i < local_size; ++i) rels[r][i].key = keys[orc_perm(i, local_size, seed + r)];
  }
  free(tmp);
  free(keys);
  return 0;
}

int orc_distribute(orc_rng **rngs, orc_tuple **rels, uint64_t local_size, uint32_t nranks) {
  uint64_t sec = local_size / nranks;
  uint64_t last = local_size - (uint64_t)(nranks - 1) * sec;
  orc_tuple *tmp = (orc_tuple *)malloc((last ? last : 1) * sizeof(orc_tuple));
  if (!tmp) return -1;
  for (uint32_t a = 0; a < nranks; ++a) {
    for (uint32_t b = a + 1; b < nranks; ++b) {
      uint32_t section = (a + b) % nranks;
      uint64_t start = (uint64_t)section * sec;
      uint64_t size = (section == nranks - 1) ? last : sec;
      memcpy(tmp, rels[a] + start, size * sizeof(orc_tuple));
      memcpy(rels[a] + start, rels[b] + start, size * sizeof(orc_tuple));
      memcpy(rels[b] + start, tmp, size * sizeof(orc_tuple));
    }
  }
  free(tmp);
  for (uint32_t r = 0; r < nranks; ++r) orc_random_order(rngs[r], rels[r], local_size);
  return 0;
}

/* ------------------------------------------------------------------------ */
/* counter-based generator (harness-owned, no reference counterpart)          */
/* ------------------------------------------------------------------------ */

static inline uint32_t bits_for(uint64_t domain) {
  uint32_t b = 0;
  while (b < 64 && (b == 64 ? 0 : ((uint64_t)1 << b)) < domain) ++b;
  return b;
}

static inline uint64_t splitmix(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

/* A bijection on [0, 2^b): odd-multiply/add and xor-shift-right rounds are each
 * invertible mod 2^b.  Cycle-walk back into [0, domain). */
uint64_t orc_perm(uint64_t x, uint64_t domain, uint64_t seed) {
  if (domain <= 1) return 0;
  uint32_t b = bits_for(domain);
  uint64_t mask = (b >= 64) ? ~0ull : (((uint64_t)1 << b) - 1);
  uint32_t s = b / 2 ? b / 2 : 1;
  uint64_t m0 = splitmix(seed) | 1, a0 = splitmix(seed + 1);
  uint64_t m1 = splitmix(seed + 2) | 1, a1 = splitmix(seed + 3);
  uint64_t m2 = splitmix(seed + 4) | 1, a2 = splitmix(seed + 5);
  do {
    x = (x * m0 + a0) & mask;
    x ^= x >> s;
    x = (x * m1 + a1) & mask;
    x ^= x >> s;
    x = (x * m2 + a2) & mask;
    x ^= x >> s;
  } while (x >= domain);
  return x;
}

void orc_fill_perm(orc_tuple *t, uint64_t n, uint64_t first, uint64_t domain, uint64_t seed,
                   uint64_t modulo, uint64_t start_key, uint64_t start_rid) {
  for (uint64_t i = 0; i < n; ++i) {
    uint64_t k = orc_perm(first + i, domain, seed);
    if (modulo) k %= modulo;
    t[i].key = start_key + k;
    t[i].rid = start_rid + i;
  }
}

/* Zipf by inverse-CDF on the continuous approximation (Gray et al.-style
 * closed form would need zeta; for a test generator the rejection-free
 * approximation below is enough: rank = floor(exp(u * ln(domain+1))) - 1 for
 * theta == 1, general theta via the power-law inverse).  Ranks are then sent
 * through orc_perm so that hot keys do not all share low bits. */
void orc_fill_zipf(orc_tuple *t, uint64_t n, uint64_t first, uint64_t domain, double theta, uint64_t seed,
                   uint64_t start_rid) {
  double dn = (double)domain;
  for (uint64_t i = 0; i < n; ++i) {
    uint64_t h = splitmix(seed ^ splitmix(first + i));
    double u = (double)(h >> 11) * (1.0 / 9007199254740992.0); /* [0,1) */
    double rank;
    if (fabs(theta - 1.0) < 1e-9) {
      rank = exp(u * log(dn + 1.0)) - 1.0;
    } else {
      double e = 1.0 - theta;
      rank = pow(u * (pow(dn + 1.0, e) - 1.0) + 1.0, 1.0 / e) - 1.0;
    }
    uint64_t r = (uint64_t)rank;
    if (r >= domain) r = domain - 1;
    t[i].key = orc_perm(r, domain, seed + 77);
    t[i].rid = start_rid + i;
  }
}

/* ------------------------------------------------------------------------ */
/* histogram phase                                                           */
/* ------------------------------------------------------------------------ */

/* histograms/LocalHistogram.cpp:35-53: values[key & (P-1)]++ */
void orc_local_histogram(const orc_tuple *t, uint64_t n, const orc_config *cfg, uint64_t *hist) {
  uint64_t parts = (uint64_t)1 << cfg->net_bits;
  memset(hist, 0, parts * sizeof(uint64_t));
  for (uint64_t i = 0; i < n; ++i) ++hist[t[i].key & (parts - 1)];
}

/* histograms/GlobalHistogram.cpp:31-48: MPI_Allreduce(SUM) over ranks */
void orc_global_histogram(const uint64_t *locals, uint32_t nranks, uint32_t parts, uint64_t *global) {
  for (uint32_t p = 0; p < parts; ++p) {
    uint64_t s = 0;
    for (uint32_t r = 0; r < nranks; ++r) s += locals[(uint64_t)r * parts + p];
    global[p] = s;
  }
}

/* histograms/AssignmentMap.cpp:34-49: static round-robin */
void orc_assignment(uint32_t nranks, uint32_t parts, uint32_t *assignment) {
  for (uint32_t p = 0; p < parts; ++p) assignment[p] = p % nranks;
}

typedef struct {
  uint64_t w;
  uint32_t p;
} weighted_part;

static int cmp_weight_desc(const void *a, const void *b) {
  const weighted_part *x = (const weighted_part *)a, *y = (const weighted_part *)b;
  if (x->w != y->w) return x->w > y->w ? -1 : 1;
  return x->p < y->p ? -1 : (x->p > y->p);
}

void orc_assignment_balanced(const uint64_t *global_inner, const uint64_t *global_outer, uint32_t nranks,
                             uint32_t parts, uint32_t *assignment) {
  weighted_part *wp = (weighted_part *)malloc(parts * sizeof(weighted_part));
  uint64_t *load = (uint64_t *)calloc(nranks, sizeof(uint64_t));
  uint32_t *room = (uint32_t *)malloc(nranks * sizeof(uint32_t));
  for (uint32_t p = 0; p < parts; ++p) {
    wp[p].w = global_inner[p] + global_outer[p];
    wp[p].p = p;
  }
  qsort(wp, parts, sizeof(weighted_part), cmp_weight_desc);
  for (uint32_t r = 0; r < nranks; ++r) room[r] = parts / nranks + (r < parts % nranks ? 1u : 0u);
  for (uint32_t i = 0; i < parts; ++i) {
    uint32_t best = nranks;
    for (uint32_t r = 0; r < nranks; ++r)
      if (room[r] > 0 && (best == nranks || load[r] < load[best])) best = r;
    assignment[wp[i].p] = best;
    load[best] += wp[i].w;
    room[best] -= 1;
  }
  free(wp);
  free(load);
  free(room);
}

/* histograms/OffsetMap.cpp:59-73 (base: running sum per owner),
 * :75-85 (relative: MPI_Exscan over ranks; rank 0 keeps calloc's zero),
 * :87-93 (absolute = base + relative).  Units: tuples. */
void orc_offsets(const uint64_t *locals, const uint64_t *global, const uint32_t *assignment,
                 uint32_t nranks, uint32_t rank, uint32_t parts, uint64_t *base, uint64_t *rel,
                 uint64_t *abs) {
  uint64_t *cur = (uint64_t *)calloc(nranks, sizeof(uint64_t));
  for (uint32_t p = 0; p < parts; ++p) {
    uint32_t owner = assignment[p];
    base[p] = cur[owner];
    cur[owner] += global[p];
  }
  free(cur);
  for (uint32_t p = 0; p < parts; ++p) {
    uint64_t s = 0;
    for (uint32_t r = 0; r < rank; ++r) s += locals[(uint64_t)r * parts + p];
    rel[p] = s;
    abs[p] = base[p] + s;
  }
}

/* data/Window.cpp:168-178 */
uint64_t orc_window_size(const uint64_t *global, const uint32_t *assignment, uint32_t parts, uint32_t rank) {
  uint64_t s = 0;
  for (uint32_t p = 0; p < parts; ++p)
    if (assignment[p] == rank) s += global[p];
  return s;
}

/* ------------------------------------------------------------------------ */
/* network partitioning                                                      */
/* ------------------------------------------------------------------------ */

/* tasks/NetworkPartitioning.cpp:128-129:
 * value = rid + ((key >> F_net) << (F_net + PAYLOAD_BITS)) */
uint64_t orc_compress(uint64_t key, uint64_t rid, const orc_config *cfg) {
  return rid + ((key >> cfg->net_bits) << (cfg->net_bits + cfg->payload_bits));
}

/* tasks/NetworkPartitioning.cpp:116-173 + :186-209 + data/Window.cpp:86-144.
 * The reference appends each tuple to its partition's cache line, then to a
 * 64 KiB staging buffer, and puts full buffers (and the remainders, in
 * partition order) at writeOffsets[p] + writeCounters[p].  Since every put of
 * partition p lands right after the previous one, the bytes that end up in
 * the windows are: partition p of rank `assignment[p]`, starting at this
 * rank's absolute offset, holds this rank's tuples of p in scan order. */
void orc_network_partition(const orc_tuple *t, uint64_t n, const orc_config *cfg,
                           const uint32_t *assignment, const uint64_t *abs_offsets,
                           uint64_t **windows) {
  uint64_t parts = (uint64_t)1 << cfg->net_bits;
  uint64_t *written = (uint64_t *)calloc(parts, sizeof(uint64_t));
  for (uint64_t i = 0; i < n; ++i) {
    uint64_t p = t[i].key & (parts - 1);
    windows[assignment[p]][abs_offsets[p] + written[p]++] = orc_compress(t[i].key, t[i].rid, cfg);
  }
  free(written);
}

/* ------------------------------------------------------------------------ */
/* local partitioning                                                        */
/* ------------------------------------------------------------------------ */

static inline uint64_t local_digit(uint64_t value, const orc_config *cfg) {
  uint32_t shift = cfg->net_bits + cfg->payload_bits;
  uint64_t mask = (((uint64_t)1 << cfg->loc_bits) - 1) << shift;
  return (value & mask) >> shift;
}

/* tasks/LocalPartitioning.cpp:138-163 */
void orc_local_part_histogram(const uint64_t *part, uint64_t n, const orc_config *cfg, uint64_t *hist) {
  uint64_t sub = (uint64_t)1 << cfg->loc_bits;
  memset(hist, 0, sub * sizeof(uint64_t));
  for (uint64_t i = 0; i < n; ++i) ++hist[local_digit(part[i], cfg)];
}

/* tasks/LocalPartitioning.cpp:165-192: exclusive sum, each start padded up to a
 * multiple of 8 tuples (one 64-byte cache line) */
void orc_local_part_prefix(const uint64_t *hist, uint32_t subparts, uint64_t *prefix) {
  uint64_t sum = 0;
  for (uint32_t p = 0; p < subparts; ++p) {
    prefix[p] = sum;
    sum += hist[p];
    if (sum % 8 != 0) sum += 8 - (sum % 8);
  }
}

/* tasks/LocalPartitioning.cpp:194-250: stable scatter (cache-line buffering only
 * delays the stores; slot order per sub-partition is input order).  `out` holds
 * n + subparts*8 values; padding slots are left untouched. */
void orc_local_part_scatter(const uint64_t *part, uint64_t n, const orc_config *cfg,
                            const uint64_t *prefix, uint64_t *out) {
  uint64_t sub = (uint64_t)1 << cfg->loc_bits;
  uint64_t *slot = (uint64_t *)malloc(sub * sizeof(uint64_t));
  memcpy(slot, prefix, sub * sizeof(uint64_t));
  for (uint64_t i = 0; i < n; ++i) out[slot[local_digit(part[i], cfg)]++] = part[i];
  free(slot);
}

/* ------------------------------------------------------------------------ */
/* build / probe                                                             */
/* ------------------------------------------------------------------------ */

/* Checksum term per matching pair (project-defined; the reference has none). */
uint64_t orc_mix(uint64_t key, uint64_t rid_r, uint64_t rid_s) {
  uint32_t k32 = (uint32_t)key ^ (uint32_t)(key >> 32);
  uint32_t rr = (uint32_t)rid_r ^ (uint32_t)(rid_r >> 32);
  uint32_t rs = (uint32_t)rid_s ^ (uint32_t)(rid_s >> 32);
  uint64_t a = (uint64_t)(rr ^ 0x9E3779B9u) * (uint64_t)(rs ^ 0x85EBCA6Bu);
  uint64_t b = (uint64_t)(k32 ^ 0xC2B2AE35u) * (uint64_t)(uint32_t)(rr + rs + 0x27D4EB2Fu);
  return a + b;
}

/* tasks/BuildProbe.cpp:16-25 -- 32-bit-safe round-up, kept as is */
static uint64_t next_pow2_ref(uint64_t v) {
  v--;
  v |= v >> 1;
  v |= v >> 2;
  v |= v >> 4;
  v |= v >> 8;
  v |= v >> 16;
  v++;
  return v;
}

/* tasks/BuildProbe.cpp:47-121.  Bucket-chained table: bucket[] holds 1-based
 * index of the newest tuple, next[] the previous one (LIFO chains); hash bits
 * start at F_net+PAYLOAD_BITS+F_loc whether or not the local pass ran (:55-57);
 * a probe tuple matches when value>>keyShift is equal (:101-102), every
 * duplicate counts.  `partition` (not known to the reference task) restores
 * the key's low bits for the checksum. */
/* tasks/BuildProbe.cpp:47-121.  The reference only counts (`:97-106`); with a sink, every match
 * additionally appends (rid_R, rid_S) -- the materialised operator output. */
void orc_build_probe_pairs(const uint64_t *r, uint64_t nr, const uint64_t *s, uint64_t ns,
                           const orc_config *cfg, uint32_t partition, orc_result *acc, orc_pair_sink *sink) {
  if (nr == 0 || ns == 0) return;
  uint32_t key_shift = cfg->net_bits + cfg->payload_bits;
  uint32_t shift_bits = key_shift + cfg->loc_bits;
  uint64_t n = next_pow2_ref(nr);
  uint64_t mask = (n - 1) << shift_bits;
  uint64_t rid_mask = (key_shift >= 64) ? ~0ull : (((uint64_t)1 << key_shift) - 1);

  uint64_t *next = (uint64_t *)calloc(nr, sizeof(uint64_t));
  uint64_t *bucket = (uint64_t *)calloc(n, sizeof(uint64_t));
  for (uint64_t t = 0; t < nr;) {
    uint64_t idx = (r[t] & mask) >> shift_bits;
    next[t] = bucket[idx];
    bucket[idx] = ++t;
  }
  for (uint64_t t = 0; t < ns; ++t) {
    uint64_t idx = (s[t] & mask) >> shift_bits;
    for (uint64_t hit = bucket[idx]; hit > 0; hit = next[hit - 1]) {
      if ((s[t] >> key_shift) == (r[hit - 1] >> key_shift)) {
        uint64_t key = ((s[t] >> key_shift) << cfg->net_bits) | partition;
        uint64_t rid_r = r[hit - 1] & rid_mask;
        uint64_t rid_s = s[t] & rid_mask;
        acc->matches += 1;
        acc->sum_key += key;
        acc->sum_rid_r += rid_r;
        acc->sum_rid_s += rid_s;
        acc->sum_mix += orc_mix(key, rid_r, rid_s);
        if (sink) {
          if (sink->n < sink->cap) {
            sink->pairs[2 * sink->n] = rid_r;
            sink->pairs[2 * sink->n + 1] = rid_s;
          }
          sink->n += 1;
        }
      }
    }
  }
  free(next);
  free(bucket);
}

void orc_build_probe(const uint64_t *r, uint64_t nr, const uint64_t *s, uint64_t ns,
                     const orc_config *cfg, uint32_t partition, orc_result *acc) {
  orc_build_probe_pairs(r, nr, s, ns, cfg, partition, acc, NULL);
}

/* ------------------------------------------------------------------------ */
/* operators/HashJoin.cpp:45-220 for all ranks, simulated in one process        */
/* ------------------------------------------------------------------------ */

static void join_partition(const uint64_t *rp, uint64_t nr, const uint64_t *sp, uint64_t ns,
                           const orc_config *cfg, uint32_t p, orc_result *acc, orc_pair_sink *sink) {
  if (!cfg->two_level) {
    /* HashJoin.cpp:151-156: one BuildProbe per owned network partition */
    orc_build_probe_pairs(rp, nr, sp, ns, cfg, p, acc, sink);
    return;
  }
  /* tasks/LocalPartitioning.cpp:59-136 */
  uint32_t sub = 1u << cfg->loc_bits;
  uint64_t *hr = (uint64_t *)malloc(sub * sizeof(uint64_t)), *hs = (uint64_t *)malloc(sub * sizeof(uint64_t));
  uint64_t *pr = (uint64_t *)malloc(sub * sizeof(uint64_t)), *ps = (uint64_t *)malloc(sub * sizeof(uint64_t));
  orc_local_part_histogram(rp, nr, cfg, hr);
  orc_local_part_histogram(sp, ns, cfg, hs);
  orc_local_part_prefix(hr, sub, pr);
  orc_local_part_prefix(hs, sub, ps);
  uint64_t *ro = (uint64_t *)malloc((nr + (uint64_t)sub * 8) * sizeof(uint64_t));
  uint64_t *so = (uint64_t *)malloc((ns + (uint64_t)sub * 8) * sizeof(uint64_t));
  orc_local_part_scatter(rp, nr, cfg, pr, ro);
  orc_local_part_scatter(sp, ns, cfg, ps, so);
  for (uint32_t q = 0; q < sub; ++q)
    if (hr[q] > 0 && hs[q] > 0) orc_build_probe_pairs(ro + pr[q], hr[q], so + ps[q], hs[q], cfg, p, acc, sink);
  free(hr); free(hs); free(pr); free(ps); free(ro); free(so);
}

int orc_join(const orc_tuple *const *inner, const uint64_t *inner_n, const orc_tuple *const *outer,
             const uint64_t *outer_n, uint32_t nranks, const orc_config *cfg, orc_result *per_rank) {
  return orc_join_pairs(inner, inner_n, outer, outer_n, nranks, cfg, per_rank, NULL);
}

int orc_join_pairs(const orc_tuple *const *inner, const uint64_t *inner_n, const orc_tuple *const *outer,
                   const uint64_t *outer_n, uint32_t nranks, const orc_config *cfg, orc_result *per_rank,
                   orc_pair_sink *sinks) {
  uint32_t parts = 1u << cfg->net_bits;
  uint64_t *loc[2], *glob[2];
  uint32_t *assign = (uint32_t *)malloc(parts * sizeof(uint32_t));
  uint64_t **win[2];
  uint64_t *base[2];
  const orc_tuple *const *rel[2] = {inner, outer};
  const uint64_t *rel_n[2] = {inner_n, outer_n};

  for (int k = 0; k < 2; ++k) {
    loc[k] = (uint64_t *)malloc((uint64_t)nranks * parts * sizeof(uint64_t));
    glob[k] = (uint64_t *)malloc(parts * sizeof(uint64_t));
    for (uint32_t r = 0; r < nranks; ++r)
      orc_local_histogram(rel[k][r], rel_n[k][r], cfg, loc[k] + (uint64_t)r * parts);
    orc_global_histogram(loc[k], nranks, parts, glob[k]);
  }
  if (cfg->balanced) orc_assignment_balanced(glob[0], glob[1], nranks, parts, assign);
  else orc_assignment(nranks, parts, assign);
  for (int k = 0; k < 2; ++k) {
    win[k] = (uint64_t **)malloc(nranks * sizeof(uint64_t *));
    for (uint32_t r = 0; r < nranks; ++r) {
      uint64_t ws = orc_window_size(glob[k], assign, parts, r);
      win[k][r] = (uint64_t *)malloc((ws ? ws : 1) * sizeof(uint64_t));
    }
    base[k] = (uint64_t *)malloc(parts * sizeof(uint64_t));
    uint64_t *rl = (uint64_t *)malloc(parts * sizeof(uint64_t));
    uint64_t *ab = (uint64_t *)malloc(parts * sizeof(uint64_t));
    for (uint32_t r = 0; r < nranks; ++r) {
      orc_offsets(loc[k], glob[k], assign, nranks, r, parts, base[k], rl, ab);
      orc_network_partition(rel[k][r], rel_n[k][r], cfg, assign, ab, win[k]);
    }
    free(rl);
    free(ab);
  }
  for (uint32_t r = 0; r < nranks; ++r) {
    memset(&per_rank[r], 0, sizeof(orc_result));
    for (uint32_t p = 0; p < parts; ++p) {
      if (assign[p] != r) continue;
      join_partition(win[0][r] + base[0][p], glob[0][p], win[1][r] + base[1][p], glob[1][p], cfg, p,
                     &per_rank[r], sinks ? &sinks[r] : NULL);
    }
  }
  for (int k = 0; k < 2; ++k) {
    for (uint32_t r = 0; r < nranks; ++r) free(win[k][r]);
    free(win[k]); free(base[k]); free(loc[k]); free(glob[k]);
  }
  free(assign);
  return 0;
}

/* BASELINE config 5's 8-byte input layout: key in the bits above `rid_bits`. */
uint64_t orc_pack8(uint64_t key, uint64_t rid, uint32_t rid_bits) { return (key << rid_bits) | rid; }

/* ------------------------------------------------------------------------ */
/* independent checker: sort both sides by key, merge                         */
/* ------------------------------------------------------------------------ */

static int cmp_key(const void *a, const void *b) {
  uint64_t x = ((const orc_tuple *)a)->key, y = ((const orc_tuple *)b)->key;
  return (x > y) - (x < y);
}

void orc_bruteforce_join(const orc_tuple *r, uint64_t nr, const orc_tuple *s, uint64_t ns, orc_result *out) {
  memset(out, 0, sizeof(*out));
  orc_tuple *a = (orc_tuple *)malloc((nr ? nr : 1) * sizeof(orc_tuple));
  orc_tuple *b = (orc_tuple *)malloc((ns ? ns : 1) * sizeof(orc_tuple));
  memcpy(a, r, nr * sizeof(orc_tuple));
  memcpy(b, s, ns * sizeof(orc_tuple));
  qsort(a, nr, sizeof(orc_tuple), cmp_key);
  qsort(b, ns, sizeof(orc_tuple), cmp_key);
  uint64_t i = 0, j = 0;
  while (i < nr && j < ns) {
    if (a[i].key < b[j].key) { ++i; continue; }
    if (a[i].key > b[j].key) { ++j; continue; }
    uint64_t k = a[i].key, i2 = i, j2 = j;
    while (i2 < nr && a[i2].key == k) ++i2;
    while (j2 < ns && b[j2].key == k) ++j2;
    for (uint64_t x = i; x < i2; ++x)
      for (uint64_t y = j; y < j2; ++y) {
        out->matches += 1;
        out->sum_key += k;
        out->sum_rid_r += a[x].rid;
        out->sum_rid_s += b[y].rid;
        out->sum_mix += orc_mix(k, a[x].rid, b[y].rid);
      }
    i = i2;
    j = j2;
  }
  free(a);
  free(b);
}
