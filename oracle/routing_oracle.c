/* TEST INFRASTRUCTURE ONLY (oracle) -- plain-C restatement of the reference's request-routing
 * and cache-key integer arithmetic.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg may link or call this.  The product's own implementation lives in
 * distributed-inference-engine-cpp_b200/host/ and is checked against this and against the
 * reference's sources compiled into oracle/_ref (see oracle/Makefile).
 *
 * Pinned: checked in tests/test_routing_oracle.py against the golden vectors harvested from the
 * reference's own sources and, when oracle/_ref was built, against the
 * reference objects themselves.
 *
 * What each function follows (paths relative to /root/reference):
 *   ro_fnv1a            <- ConsistentHash::hash, src/consistent_hash.cpp:6-14.  NB each `char` is
 *                          sign-extended before the XOR (:10), unlike canonical FNV-1a.
 *   ro_ring_build       <- ConsistentHash::addNode, src/consistent_hash.cpp:16-23 (vnode key
 *                          node + "#" + i; a later insert overwrites an equal hash, :21)
 *   ro_ring_lookup      <- ConsistentHash::getNode, src/consistent_hash.cpp:34-45 (lower_bound,
 *                          wrap to begin; empty ring -> -1 standing for "")
 *   ro_hash_float       <- std::hash<float> of libstdc++ as used by VectorHash
 *                          (include/lru_cache.h:87): 0 for +-0, else _Hash_bytes(&f,4,0xc70f6907)
 *                          (/usr/include/c++/13/bits/functional_hash.h:205,240-248; algorithm:
 *                          libsupc++ hash_bytes.cc, 64-bit Murmur-style mix)
 *   ro_vector_hash      <- VectorHash::operator(), include/lru_cache.h:84-96
 */
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

uint32_t ro_fnv1a(const char *s, size_t n) {
  uint32_t h = 2166136261u;
  for (size_t i = 0; i < n; ++i) {
    h ^= (uint32_t)(int32_t)(signed char)s[i]; /* sign extension, as `char` on x86-64 */
    h *= 16777619u;
  }
  return h;
}

typedef struct {
  uint32_t h;
  int32_t owner;
  uint32_t seq; /* insertion order, to resolve equal hashes the way std::map::operator[] does */
} ro_slot;

static int slot_cmp(const void *a, const void *b) {
  const ro_slot *x = (const ro_slot *)a, *y = (const ro_slot *)b;
  if (x->h != y->h) return x->h < y->h ? -1 : 1;
  return x->seq < y->seq ? -1 : (x->seq > y->seq);
}

/* names: n_nodes NUL-terminated strings.  out_h/out_owner need room for n_nodes*vnodes entries.
 * Returns the number of distinct ring positions, sorted ascending by hash. */
size_t ro_ring_build(const char *const *names, size_t n_nodes, int vnodes, uint32_t *out_h,
                     int32_t *out_owner) {
  size_t total = n_nodes * (size_t)vnodes, k = 0;
  ro_slot *s = (ro_slot *)malloc(sizeof(ro_slot) * (total ? total : 1));
  char buf[512];
  for (size_t n = 0; n < n_nodes; ++n)
    for (int i = 0; i < vnodes; ++i) {
      int len = snprintf(buf, sizeof buf, "%s#%d", names[n], i);
      s[k].h = ro_fnv1a(buf, (size_t)len);
      s[k].owner = (int32_t)n;
      s[k].seq = (uint32_t)k;
      ++k;
    }
  qsort(s, total, sizeof(ro_slot), slot_cmp);
  size_t m = 0;
  for (size_t i = 0; i < total; ++i) {
    if (m && out_h[m - 1] == s[i].h) {
      out_owner[m - 1] = s[i].owner; /* later insert wins */
    } else {
      out_h[m] = s[i].h;
      out_owner[m] = s[i].owner;
      ++m;
    }
  }
  free(s);
  return m;
}

int32_t ro_ring_lookup(const uint32_t *ring_h, const int32_t *ring_owner, size_t m, const char *key,
                       size_t key_len) {
  if (m == 0) return -1;
  uint32_t h = ro_fnv1a(key, key_len);
  size_t lo = 0, hi = m; /* first position with ring_h >= h */
  while (lo < hi) {
    size_t mid = (lo + hi) / 2;
    if (ring_h[mid] < h) lo = mid + 1; else hi = mid;
  }
  if (lo == m) lo = 0;
  return ring_owner[lo];
}

static uint64_t shift_mix(uint64_t v) { return v ^ (v >> 47); }

uint64_t ro_hash_float(float f) {
  if (f == 0.0f) return 0; /* both +0 and -0 */
  const uint64_t mul = (((uint64_t)0xc6a4a793UL) << 32) + (uint64_t)0x5bd1e995UL;
  uint32_t bits;
  memcpy(&bits, &f, 4);
  uint64_t hash = (uint64_t)0xc70f6907UL ^ (4 * mul);
  /* len 4 < 8: no aligned words; tail bytes loaded little-endian */
  hash ^= (uint64_t)bits;
  hash *= mul;
  hash = shift_mix(hash) * mul;
  hash = shift_mix(hash);
  return hash;
}

uint64_t ro_vector_hash(const float *v, size_t n) {
  uint64_t h = 0;
  if (n) {
    h ^= ro_hash_float(v[0]) + 0x9e3779b9 + (h << 6) + (h >> 2);
    h ^= ro_hash_float(v[n / 2]) + 0x9e3779b9 + (h << 6) + (h >> 2);
    h ^= ro_hash_float(v[n - 1]) + 0x9e3779b9 + (h << 6) + (h >> 2);
  }
  return h;
}
