// TEST INFRASTRUCTURE ONLY (oracle).  C shim over the REFERENCE'S OWN sources, compiled where they
// lie under /root/reference by oracle/Makefile into oracle/_ref/libref_routing.so (git-ignored).
// Nothing from /root/reference is copied into this repository: this file only #includes the
// reference headers at build time and forwards calls.  Used by tests to pin
// oracle/routing_oracle.c and the product's host-side restatement to the real thing.
//
// Wrapped reference types: ConsistentHash (include/consistent_hash.h, src/consistent_hash.cpp),
// LRUCache + VectorHash (include/lru_cache.h), BatchProcessor (include/batch_processor.h).
#define private public  // to reach ConsistentHash::hash (src/consistent_hash.cpp:6); test shim only
#include "consistent_hash.h"
#undef private
#include "batch_processor.h"
#include "lru_cache.h"

#include <chrono>
#include <cstring>
#include <thread>
#include <vector>

using FVec = std::vector<float>;
using RefLRU = LRUCache<FVec, FVec, VectorHash>;

extern "C" {

void *ref_ring_new(int vnodes) { return new ConsistentHash(vnodes); }
void ref_ring_free(void *r) { delete static_cast<ConsistentHash *>(r); }
void ref_ring_add(void *r, const char *node) { static_cast<ConsistentHash *>(r)->addNode(node); }
void ref_ring_remove(void *r, const char *node) { static_cast<ConsistentHash *>(r)->removeNode(node); }
uint32_t ref_ring_hash(void *r, const char *key, size_t n) {
  return static_cast<ConsistentHash *>(r)->hash(std::string(key, n));
}
// writes the owning node name (NUL-terminated) into out; returns its length
int ref_ring_get(void *r, const char *key, size_t n, char *out, size_t cap) {
  std::string s = static_cast<ConsistentHash *>(r)->getNode(std::string(key, n));
  std::snprintf(out, cap, "%s", s.c_str());
  return (int)s.size();
}
// all nodes in ring order, joined by '\n'
int ref_ring_all(void *r, char *out, size_t cap) {
  std::string j;
  for (auto &s : static_cast<ConsistentHash *>(r)->getAllNodes()) { j += s; j += '\n'; }
  std::snprintf(out, cap, "%s", j.c_str());
  return (int)j.size();
}

uint64_t ref_hash_float(float f) { return std::hash<float>{}(f); }
uint64_t ref_vector_hash(const float *v, size_t n) { return VectorHash{}(FVec(v, v + n)); }

void *ref_lru_new(size_t cap) { return new RefLRU(cap); }
void ref_lru_free(void *c) { delete static_cast<RefLRU *>(c); }
// returns value length on hit (copied into out, up to cap floats), -1 on miss
long ref_lru_get(void *c, const float *k, size_t n, float *out, size_t cap) {
  auto v = static_cast<RefLRU *>(c)->get(FVec(k, k + n));
  if (!v) return -1;
  std::memcpy(out, v->data(), sizeof(float) * std::min(cap, v->size()));
  return (long)v->size();
}
void ref_lru_put(void *c, const float *k, size_t n, const float *v, size_t m) {
  static_cast<RefLRU *>(c)->put(FVec(k, k + n), FVec(v, v + m));
}
size_t ref_lru_size(void *c) { return static_cast<RefLRU *>(c)->size(); }
size_t ref_lru_hits(void *c) { return static_cast<RefLRU *>(c)->getHits(); }
size_t ref_lru_misses(void *c) { return static_cast<RefLRU *>(c)->getMisses(); }

// BatchProcessor<int,int> with a callback that sleeps callback_ms and returns x+1 for each x.
// `submitters` threads each submit `per_thread` requests.  out[0..4] = total_requests,
// total_batches, timeout_batches, full_batches, wrong_answers; *avg = avg_batch_size.
void ref_batch_run(int max_batch, int timeout_ms, int callback_ms, int submitters, int per_thread,
                   long long *out, double *avg) {
  BatchProcessor<int, int> bp((size_t)max_batch, std::chrono::milliseconds(timeout_ms),
                              [callback_ms](const std::vector<int> &xs) {
                                std::this_thread::sleep_for(std::chrono::milliseconds(callback_ms));
                                std::vector<int> ys;
                                for (int x : xs) ys.push_back(x + 1);
                                return ys;
                              });
  bp.start();
  std::atomic<long long> wrong{0};
  std::vector<std::thread> ts;
  for (int t = 0; t < submitters; ++t)
    ts.emplace_back([&, t] {
      for (int i = 0; i < per_thread; ++i) {
        int x = t * 1000 + i;
        if (bp.process(x) != x + 1) wrong++;
      }
    });
  for (auto &t : ts) t.join();
  auto m = bp.getMetrics();
  bp.stop();
  out[0] = m.total_requests; out[1] = m.total_batches; out[2] = m.timeout_batches;
  out[3] = m.full_batches; out[4] = wrong.load();
  *avg = m.avg_batch_size;
}

}  // extern "C"
