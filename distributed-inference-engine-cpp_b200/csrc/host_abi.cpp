// C ABI over the host-side restatements (routing ring, cache keys, LRU, batching) so that the
// Python parity tests can drive them exactly like the reference objects behind oracle/_ref.
#include <chrono>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

#include "../../include/b200rn.h"
#include "../host/batch_processor.h"
#include "../host/consistent_hash.h"
#include "../host/lru_cache.h"

using FVec = std::vector<float>;
using Cache = LRUCache<FVec, FVec, VectorHash>;

struct b200rn_ring { ConsistentHash ring; explicit b200rn_ring(int v) : ring(v) {} };
struct b200rn_lru { Cache cache; explicit b200rn_lru(size_t c) : cache(c) {} };

extern "C" {

uint32_t b200rn_fnv1a(const char* key, size_t len) { return ConsistentHash::hashKey(key, len); }
b200rn_ring_t* b200rn_ring_create(int virtual_nodes) { return new b200rn_ring(virtual_nodes); }
void b200rn_ring_destroy(b200rn_ring_t* r) { delete r; }
void b200rn_ring_add(b200rn_ring_t* r, const char* node) { r->ring.addNode(node); }
void b200rn_ring_remove(b200rn_ring_t* r, const char* node) { r->ring.removeNode(node); }
int b200rn_ring_get(const b200rn_ring_t* r, const char* key, size_t len, char* out, size_t cap) {
  std::string s = r->ring.getNode(std::string(key, len));
  if (out && cap) std::snprintf(out, cap, "%s", s.c_str());
  return int(s.size());
}
int b200rn_ring_nodes(const b200rn_ring_t* r, char* out, size_t cap) {
  std::string j;
  for (auto& s : r->ring.getAllNodes()) { j += s; j += '\n'; }
  if (out && cap) std::snprintf(out, cap, "%s", j.c_str());
  return int(j.size());
}
uint64_t b200rn_hash_float(float f) { return VectorHash::hashFloat(f); }
uint64_t b200rn_vector_hash(const float* v, size_t n) { return VectorHash{}(FVec(v, v + n)); }

b200rn_lru_t* b200rn_lru_create(size_t capacity) { return new b200rn_lru(capacity); }
void b200rn_lru_destroy(b200rn_lru_t* c) { delete c; }
long b200rn_lru_get(b200rn_lru_t* c, const float* key, size_t n, float* out, size_t cap) {
  auto v = c->cache.get(FVec(key, key + n));
  if (!v) return -1;
  if (out) std::memcpy(out, v->data(), sizeof(float) * std::min(cap, v->size()));
  return long(v->size());
}
void b200rn_lru_put(b200rn_lru_t* c, const float* key, size_t n, const float* val, size_t m) {
  c->cache.put(FVec(key, key + n), FVec(val, val + m));
}
size_t b200rn_lru_size(const b200rn_lru_t* c) { return c->cache.size(); }
size_t b200rn_lru_hits(const b200rn_lru_t* c) { return c->cache.getHits(); }
size_t b200rn_lru_misses(const b200rn_lru_t* c) { return c->cache.getMisses(); }

void b200rn_batch_selftest(int max_batch, int timeout_ms, int callback_ms, int submitters, int per_thread,
                           long long* out, double* avg_batch) {
  BatchProcessor<int, int> bp(size_t(max_batch), std::chrono::milliseconds(timeout_ms),
                              [callback_ms](const std::vector<int>& xs) {
                                std::this_thread::sleep_for(std::chrono::milliseconds(callback_ms));
                                std::vector<int> ys;
                                ys.reserve(xs.size());
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
        if (bp.process(x) != x + 1) ++wrong;
      }
    });
  for (auto& t : ts) t.join();
  auto m = bp.getMetrics();
  bp.stop();
  out[0] = m.total_requests; out[1] = m.total_batches; out[2] = m.timeout_batches; out[3] = m.full_batches;
  out[4] = wrong.load();
  *avg_batch = m.avg_batch_size;
}

void b200rn_batch_selftest2(int max_batch, int timeout_ms, int callback_ms, int submitters, int per_thread, int fill_wait,
                            int in_flight, long long* out, double* avg_batch) {
  std::atomic<int> running{0}, peak{0};
  BatchProcessor<int, int> bp(size_t(max_batch), std::chrono::milliseconds(timeout_ms),
                              [&, callback_ms](const std::vector<int>& xs) {
                                const int now = ++running;
                                int seen = peak.load();
                                while (now > seen && !peak.compare_exchange_weak(seen, now)) {}
                                std::this_thread::sleep_for(std::chrono::milliseconds(callback_ms));
                                --running;
                                std::vector<int> ys;
                                ys.reserve(xs.size());
                                for (int x : xs) ys.push_back(x + 1);
                                return ys;
                              });
  bp.setFillWait(fill_wait != 0);
  bp.setInFlight(in_flight);
  bp.start();
  std::atomic<long long> wrong{0};
  std::vector<std::thread> ts;
  for (int t = 0; t < submitters; ++t)
    ts.emplace_back([&, t] {
      for (int i = 0; i < per_thread; ++i) {
        int x = t * 1000 + i;
        if (bp.process(x) != x + 1) ++wrong;
      }
    });
  for (auto& t : ts) t.join();
  auto m = bp.getMetrics();
  bp.stop();
  out[0] = m.total_requests; out[1] = m.total_batches; out[2] = m.timeout_batches; out[3] = m.full_batches;
  out[4] = wrong.load(); out[5] = peak.load();
  *avg_batch = m.avg_batch_size;
}

}  // extern "C"
