// LRU cache for inference results, keyed by the raw request vector.
//
// Same public surface, hit/miss/eviction behaviour and cache keys as the reference's LRUCache and
// VectorHash (reference: include/lru_cache.h:14-96): get() promotes and counts a hit or a miss,
// put() on a present key overwrites + promotes, put() on a new key at capacity evicts the least
// recently used entry first; key equality is Key::operator== (for vector<float>: -0.0 == +0.0, NaN
// never equal).  The storage is different: each key is held ONCE (the reference keeps every 602 KB
// image key twice, once in its list and once in its map).
#pragma once
#include <atomic>
#include <cstdint>
#include <cstring>
#include <functional>
#include <mutex>
#include <optional>
#include <unordered_map>
#include <vector>

template <typename Key, typename Value, typename Hash = std::hash<Key>>
class LRUCache {
  // One heap item per entry holds the key and the value once; the recency list is threaded
  // through the items and the index maps a pointer to the stored key onto its item.
  struct Item {
    Key key;
    Value value;
    Item* newer = nullptr;
    Item* older = nullptr;
  };
  struct KeyPtrHash {
    Hash h;
    size_t operator()(const Key* k) const { return h(*k); }
  };
  struct KeyPtrEq {
    // identity first, so an entry can always find itself (keys holding NaN are not == themselves)
    bool operator()(const Key* a, const Key* b) const { return a == b || *a == *b; }
  };

 public:
  explicit LRUCache(size_t capacity) : capacity_(capacity) {}
  ~LRUCache() { drop_all(); }
  LRUCache(const LRUCache&) = delete;
  LRUCache& operator=(const LRUCache&) = delete;

  std::optional<Value> get(const Key& key) {
    std::lock_guard<std::mutex> g(mu_);
    auto it = index_.find(&key);
    if (it == index_.end()) {
      ++misses_;
      return std::nullopt;
    }
    touch(it->second);
    ++hits_;
    return it->second->value;
  }

  void put(const Key& key, const Value& value) { put(Key(key), value); }

  // the key is moved into the entry (the caller's copy of a 602 KB sample is the one the cache keeps)
  void put(Key&& key, const Value& value) {
    std::lock_guard<std::mutex> g(mu_);
    auto it = index_.find(&key);
    if (it != index_.end()) {
      it->second->value = value;
      touch(it->second);
      return;
    }
    if (count_ >= capacity_ && oldest_) {
      Item* victim = oldest_;
      unlink(victim);
      index_.erase(&victim->key);
      delete victim;
      --count_;
    }
    Item* item = new Item{std::move(key), value, nullptr, nullptr};
    index_.emplace(&item->key, item);
    pushNewest(item);
    ++count_;
  }

  void clear() {
    std::lock_guard<std::mutex> g(mu_);
    drop_all();
    hits_ = 0;
    misses_ = 0;
  }

  size_t size() const {
    std::lock_guard<std::mutex> g(mu_);
    return count_;
  }
  size_t capacity() const { return capacity_; }
  size_t getHits() const { return hits_.load(); }
  size_t getMisses() const { return misses_.load(); }
  double getHitRate() const {
    const size_t h = hits_.load(), total = h + misses_.load();
    return total ? static_cast<double>(h) / static_cast<double>(total) : 0.0;
  }

 private:
  void drop_all() {
    index_.clear();
    for (Item* p = newest_; p;) {
      Item* next = p->older;
      delete p;
      p = next;
    }
    newest_ = oldest_ = nullptr;
    count_ = 0;
  }
  void unlink(Item* n) {
    if (n->newer) n->newer->older = n->older; else newest_ = n->older;
    if (n->older) n->older->newer = n->newer; else oldest_ = n->newer;
    n->newer = n->older = nullptr;
  }
  void pushNewest(Item* n) {
    n->older = newest_;
    n->newer = nullptr;
    if (newest_) newest_->newer = n;
    newest_ = n;
    if (!oldest_) oldest_ = n;
  }
  void touch(Item* n) {
    if (n == newest_) return;
    unlink(n);
    pushNewest(n);
  }

  size_t capacity_;
  size_t count_ = 0;
  std::unordered_map<const Key*, Item*, KeyPtrHash, KeyPtrEq> index_;
  Item* newest_ = nullptr;
  Item* oldest_ = nullptr;
  mutable std::mutex mu_;
  std::atomic<size_t> hits_{0};
  std::atomic<size_t> misses_{0};
};

// Cache key hash for a request vector: boost-style combine over three sampled elements
// (first, middle, last), each hashed the way libstdc++'s std::hash<float> does on x86-64
// (0 for +-0, otherwise the 64-bit Murmur-style _Hash_bytes of the 4 value bytes with seed
// 0xc70f6907).  Written out explicitly so the keys stay bit-identical to the reference's
// (include/lru_cache.h:84-96) under any standard library.
struct VectorHash {
  static uint64_t hashFloat(float f) {
    if (f == 0.0f) return 0;
    constexpr uint64_t kMul = (uint64_t(0xc6a4a793UL) << 32) + uint64_t(0x5bd1e995UL);
    uint32_t bits;
    std::memcpy(&bits, &f, sizeof bits);
    uint64_t h = uint64_t(0xc70f6907UL) ^ (4 * kMul);
    h ^= uint64_t(bits);
    h *= kMul;
    h = (h ^ (h >> 47)) * kMul;
    h ^= h >> 47;
    return h;
  }
  static uint64_t mix(uint64_t seed, uint64_t v) { return seed ^ (v + 0x9e3779b9 + (seed << 6) + (seed >> 2)); }

  size_t operator()(const std::vector<float>& v) const {
    uint64_t h = 0;
    if (!v.empty()) {
      h = mix(h, hashFloat(v.front()));
      h = mix(h, hashFloat(v[v.size() / 2]));
      h = mix(h, hashFloat(v.back()));
    }
    return static_cast<size_t>(h);
  }
};
