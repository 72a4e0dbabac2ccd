#include "consistent_hash.h"

#include <algorithm>
#include <set>

uint32_t ConsistentHash::hashKey(const char* data, size_t len) {
  uint32_t h = 0x811C9DC5u;                      // FNV offset basis 2166136261
  for (size_t i = 0; i < len; ++i) {
    // the reference XORs a (signed) char converted to uint32_t: bytes >= 0x80 flip the top 24 bits
    h ^= static_cast<uint32_t>(static_cast<int32_t>(static_cast<signed char>(data[i])));
    h *= 0x01000193u;                            // FNV prime 16777619
  }
  return h;
}

size_t ConsistentHash::lowerBound(uint32_t h) const {
  size_t lo = 0, hi = ring_.size();
  while (lo < hi) {
    size_t mid = lo + (hi - lo) / 2;
    if (ring_[mid].first < h) lo = mid + 1; else hi = mid;
  }
  return lo;
}

void ConsistentHash::addNode(const std::string& node) {
  std::lock_guard<std::mutex> g(mu_);
  for (int i = 0; i < vnodes_; ++i) {
    const std::string v = node + "#" + std::to_string(i);
    const uint32_t h = hashKey(v.data(), v.size());
    const size_t at = lowerBound(h);
    if (at < ring_.size() && ring_[at].first == h) ring_[at].second = node;   // later insert wins
    else ring_.insert(ring_.begin() + static_cast<std::ptrdiff_t>(at), {h, node});
  }
}

void ConsistentHash::removeNode(const std::string& node) {
  std::lock_guard<std::mutex> g(mu_);
  for (int i = 0; i < vnodes_; ++i) {
    const std::string v = node + "#" + std::to_string(i);
    const uint32_t h = hashKey(v.data(), v.size());
    const size_t at = lowerBound(h);
    // like the reference, the position is dropped whoever owns it now
    if (at < ring_.size() && ring_[at].first == h) ring_.erase(ring_.begin() + static_cast<std::ptrdiff_t>(at));
  }
}

std::string ConsistentHash::getNode(const std::string& key) const {
  std::lock_guard<std::mutex> g(mu_);
  if (ring_.empty()) return std::string();
  size_t at = lowerBound(hashKey(key.data(), key.size()));
  if (at == ring_.size()) at = 0;
  return ring_[at].second;
}

std::vector<std::string> ConsistentHash::getAllNodes() const {
  std::lock_guard<std::mutex> g(mu_);
  std::vector<std::string> out;
  std::set<std::string> seen;
  for (const auto& slot : ring_)
    if (seen.insert(slot.second).second) out.push_back(slot.second);
  return out;
}

std::map<std::string, int> ConsistentHash::getDistribution(const std::vector<std::string>& keys) const {
  std::map<std::string, int> counts;
  for (const auto& k : keys) ++counts[getNode(k)];
  return counts;
}
