// Consistent-hash ring mapping request keys onto GPU workers.
//
// Same public surface and bit-exact placement as the reference's ConsistentHash
// (reference: include/consistent_hash.h:10-24, src/consistent_hash.cpp:6-70): FNV-1a/32 with each
// `char` sign-extended before the XOR, 150 virtual nodes "node#i" per node by default, lookup =
// first ring position >= hash(key) with wrap-around, a later insert taking over an equal position.
// Storage differs: a sorted flat array searched by bisection instead of a red-black tree, so a
// lookup touches one contiguous block (the ring is read on every routed request and almost never
// written).
#pragma once
#include <cstdint>
#include <map>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

class ConsistentHash {
 public:
  explicit ConsistentHash(int virtual_nodes = 150) : vnodes_(virtual_nodes) {}

  void addNode(const std::string& node);
  void removeNode(const std::string& node);
  std::string getNode(const std::string& key) const;            // "" when the ring is empty
  std::vector<std::string> getAllNodes() const;                  // distinct nodes in ring order
  std::map<std::string, int> getDistribution(const std::vector<std::string>& keys) const;

  static uint32_t hashKey(const char* data, size_t len);        // the ring's hash function

 private:
  size_t lowerBound(uint32_t h) const;
  int vnodes_;
  std::vector<std::pair<uint32_t, std::string>> ring_;           // sorted by hash, unique hashes
  mutable std::mutex mu_;
};
