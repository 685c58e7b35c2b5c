// Test infrastructure only (oracle build). Minimal stand-in for
// <boost/dynamic_bitset.hpp>: Boost is not installed in this image and the
// reference (CNNS/src/nsg/index_nsg.cpp:7, CNNS/include/nsg/index_nsg.h:12)
// only needs construct(n, 0), operator[], reset() and size(). No arithmetic
// lives here; it only records which node ids a search has visited.
#pragma once
#include <algorithm>
#include <cstddef>
#include <vector>
namespace boost {
template <typename Block = unsigned long>
class dynamic_bitset {
 public:
  dynamic_bitset() {}
  dynamic_bitset(std::size_t n, unsigned long = 0) : bits_(n, 0) {}
  struct ref {
    unsigned char &b;
    ref &operator=(bool v) { b = v ? 1 : 0; return *this; }
    ref &operator=(int v) { b = v ? 1 : 0; return *this; }
    operator bool() const { return b != 0; }
    bool operator==(bool v) const { return (b != 0) == v; }
  };
  ref operator[](std::size_t i) { return ref{bits_[i]}; }
  bool operator[](std::size_t i) const { return bits_[i] != 0; }
  void reset() { std::fill(bits_.begin(), bits_.end(), 0); }
  std::size_t size() const { return bits_.size(); }
 private:
  std::vector<unsigned char> bits_;
};
}  // namespace boost
