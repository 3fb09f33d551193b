// TEST INFRASTRUCTURE ONLY — a stand-in for oneTBB's header so that the REFERENCE's lib/sparse
// headers (DenseBlock.hpp:6,72) compile in this container, where oneTBB (a FetchContent download of
// the reference's build, cmake/libHPC_preset.cmake:84-90) is absent.  Only what those headers use.
// Not thread-safe growth: the reference only resizes in the constructor (DenseBlock.hpp:22).
#pragma once
#include <vector>
namespace tbb {
template <typename T>
struct concurrent_vector {
  // std::vector<bool> is a bitset without element references; the reference stores bool leaves
  using Store = std::vector<std::conditional_t<std::is_same_v<T, bool>, unsigned char, T>>;
  void resize(std::size_t n) { v.resize(n); }
  std::size_t size() const { return v.size(); }
  auto &operator[](std::size_t i) { return reinterpret_cast<T &>(v[i]); }
  const auto &operator[](std::size_t i) const { return reinterpret_cast<const T &>(v[i]); }
  Store v;
};
}  // namespace tbb
