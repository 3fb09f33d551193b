// TEST INFRASTRUCTURE ONLY — stand-in for tbb::parallel_for(first, last, f) (HashBlock.hpp:113): serial.
#pragma once
namespace tbb {
template <typename Index, typename F>
void parallel_for(Index first, Index last, const F &f) {
  for (Index i = first; i < last; ++i) f(i);
}
}  // namespace tbb
