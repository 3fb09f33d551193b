// hpcx/HPCHighDimensionFlatArray.hpp — hpc::HPCHighDimensionFlatArray, the ghost-padded dense
// N-D grid container of libHPC, as the host side of the libhpcx stencil path.
//
// Drop-in for the reference's lib/hpc/include/HPCHighDimensionFlatArray.hpp:54-224: same
// template signature <Dimension, _Ty, Low_Bound, High_Bound = Low_Bound, Alignment = 16,
// Alloc = AlignedAllocator<_Ty, Alignment>>, same public members (variadic ctor :76-83,
// shrink_to_fit :88, data :93-98, at :107-109, operator() :118-125), same protected helpers
// (resize :135-144, padded_index :151-153, compute_stride_and_total :161-171,
// unsafe_linearize :180-187, safe_linearize :197-208) and therefore the same flat layout:
//     stride[i] = prod_{j>i}(dim[j] + Low_Bound + High_Bound)
//     offset(idx) = sum_i stride[i] * (idx[i] + Low_Bound),   idx[i] in [-Low_Bound, dim[i]+High_Bound)
// The buffer is value-initialised, so ghost cells start at zero.  tests/test_cpp_host.py
// compares strides, totals, offsets and the out_of_range behaviour of at() with the
// reference header compiled separately (oracle/_ref/libhpcref.so), index by index.
//
// Additions (do not change the layout): const overloads of at()/operator(), dim()/stride()/
// size() accessors and the low_bound/high_bound constants, which the device mirror needs to
// describe the array across the C ABI (hpcx_grid_desc in hpcx.h).
#pragma once
#ifndef HPCX_HIGH_DIMENSION_FLAT_ARRAY_HPP_
#define HPCX_HIGH_DIMENSION_FLAT_ARRAY_HPP_

#include <array>
#include <cassert>
#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <type_traits>
#include <utility>
#include <vector>

#include "AlignedAlloc.hpp"

namespace hpc {

template <std::size_t Dimension, typename _Ty, std::size_t Low_Bound = 0,
          std::size_t High_Bound = Low_Bound, std::size_t Alignment = 16,
          class Alloc = AlignedAllocator<_Ty, Alignment>>
class HPCHighDimensionFlatArray {
  static_assert(Dimension > 0, "Dimension must larger than zero");
  static_assert(std::is_same_v<std::remove_cv_t<std::remove_reference_t<_Ty>>, _Ty>,
                "_Ty must not be cvref");

 public:
  using value_type = _Ty;
  using extent_type = std::array<std::size_t, Dimension>;
  using index_type = std::array<std::intptr_t, Dimension>;
  static constexpr std::size_t dimension = Dimension;
  static constexpr std::size_t low_bound = Low_Bound;
  static constexpr std::size_t high_bound = High_Bound;

  // a(ny, nx): logical sizes, slowest dimension first; ghost cells are added on both sides.
  template <typename... DimForEachLayer,
            std::enable_if_t<(sizeof...(DimForEachLayer) == Dimension &&
                              std::conjunction_v<std::is_integral<DimForEachLayer>...>),
                             int> = 0>
  explicit HPCHighDimensionFlatArray(const DimForEachLayer &...dims)
      : HPCHighDimensionFlatArray(extent_type{static_cast<std::size_t>(dims)...}) {}

  void shrink_to_fit() { _flat.shrink_to_fit(); }

  constexpr _Ty *data() noexcept { return _flat.data(); }
  constexpr const _Ty *data() const noexcept { return _flat.data(); }

  // Bounds-checked access over the padded index range; throws std::out_of_range outside it.
  _Ty &at(const index_type &indices) { return _flat[safe_linearize(indices)]; }
  const _Ty &at(const index_type &indices) const { return _flat[safe_linearize(indices)]; }

  // Unchecked access; ghost cells are reached with negative / >= dim indices.
  template <typename... Indicies,
            std::enable_if_t<(sizeof...(Indicies) == Dimension &&
                              std::conjunction_v<std::is_integral<Indicies>...>),
                             int> = 0>
  _Ty &operator()(const Indicies &...idxs) noexcept {
    return _flat[unsafe_linearize({static_cast<std::intptr_t>(idxs)...})];
  }
  template <typename... Indicies,
            std::enable_if_t<(sizeof...(Indicies) == Dimension &&
                              std::conjunction_v<std::is_integral<Indicies>...>),
                             int> = 0>
  const _Ty &operator()(const Indicies &...idxs) const noexcept {
    return _flat[unsafe_linearize({static_cast<std::intptr_t>(idxs)...})];
  }

  // ---- additions
  const extent_type &dim() const noexcept { return _dim; }        // logical extents
  const extent_type &stride() const noexcept { return _stride; }  // padded row-major strides
  std::size_t size() const noexcept { return _flat.size(); }      // padded element count

 protected:
  void resize(const extent_type &dim, const _Ty &value = _Ty{}) {
    const auto layout = compute_stride_and_total(dim);
    assert(layout.second > 0 && "resize() attempted to create zero-sized flat array!");
    _dim = dim;
    _stride = layout.first;
    _flat.assign(layout.second, value);
  }

  std::intptr_t padded_index(std::intptr_t val) const noexcept {
    return val + static_cast<std::intptr_t>(Low_Bound);
  }

  static constexpr auto compute_stride_and_total(const extent_type &dims) noexcept
      -> std::pair<extent_type, std::size_t> {
    extent_type stride = {};
    std::size_t running = 1;
    for (std::size_t k = 0; k < Dimension; ++k) {
      const std::size_t i = Dimension - 1 - k;  // fastest dimension first
      stride[i] = running;
      running *= dims[i] + Low_Bound + High_Bound;
    }
    return {stride, running};
  }

  std::intptr_t unsafe_linearize(const index_type &insert) const noexcept {
    std::intptr_t offset = 0;
    for (std::size_t i = 0; i < Dimension; ++i)
      offset += static_cast<std::intptr_t>(_stride[i]) * padded_index(insert[i]);
    return offset;
  }

  std::intptr_t safe_linearize(const index_type &insert) const {
    for (std::size_t i = 0; i < Dimension; ++i) {
      const std::intptr_t lowest = -static_cast<std::intptr_t>(Low_Bound);
      const std::intptr_t end = static_cast<std::intptr_t>(_dim[i] + High_Bound);
      if (insert[i] < lowest || insert[i] >= end)
        throw std::out_of_range("invalid index, out of boundary");
    }
    return unsafe_linearize(insert);
  }

 private:
  std::vector<_Ty, Alloc> _flat;
  extent_type _dim{};
  extent_type _stride{};

  explicit HPCHighDimensionFlatArray(const extent_type &extents) { resize(extents); }
};

}  // namespace hpc
#endif  // HPCX_HIGH_DIMENSION_FLAT_ARRAY_HPP_
