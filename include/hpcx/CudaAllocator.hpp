// hpcx/CudaAllocator.hpp — the reference's CUDA allocator for STL containers, over the C ABI.
//
// Reference: lib/gpu/util/include/cudaAllocator.hpp:9-26 (strategies CudaMemGPUOnly /
// CudaMemManaged), :39-83 (CudaAllocator<_Ty, _Strategy>), used as
//   std::vector<uint32_t, CudaAllocator<uint32_t, CudaMemManaged>>
// to hand buffers to kernels (lib/gpu/radix_gpu/include/radix_sort_gpu.h:22-23).  Same names,
// same template shape and the same exceptions (std::bad_array_new_length, std::bad_alloc), so
// code written against the reference compiles against this header unchanged.  Differences, all
// deliberate: the strategies call hpcx_mem_alloc / hpcx_mem_free (libhpcx.so) instead of the CUDA
// runtime, so a host translation unit needs neither nvcc nor <cuda_runtime.h>; a failure other
// than out-of-memory throws std::system_error instead of being ignored (the reference's
// CHECK_CUDA is commented out, :68); and deallocate() frees (the reference's is a no-op, :71-73).
//
// A vector built on CudaMemManaged is what the library's entry points call HPCX_MEM_MANAGED:
//   std::vector<double, CudaAllocator<double, CudaMemManaged>> x(n), y(n);
//   hpcx_spmv(A, x.data(), HPCX_MEM_MANAGED, y.data(), HPCX_MEM_MANAGED, nullptr);
#pragma once
#ifndef HPCX_CUDA_ALLOCATOR_HPP_
#define HPCX_CUDA_ALLOCATOR_HPP_

#include <cstddef>
#include <limits>
#include <new>
#include <type_traits>
#include <utility>

#include "Runtime.hpp"

struct CudaMemStrategy {
  virtual hpcx_status malloc(void **ptr, std::size_t size) = 0;
  virtual hpcx_status free(void *ptr) = 0;
  virtual ~CudaMemStrategy() = default;
};

struct CudaMemGPUOnly : public CudaMemStrategy {
  static constexpr int32_t mem = HPCX_MEM_DEVICE;
  hpcx_status malloc(void **ptr, std::size_t size) override { return hpcx_mem_alloc(mem, size, ptr); }
  hpcx_status free(void *ptr) override { return hpcx_mem_free(mem, ptr); }
};

struct CudaMemManaged : public CudaMemStrategy {
  static constexpr int32_t mem = HPCX_MEM_MANAGED;
  hpcx_status malloc(void **ptr, std::size_t size) override { return hpcx_mem_alloc(mem, size, ptr); }
  hpcx_status free(void *ptr) override { return hpcx_mem_free(mem, ptr); }
};

// Not in the reference: page-locked host memory, what the copy pipelines of hpcx_spmv /
// hpcx_grid_upload need for their transfers to be asynchronous.
struct CudaMemPinned : public CudaMemStrategy {
  static constexpr int32_t mem = HPCX_MEM_HOST;
  hpcx_status malloc(void **ptr, std::size_t size) override { return hpcx_mem_alloc(mem, size, ptr); }
  hpcx_status free(void *ptr) override { return hpcx_mem_free(mem, ptr); }
};

template <typename _Strategy, class = void>
struct has_malloc_and_free : std::false_type {};
template <typename _Strategy>
struct has_malloc_and_free<
    _Strategy, std::void_t<decltype(std::declval<std::decay_t<_Strategy>>().malloc((void **)nullptr,
                                                                                   std::declval<std::size_t>())),
                           decltype(std::declval<std::decay_t<_Strategy>>().free((void *)nullptr))>>
    : std::true_type {};

template <typename _Ty, typename _Strategy>
struct CudaAllocator {
  static_assert(has_malloc_and_free<_Strategy>::value,
                "Strategy must provide malloc(void**, size_t) and free(void*)");
  using value_type = _Ty;
  using pointer = _Ty *;
  using const_pointer = const _Ty *;
  using size_type = std::size_t;
  using difference_type = std::ptrdiff_t;
  using propagate_on_container_move_assignment = std::true_type;
  using is_always_equal = std::false_type;
  template <class U>
  struct rebind {
    using other = CudaAllocator<U, _Strategy>;
  };

  explicit CudaAllocator(_Strategy s = {}) noexcept : strat_(std::move(s)) {}
  template <class U>
  constexpr CudaAllocator(const CudaAllocator<U, _Strategy> &o) noexcept : strat_(o.strategy()) {}

  _Ty *allocate(std::size_t count) {
    if (sizeof(_Ty) > 1 && count > std::numeric_limits<std::size_t>::max() / sizeof(_Ty))
      throw std::bad_array_new_length{};
    void *ptr = nullptr;
    const hpcx_status st = strat_.malloc(&ptr, count * sizeof(_Ty));
    if (st != HPCX_OK) hpcx::throwStatus(st, "CudaAllocator::allocate", __FILE__, __LINE__);  // HPCX_OOM -> std::bad_alloc
    return static_cast<_Ty *>(ptr);
  }
  void deallocate(_Ty *ptr, std::size_t = 0) noexcept { (void)strat_.free(ptr); }
  // Trivial default construction is skipped, as in the reference (:74-78): a vector of PODs in
  // device-only memory must not be value-initialised from the host.
  template <typename... Args>
  void construct(_Ty *p, Args &&...args) {
    if constexpr (!(sizeof...(args) == 0 && std::is_trivial_v<_Ty>))
      ::new (reinterpret_cast<void *>(p)) _Ty(std::forward<Args>(args)...);
  }

  const _Strategy &strategy() const noexcept { return strat_; }

 private:
  _Strategy strat_;
};

template <class T, class U, typename _Strategy>
bool operator==(const CudaAllocator<T, _Strategy> &, const CudaAllocator<U, _Strategy> &) {
  return true;
}
template <class T, class U, typename _Strategy>
bool operator!=(const CudaAllocator<T, _Strategy> &, const CudaAllocator<U, _Strategy> &) {
  return false;
}

namespace hpcx {
// HPCX_MEM_* of the storage behind a standard container's allocator: vectors on CudaAllocator
// go to the kernels as they are (managed / device pointers), everything else is host memory that
// the library stages through its copy pipeline.
template <class Alloc>
struct memKindOf : std::integral_constant<int32_t, HPCX_MEM_HOST> {};
template <class T>
struct memKindOf<CudaAllocator<T, CudaMemManaged>> : std::integral_constant<int32_t, HPCX_MEM_MANAGED> {};
template <class T>
struct memKindOf<CudaAllocator<T, CudaMemGPUOnly>> : std::integral_constant<int32_t, HPCX_MEM_DEVICE> {};
}  // namespace hpcx

#endif  // HPCX_CUDA_ALLOCATOR_HPP_
