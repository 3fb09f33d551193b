// hpcx/AlignedAlloc.hpp — hpc::AlignedAllocator<T, Align>, the aligned STL allocator every
// libHPC grid container stores its flat buffer in.
//
// Drop-in for the reference's lib/hpc/include/AlignedAlloc.hpp:29-158 (same class name,
// template parameters, default Align = 64, member typedefs, rebind, allocate/deallocate/
// construct/destroy, and the "equal iff same alignment" comparison of :148-158), written
// fresh for this library.  Two defects of the reference are not reproduced:
//   * its detail:: helpers are non-inline definitions in a header (:13-26), an ODR violation
//     as soon as two translation units include it — here they are inline;
//   * std::aligned_alloc formally requires size % align == 0 — the request is rounded up.
// Allocation failure throws std::bad_alloc exactly like the reference (:78-80).
#pragma once
#ifndef HPCX_ALIGNED_ALLOC_HPP_
#define HPCX_ALIGNED_ALLOC_HPP_

#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <memory>
#include <new>
#include <type_traits>
#include <utility>

namespace hpc {
namespace detail {

inline void *allocate_aligned_memory(std::size_t align, std::size_t size) {
  if (align < sizeof(void *)) align = sizeof(void *);
  const std::size_t padded = (size + align - 1) / align * align;
#ifdef _MSC_VER
  return _aligned_malloc(padded ? padded : align, align);
#else
  return std::aligned_alloc(align, padded ? padded : align);
#endif
}

inline void deallocate_aligned_memory(void *ptr) noexcept {
#ifdef _MSC_VER
  _aligned_free(ptr);
#else
  std::free(ptr);
#endif
}

// Members shared by the T and const T flavours of the allocator.
template <typename Value, typename Pointer, std::size_t Align>
struct aligned_allocator_core {
  using value_type = Value;
  using pointer = Pointer;
  using const_pointer = const Value *;
  using const_reference = const Value &;
  using size_type = std::size_t;
  using difference_type = std::ptrdiff_t;
  using propagate_on_container_move_assignment = std::true_type;

  size_type max_size() const noexcept {
    return (~size_type(0) - size_type(Align)) / sizeof(Value);
  }
  const_pointer address(const_reference x) const noexcept { return std::addressof(x); }

  pointer allocate(size_type n, const void * /*hint*/ = nullptr) {
    void *p = allocate_aligned_memory(Align, n * sizeof(Value));
    if (!p) throw std::bad_alloc();
    return static_cast<pointer>(p);
  }
  void deallocate(pointer p, size_type) noexcept {
    deallocate_aligned_memory(const_cast<void *>(static_cast<const void *>(p)));
  }
  template <class U, class... Args>
  void construct(U *p, Args &&...args) {
    ::new (static_cast<void *>(p)) U(std::forward<Args>(args)...);
  }
  void destroy(pointer p) { p->~Value(); }
};

}  // namespace detail

template <typename T, std::size_t Align = 64>
class AlignedAllocator : public detail::aligned_allocator_core<T, T *, Align> {
  static_assert((Align & (Align - 1)) == 0, "Align must be a power of two");

 public:
  using reference = T &;
  template <class U>
  struct rebind {
    using other = AlignedAllocator<U, Align>;
  };
  AlignedAllocator() noexcept = default;
  template <class U>
  AlignedAllocator(const AlignedAllocator<U, Align> &) noexcept {}
  using detail::aligned_allocator_core<T, T *, Align>::address;
  T *address(reference x) const noexcept { return std::addressof(x); }
};

template <std::size_t Align>
class AlignedAllocator<void, Align> {
 public:
  using pointer = void *;
  using const_pointer = const void *;
  using value_type = void;
  template <class U>
  struct rebind {
    using other = AlignedAllocator<U, Align>;
  };
};

template <typename T, std::size_t Align>
class AlignedAllocator<const T, Align> : public detail::aligned_allocator_core<T, const T *, Align> {
 public:
  using reference = const T &;
  template <class U>
  struct rebind {
    using other = AlignedAllocator<U, Align>;
  };
  AlignedAllocator() noexcept = default;
  template <class U>
  AlignedAllocator(const AlignedAllocator<U, Align> &) noexcept {}
};

template <typename T, std::size_t TAlign, typename U, std::size_t UAlign>
inline bool operator==(const AlignedAllocator<T, TAlign> &, const AlignedAllocator<U, UAlign> &) noexcept {
  return TAlign == UAlign;
}
template <typename T, std::size_t TAlign, typename U, std::size_t UAlign>
inline bool operator!=(const AlignedAllocator<T, TAlign> &, const AlignedAllocator<U, UAlign> &) noexcept {
  return TAlign != UAlign;
}

}  // namespace hpc
#endif  // HPCX_ALIGNED_ALLOC_HPP_
