// hpcx/Runtime.hpp — C++ face of the libhpcx context and of its error convention.
//
// The C ABI (hpcx.h) never throws; this header turns its status codes back into the
// exceptions the reference's callers already handle:
//   HPCX_OUT_OF_RANGE -> std::out_of_range   (HPCHighDimensionFlatArray.hpp:201-203)
//   HPCX_OOM          -> std::bad_alloc      (AlignedAlloc.hpp:78-80, cudaAllocator.hpp:59-67)
//   anything else     -> std::system_error with an error category and a "file:line" prefix,
//                        the shape of cudahelper::throwCudaError
//                        (lib/gpu/util/include/cudaHelper.cuh:10-27) and checkCuda (:96-101).
// There is no CPU fallback behind any of these calls: without libhpcx.so the program does
// not link, and without an sm_100 device hpcx::Context's constructor throws.
#pragma once
#ifndef HPCX_RUNTIME_HPP_
#define HPCX_RUNTIME_HPP_

#include <cstddef>
#include <cstdint>
#include <new>
#include <stdexcept>
#include <string>
#include <system_error>
#include <type_traits>

#include "../hpcx.h"

namespace hpcx {

inline const std::error_category &errorCategory() noexcept {
  struct Category final : std::error_category {
    const char *name() const noexcept override { return "hpcx"; }
    std::string message(int v) const override {
      switch (v) {
        case HPCX_OK: return "ok";
        case HPCX_INVALID_ARG: return "invalid argument";
        case HPCX_OOM: return "out of memory";
        case HPCX_CUDA_ERROR: return "CUDA error";
        case HPCX_NCCL_ERROR: return "NCCL error";
        case HPCX_UNSUPPORTED: return "unsupported";
        case HPCX_OUT_OF_RANGE: return "out of range";
        default: return "unknown hpcx status";
      }
    }
  };
  static Category instance;
  return instance;
}

[[noreturn]] inline void throwStatus(hpcx_status st, const char *detail, const char *file, int line) {
  const std::string where = std::string(file ? file : "[??]") + ":" + std::to_string(line);
  const std::string what = detail && *detail ? where + ": " + detail : where;
  if (st == HPCX_OUT_OF_RANGE) throw std::out_of_range(detail && *detail ? detail : "invalid index, out of boundary");
  if (st == HPCX_OOM) throw std::bad_alloc();
  throw std::system_error(std::error_code(static_cast<int>(st), errorCategory()), what);
}

inline hpcx_status checkStatus(hpcx_ctx ctx, hpcx_status st, const char *file, int line) {
  if (st != HPCX_OK) throwStatus(st, ctx ? hpcx_last_error_string(ctx) : "", file, line);
  return st;
}
#define HPCX_CHECK(ctx, expr) ::hpcx::checkStatus((ctx), (expr), __FILE__, __LINE__)

template <typename T>
constexpr int32_t dtypeOf() {
  static_assert(std::is_same_v<T, float> || std::is_same_v<T, double>,
                "libhpcx computes in float or double");
  return std::is_same_v<T, double> ? HPCX_F64 : HPCX_F32;
}

// One context per process and device (one process per GPU, hpcx.h).
class Context {
 public:
  explicit Context(int device = 0) {
    const hpcx_status st = hpcx_init(device, &ctx_);
    if (st != HPCX_OK) {
      const std::string msg = ctx_ ? hpcx_last_error_string(ctx_) : "hpcx_init failed";
      if (ctx_) hpcx_finalize(ctx_);
      ctx_ = nullptr;
      throwStatus(st, msg.c_str(), __FILE__, __LINE__);
    }
  }
  ~Context() {
    if (ctx_) {
      hpcx_comm_finalize(ctx_);
      hpcx_finalize(ctx_);
    }
  }
  Context(const Context &) = delete;
  Context &operator=(const Context &) = delete;

  hpcx_ctx handle() const noexcept { return ctx_; }
  void synchronize() { HPCX_CHECK(ctx_, hpcx_synchronize(ctx_)); }
  std::int64_t launchCount() const noexcept { return hpcx_launch_count(ctx_); }

 private:
  hpcx_ctx ctx_ = nullptr;
};

}  // namespace hpcx
#endif  // HPCX_RUNTIME_HPP_
