// Shim for the reference arm: PADDLE_ENFORCE_* as abort-with-message, the error factory
// functions as printf-style string builders, and gflags DECLARE_* as plain extern globals
// (defined in baseline/harness/ref_shim.cc; FLAGS_<name> environment variables are honoured
// there like gflags would).
#pragma once
#include <cuda_runtime.h>
#include <glog/logging.h>
#include <cstdint>
#include <cstdio>
#include <string>
#include "paddle/fluid/platform/device/gpu/gpu_types.h"

namespace paddle {
namespace platform {
namespace errors {
template <typename... Args>
inline std::string ref_fmt(const char* fmt, Args... args) {
  char buf[512];
  snprintf(buf, sizeof(buf), fmt, args...);
  return std::string(buf);
}
inline std::string ref_fmt(const char* fmt) { return std::string(fmt); }
#define REF_ERR(name)                                  \
  template <typename... Args>                          \
  inline std::string name(const char* fmt, Args... a) { return ref_fmt(fmt, a...); }
REF_ERR(External)
REF_ERR(InvalidArgument)
REF_ERR(PreconditionNotMet)
REF_ERR(Unimplemented)
REF_ERR(Unavailable)
REF_ERR(Fatal)
REF_ERR(NotFound)
REF_ERR(OutOfRange)
#undef REF_ERR
}  // namespace errors
}  // namespace platform
}  // namespace paddle

#define PADDLE_ENFORCE_GPU_SUCCESS(expr)                                                   \
  do {                                                                                     \
    auto ref_e__ = (expr);                                                                 \
    if (ref_e__ != cudaSuccess) {                                                          \
      LOG(FATAL) << "CUDA error " << (int)ref_e__ << " (" << cudaGetErrorString((cudaError_t)ref_e__) \
                 << ") in " #expr;                                                         \
    }                                                                                      \
  } while (0)
#define REF_ENFORCE_CMP(a, b, op, msg)                                          \
  do {                                                                          \
    if (!((a)op(b))) { LOG(FATAL) << "enforce " #a " " #op " " #b " failed: " << (msg); } \
  } while (0)
#define PADDLE_ENFORCE_EQ(a, b, msg) REF_ENFORCE_CMP(a, b, ==, msg)
#define PADDLE_ENFORCE_NE(a, b, msg) REF_ENFORCE_CMP(a, b, !=, msg)
#define PADDLE_ENFORCE_GT(a, b, msg) REF_ENFORCE_CMP(a, b, >, msg)
#define PADDLE_ENFORCE_GE(a, b, msg) REF_ENFORCE_CMP(a, b, >=, msg)
#define PADDLE_ENFORCE_LT(a, b, msg) REF_ENFORCE_CMP(a, b, <, msg)
#define PADDLE_ENFORCE_LE(a, b, msg) REF_ENFORCE_CMP(a, b, <=, msg)
#define PADDLE_THROW(msg) LOG(FATAL) << (msg)

#define DECLARE_bool(name) extern bool FLAGS_##name
#define DECLARE_int32(name) extern int32_t FLAGS_##name
#define DECLARE_uint64(name) extern uint64_t FLAGS_##name
#define DECLARE_double(name) extern double FLAGS_##name
