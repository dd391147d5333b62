// Host-side helpers shared by the C-ABI translation units: thread-local error text, CUDA checks.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>

#include "../../include/ftk.h"

namespace ftk {

char* err_buf();  // thread-local, 512 bytes (defined in api.cu)

inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(err_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}

#define FTK_CUDA_CHECK(expr)                                                                              \
  do {                                                                                                    \
    cudaError_t e__ = (expr);                                                                             \
    if (e__ != cudaSuccess)                                                                               \
      return ::ftk::fail(FTK_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
  } while (0)

#define FTK_REQUIRE(cond, ...)                              \
  do {                                                      \
    if (!(cond)) return ::ftk::fail(FTK_EINVAL, __VA_ARGS__); \
  } while (0)

// Cached per-device facts (SM count, compute capability).
struct DeviceInfo {
  int sm_count;
  int cc_major;
  int cc_minor;
};
int get_device_info(DeviceInfo* out);  // current device; FTK_OK or error

// bf16 row-major [dim1][dim0] tensor map with row pitch pitch_bytes; box = box0 x box1 elements; 128-byte swizzle;
// out-of-bounds parts of a box read as zeros.  FTK_OK or an error code.
int make_tmap(CUtensorMap* tm, const void* base, uint64_t dim0, uint64_t dim1, uint64_t pitch_bytes, uint32_t box0,
              uint32_t box1);

// Launch timeline (ftk_timeline_*): when enabled, a pair of CUDA events brackets every tagged launch on its own
// stream.  TimelineScope records the first event at construction and the second at destruction.
bool timeline_on();
void timeline_mark(int tag, cudaStream_t st, bool begin);
struct TimelineScope {
  int tag;
  cudaStream_t st;
  bool on;
  TimelineScope(int t, cudaStream_t s) : tag(t), st(s), on(timeline_on()) {
    if (on) timeline_mark(tag, st, true);
  }
  ~TimelineScope() {
    if (on) timeline_mark(tag, st, false);
  }
};

// Process-wide tuning options (ftk_set_option): read with a relaxed atomic load on every launch, never from the
// environment.
int option(int key);

}  // namespace ftk
