// ABI version, thread-local error text, cached device facts.
#include "ftk_host.h"

namespace ftk {

char* err_buf() {
  static thread_local char buf[512] = {0};
  return buf;
}

int get_device_info(DeviceInfo* out) {
  static DeviceInfo cache[64];
  static bool have[64] = {false};
  int dev = 0;
  FTK_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return fail(FTK_EINVAL, "device ordinal %d out of range", dev);
  if (!have[dev]) {
    DeviceInfo di;
    FTK_CUDA_CHECK(cudaDeviceGetAttribute(&di.sm_count, cudaDevAttrMultiProcessorCount, dev));
    FTK_CUDA_CHECK(cudaDeviceGetAttribute(&di.cc_major, cudaDevAttrComputeCapabilityMajor, dev));
    FTK_CUDA_CHECK(cudaDeviceGetAttribute(&di.cc_minor, cudaDevAttrComputeCapabilityMinor, dev));
    cache[dev] = di;  // benign race: every writer stores the same values
    have[dev] = true;
  }
  *out = cache[dev];
  return FTK_OK;
}

}  // namespace ftk

extern "C" int ftk_version(void) { return FTK_ABI_VERSION; }

extern "C" const char* ftk_last_error(void) { return ftk::err_buf(); }
