// ABI version, thread-local error text, cached device facts.
#include <atomic>

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

namespace {
std::atomic<int> g_options[FTK_OPT_COUNT_] = {{2}, {0}, {0}, {0}};  // cta_group 2, gather auto/auto, no SM reserve
}

int option(int key) { return g_options[key].load(std::memory_order_relaxed); }

}  // namespace ftk

extern "C" int ftk_set_option(int32_t key, int32_t value) {
  switch (key) {
    case FTK_OPT_CTA_GROUP:
      FTK_REQUIRE(value == 1 || value == 2, "ftk_set_option: cta_group must be 1 or 2, got %d", value);
      break;
    case FTK_OPT_GATHER_VARIANT:
      FTK_REQUIRE(value == 0 || value == 1, "ftk_set_option: gather variant must be 0 (auto) or 1 (direct), got %d", value);
      break;
    case FTK_OPT_GATHER_STORES:
      FTK_REQUIRE(value >= 0 && value <= 2, "ftk_set_option: gather stores must be 0 (auto), 1 (streaming) or 2 (plain), got %d",
                  value);
      break;
    case FTK_OPT_SM_RESERVE:
      FTK_REQUIRE(value >= 0 && value <= 64, "ftk_set_option: SM reserve %d out of range [0, 64]", value);
      break;
    default:
      return ftk::fail(FTK_EINVAL, "ftk_set_option: unknown key %d", key);
  }
  ftk::g_options[key].store(value, std::memory_order_relaxed);
  return FTK_OK;
}

extern "C" int ftk_set_sm_reserve(int32_t n_sms) { return ftk_set_option(FTK_OPT_SM_RESERVE, n_sms); }

extern "C" int ftk_version(void) { return FTK_ABI_VERSION; }

extern "C" const char* ftk_last_error(void) { return ftk::err_buf(); }
