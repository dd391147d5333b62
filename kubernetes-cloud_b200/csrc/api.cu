// ABI version, thread-local error text, cached device facts.
#include <atomic>
#include <mutex>
#include <vector>

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

// ---- TMA tensor maps (shared by the tcgen05 translation units)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int get_encode_fn(EncodeTiledFn* out) {
  static EncodeTiledFn cached = nullptr;
  if (cached == nullptr) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    FTK_CUDA_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (qres != cudaDriverEntryPointSuccess || fn == nullptr)
      return fail(FTK_EUNSUPPORTED, "cuTensorMapEncodeTiled not available from the driver");
    cached = reinterpret_cast<EncodeTiledFn>(fn);
  }
  *out = cached;
  return FTK_OK;
}

// bf16 row-major [dim1][dim0] with row pitch pitch_bytes; box = box0 x box1 elements; 128-byte swizzle.
// Out-of-bounds parts of a box are filled with zeros.
int make_tmap(CUtensorMap* tm, const void* base, uint64_t dim0, uint64_t dim1, uint64_t pitch_bytes, uint32_t box0,
              uint32_t box1) {
  EncodeTiledFn enc = nullptr;
  if (int rc = get_encode_fn(&enc)) return rc;
  // cuTensorMapEncodeTiled is a driver call and needs a context current on THIS thread; a thread that has only ever
  // been handed cached allocations (the autograd engine's thread) may not have one yet.  cudaFree(0) binds the
  // device's primary context; once per thread.
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    FTK_CUDA_CHECK(cudaFree(nullptr));
    ctx_bound = true;
  }
  FTK_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "tensor base %p is not 16-byte aligned", base);
  FTK_REQUIRE(pitch_bytes % 16 == 0, "row pitch %llu is not a multiple of 16 bytes",
              static_cast<unsigned long long>(pitch_bytes));
  FTK_REQUIRE(box0 * 2 <= 128 && box1 <= 256, "bad TMA box %u x %u", box0, box1);
  const cuuint64_t dims[2] = {dim0, dim1};
  const cuuint64_t strides[1] = {pitch_bytes};
  const cuuint32_t box[2] = {box0, box1};
  const cuuint32_t estr[2] = {1, 1};
  const CUtensorMapL2promotion promo = CU_TENSOR_MAP_L2_PROMOTION_L2_256B;  // 64 B / 128 B / none measured the same
  const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, promo,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(FTK_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", static_cast<int>(r));
  return FTK_OK;
}

namespace {
std::atomic<int> g_options[FTK_OPT_COUNT_] = {{2}, {0}, {0}, {0}};  // cta_group 2, gather auto/auto, no SM reserve
}

int option(int key) { return g_options[key].load(std::memory_order_relaxed); }

// ---- launch timeline
namespace {
struct TimelineEntry {
  int tag;
  cudaEvent_t a, b;
};
std::atomic<int> g_tl_on{0};
std::mutex g_tl_mu;
std::vector<TimelineEntry> g_tl;
constexpr size_t kTimelineCap = 1 << 16;
}  // namespace

bool timeline_on() { return g_tl_on.load(std::memory_order_relaxed) != 0; }

void timeline_mark(int tag, cudaStream_t st, bool begin) {
  std::lock_guard<std::mutex> lk(g_tl_mu);
  if (begin) {
    if (g_tl.size() >= kTimelineCap) return;
    TimelineEntry e{tag, nullptr, nullptr};
    if (cudaEventCreate(&e.a) != cudaSuccess || cudaEventCreate(&e.b) != cudaSuccess) return;
    cudaEventRecord(e.a, st);
    g_tl.push_back(e);
  } else {
    // the matching begin is the latest open entry with this tag
    for (size_t i = g_tl.size(); i-- > 0;)
      if (g_tl[i].tag == tag) {
        cudaEventRecord(g_tl[i].b, st);
        break;
      }
  }
}

}  // namespace ftk

extern "C" int ftk_set_option(int32_t key, int32_t value) {
  switch (key) {
    case FTK_OPT_CTA_GROUP:
      FTK_REQUIRE(value == 1 || value == 2, "ftk_set_option: cta_group must be 1 or 2, got %d", value);
      break;
    case FTK_OPT_GATHER_VARIANT:
      FTK_REQUIRE(value >= 0 && value <= 2, "ftk_set_option: gather variant must be 0 (auto), 1 (direct) or 2 (staged), got %d",
                  value);
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

extern "C" int ftk_timeline_enable(int32_t on) {
  std::lock_guard<std::mutex> lk(ftk::g_tl_mu);
  for (auto& e : ftk::g_tl) {
    cudaEventDestroy(e.a);
    cudaEventDestroy(e.b);
  }
  ftk::g_tl.clear();
  ftk::g_tl_on.store(on ? 1 : 0, std::memory_order_relaxed);
  return FTK_OK;
}

extern "C" int ftk_timeline_read(int32_t* tags, float* ms, float* start_ms, int32_t max_n) {
  FTK_REQUIRE(tags && ms && max_n >= 0, "ftk_timeline_read: bad arguments");
  std::lock_guard<std::mutex> lk(ftk::g_tl_mu);
  int n = 0;
  for (auto& e : ftk::g_tl) {
    if (n >= max_n) break;
    FTK_CUDA_CHECK(cudaEventSynchronize(e.b));
    float t = 0.f;
    FTK_CUDA_CHECK(cudaEventElapsedTime(&t, e.a, e.b));
    tags[n] = e.tag;
    ms[n] = t;
    if (start_ms != nullptr) {
      float s0 = 0.f;
      FTK_CUDA_CHECK(cudaEventElapsedTime(&s0, ftk::g_tl.front().a, e.a));  // works across streams
      start_ms[n] = s0;
    }
    ++n;
  }
  return n;
}

extern "C" int ftk_set_sm_reserve(int32_t n_sms) { return ftk_set_option(FTK_OPT_SM_RESERVE, n_sms); }

extern "C" int ftk_version(void) { return FTK_ABI_VERSION; }

extern "C" const char* ftk_last_error(void) { return ftk::err_buf(); }
