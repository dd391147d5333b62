// G4: data-parallel gradient all-reduce over NCCL (NVLink 5 / NVSwitch), one communicator and one
// high-priority communication stream per rank. Replaces the bucketed gradient reduction DeepSpeed ZeRO performs
// for the reference (ds_config.json:27-45, enabled at FT.py:911-923).
//
// NCCL is resolved with dlopen at ftk_dp_init time (the copy PyTorch already loaded, or FTK_NCCL_PATH), so the
// library itself has no link-time dependency on it and loads on machines without NCCL.
#include <cuda_bf16.h>
#include <dlfcn.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "ftk_host.h"

namespace {

// Minimal NCCL declarations (stable ABI since 2.x); avoids needing nccl.h at build time.
typedef struct ncclComm* ncclComm_t;
typedef struct {
  char internal[128];
} ncclUniqueId;
typedef int ncclResult_t;
enum { ncclSum = 0, ncclAvg = 4 };
enum { ncclScalarHostImmediate = 1 };
enum { ncclFloat32 = 7, ncclBfloat16 = 9 };

struct NcclApi {
  void* handle;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*);
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
  ncclResult_t (*CommDestroy)(ncclComm_t);
  ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*RedOpCreatePreMulSum)(int*, void*, int, int, ncclComm_t);  // optional (NCCL >= 2.11)
  ncclResult_t (*RedOpDestroy)(int, ncclComm_t);
  const char* (*GetErrorString)(ncclResult_t);
};

int load_nccl(NcclApi** out) {
  static NcclApi api;
  static bool loaded = false;
  if (!loaded) {
    const char* cand[4] = {getenv("FTK_NCCL_PATH"), "libnccl.so.2", "libnccl.so", nullptr};
    void* h = nullptr;
    for (int i = 0; i < 3 && h == nullptr; ++i)
      if (cand[i] != nullptr && cand[i][0] != 0) h = dlopen(cand[i], RTLD_NOW | RTLD_GLOBAL);
    if (h == nullptr)
      return ftk::fail(FTK_ENCCL, "cannot dlopen NCCL (set FTK_NCCL_PATH to libnccl.so.2): %s", dlerror());
    api.handle = h;
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(dlsym(h, "ncclCommInitRank"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(dlsym(h, "ncclCommDestroy"));
    api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(dlsym(h, "ncclAllReduce"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
    api.RedOpCreatePreMulSum =
        reinterpret_cast<decltype(api.RedOpCreatePreMulSum)>(dlsym(h, "ncclRedOpCreatePreMulSum"));
    api.RedOpDestroy = reinterpret_cast<decltype(api.RedOpDestroy)>(dlsym(h, "ncclRedOpDestroy"));
    if (!api.GetUniqueId || !api.CommInitRank || !api.CommDestroy || !api.AllReduce || !api.GetErrorString)
      return ftk::fail(FTK_ENCCL, "NCCL library lacks a required symbol");
    loaded = true;
  }
  *out = &api;
  return FTK_OK;
}

#define FTK_NCCL_CHECK(api, expr)                                                                   \
  do {                                                                                              \
    ncclResult_t r__ = (expr);                                                                      \
    if (r__ != 0) return ftk::fail(FTK_ENCCL, "%s failed: %s", #expr, (api)->GetErrorString(r__)); \
  } while (0)

// In-place scale before the sum (the reference averages gradients across ranks).
template <typename T>
__global__ void __launch_bounds__(256) prescale_kernel(T* buf, size_t n, float s);

template <>
__global__ void __launch_bounds__(256) prescale_kernel<float>(float* buf, size_t n, float s) {
  const size_t n4 = n / 4;
  float4* b4 = reinterpret_cast<float4*>(buf);
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    float4 v = b4[i];
    v.x *= s;
    v.y *= s;
    v.z *= s;
    v.w *= s;
    b4[i] = v;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) buf[n4 * 4 + threadIdx.x] *= s;
}

template <>
__global__ void __launch_bounds__(256) prescale_kernel<__nv_bfloat16>(__nv_bfloat16* buf, size_t n, float s) {
  const size_t n8 = n / 8;
  uint4* b8 = reinterpret_cast<uint4*>(buf);
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n8;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    uint4 v = b8[i];
    uint32_t* w = reinterpret_cast<uint32_t*>(&v);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      __nv_bfloat162 h = *reinterpret_cast<__nv_bfloat162*>(&w[e]);
      h = __floats2bfloat162_rn(__low2float(h) * s, __high2float(h) * s);
      w[e] = *reinterpret_cast<uint32_t*>(&h);
    }
    b8[i] = v;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 7))
    buf[n8 * 8 + threadIdx.x] = __float2bfloat16(__bfloat162float(buf[n8 * 8 + threadIdx.x]) * s);
}

}  // namespace

struct ftk_dp {
  NcclApi* api;
  ncclComm_t comm;
  cudaStream_t comm_stream;
  cudaEvent_t ev_ready;
  cudaEvent_t ev_done;
  int rank, world, sm_count;
};

static int launch_prescale(ftk_dp* dp, void* buf, size_t count, int32_t dtype, float prescale) {
  const size_t vec = count / (dtype == FTK_F32 ? 4 : 8);
  size_t blocks = (vec + 255) / 256;
  const size_t cap = static_cast<size_t>(dp->sm_count) * 8;
  if (blocks > cap) blocks = cap;
  if (blocks == 0) blocks = 1;
  if (dtype == FTK_F32)
    prescale_kernel<float><<<static_cast<unsigned>(blocks), 256, 0, dp->comm_stream>>>(static_cast<float*>(buf), count,
                                                                                     prescale);
  else
    prescale_kernel<__nv_bfloat16><<<static_cast<unsigned>(blocks), 256, 0, dp->comm_stream>>>(
        static_cast<__nv_bfloat16*>(buf), count, prescale);
  FTK_CUDA_CHECK(cudaGetLastError());
  return FTK_OK;
}

extern "C" int ftk_dp_unique_id(void* id128_host) {
  FTK_REQUIRE(id128_host != nullptr, "ftk_dp_unique_id: null output");
  NcclApi* api;
  if (int rc = load_nccl(&api)) return rc;
  ncclUniqueId id;
  FTK_NCCL_CHECK(api, api->GetUniqueId(&id));
  memcpy(id128_host, &id, 128);
  return FTK_OK;
}

extern "C" int ftk_dp_init(const void* id128_host, int32_t rank, int32_t world, ftk_dp_t** out) {
  FTK_REQUIRE(id128_host && out, "ftk_dp_init: null pointer argument");
  FTK_REQUIRE(world >= 1 && rank >= 0 && rank < world, "ftk_dp_init: bad rank %d / world %d", rank, world);
  NcclApi* api;
  if (int rc = load_nccl(&api)) return rc;
  ftk::DeviceInfo di;
  if (int rc = ftk::get_device_info(&di)) return rc;
  ftk_dp* dp = static_cast<ftk_dp*>(calloc(1, sizeof(ftk_dp)));
  if (dp == nullptr) return ftk::fail(FTK_EINVAL, "ftk_dp_init: out of host memory");
  dp->api = api;
  dp->rank = rank;
  dp->world = world;
  dp->sm_count = di.sm_count;
  ncclUniqueId id;
  memcpy(&id, id128_host, 128);
  ncclResult_t r = api->CommInitRank(&dp->comm, world, id, rank);
  if (r != 0) {
    free(dp);
    return ftk::fail(FTK_ENCCL, "ncclCommInitRank failed: %s", api->GetErrorString(r));
  }
  // from here on a failure must give back what was created: one exit path
  int lo = 0, hi = 0;
  cudaError_t e = cudaDeviceGetStreamPriorityRange(&lo, &hi);
  if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&dp->comm_stream, cudaStreamNonBlocking, hi);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&dp->ev_ready, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&dp->ev_done, cudaEventDisableTiming);
  if (e != cudaSuccess) {
    if (dp->ev_done != nullptr) cudaEventDestroy(dp->ev_done);
    if (dp->ev_ready != nullptr) cudaEventDestroy(dp->ev_ready);
    if (dp->comm_stream != nullptr) cudaStreamDestroy(dp->comm_stream);
    api->CommDestroy(dp->comm);
    free(dp);
    return ftk::fail(FTK_ECUDA, "ftk_dp_init: creating the communication stream / events failed: %s",
                     cudaGetErrorString(e));
  }
  *out = dp;
  return FTK_OK;
}

static int allreduce_on_comm_stream(ftk_dp_t* dp, void* buf, size_t count, int32_t dtype, float prescale);

extern "C" int ftk_dp_allreduce(ftk_dp_t* dp, void* buf, size_t count, int32_t dtype, float prescale,
                                void* ready_stream) {
  FTK_REQUIRE(dp && buf, "ftk_dp_allreduce: null pointer argument");
  FTK_REQUIRE(dtype == FTK_BF16 || dtype == FTK_F32, "ftk_dp_allreduce: bad dtype %d", dtype);
  FTK_REQUIRE((reinterpret_cast<uintptr_t>(buf) & 15) == 0, "ftk_dp_allreduce: buffer %p is not 16-byte aligned", buf);
  if (count == 0) return FTK_OK;
  // the bucket is ready once everything issued so far on ready_stream has run
  FTK_CUDA_CHECK(cudaEventRecord(dp->ev_ready, static_cast<cudaStream_t>(ready_stream)));
  FTK_CUDA_CHECK(cudaStreamWaitEvent(dp->comm_stream, dp->ev_ready, 0));
  return allreduce_on_comm_stream(dp, buf, count, dtype, prescale);
}

extern "C" int ftk_dp_allreduce_after(ftk_dp_t* dp, void* buf, size_t count, int32_t dtype, float prescale,
                                      void* ready_event) {
  FTK_REQUIRE(dp && buf && ready_event, "ftk_dp_allreduce_after: null pointer argument");
  FTK_REQUIRE(dtype == FTK_BF16 || dtype == FTK_F32, "ftk_dp_allreduce_after: bad dtype %d", dtype);
  FTK_REQUIRE((reinterpret_cast<uintptr_t>(buf) & 15) == 0, "ftk_dp_allreduce_after: buffer %p is not 16-byte aligned",
              buf);
  if (count == 0) return FTK_OK;
  FTK_CUDA_CHECK(cudaStreamWaitEvent(dp->comm_stream, static_cast<cudaEvent_t>(ready_event), 0));
  return allreduce_on_comm_stream(dp, buf, count, dtype, prescale);
}

static int allreduce_on_comm_stream(ftk_dp_t* dp, void* buf, size_t count, int32_t dtype, float prescale) {
  ftk::TimelineScope tl_(FTK_TL_ALLREDUCE, dp->comm_stream);
  // The scale is folded into the collective (ncclAvg for 1/world, a pre-multiplied sum otherwise): a separate
  // scaling kernel on the comm stream would have to squeeze through whatever SMs a concurrent persistent GEMM left
  // free and delays the all-reduce by about a millisecond (profiles/r01_overlap_probe.txt).
  const int nccl_dt = dtype == FTK_F32 ? ncclFloat32 : ncclBfloat16;
  if (dp->world > 1) {
    const float avg = 1.0f / static_cast<float>(dp->world);
    if (prescale == 1.0f) {
      FTK_NCCL_CHECK(dp->api, dp->api->AllReduce(buf, buf, count, nccl_dt, ncclSum, dp->comm, dp->comm_stream));
    } else if (fabsf(prescale - avg) <= 1e-7f * avg) {
      FTK_NCCL_CHECK(dp->api, dp->api->AllReduce(buf, buf, count, nccl_dt, ncclAvg, dp->comm, dp->comm_stream));
    } else if (dp->api->RedOpCreatePreMulSum != nullptr && dp->api->RedOpDestroy != nullptr) {
      int op = 0;
      float s32 = prescale;
      __nv_bfloat16 s16 = __float2bfloat16(prescale);
      void* scalar = dtype == FTK_F32 ? static_cast<void*>(&s32) : static_cast<void*>(&s16);
      FTK_NCCL_CHECK(dp->api, dp->api->RedOpCreatePreMulSum(&op, scalar, nccl_dt, ncclScalarHostImmediate, dp->comm));
      FTK_NCCL_CHECK(dp->api, dp->api->AllReduce(buf, buf, count, nccl_dt, op, dp->comm, dp->comm_stream));
      FTK_NCCL_CHECK(dp->api, dp->api->RedOpDestroy(op, dp->comm));
    } else {
      if (int rc = launch_prescale(dp, buf, count, dtype, prescale)) return rc;
      FTK_NCCL_CHECK(dp->api, dp->api->AllReduce(buf, buf, count, nccl_dt, ncclSum, dp->comm, dp->comm_stream));
    }
  } else if (prescale != 1.0f) {
    if (int rc = launch_prescale(dp, buf, count, dtype, prescale)) return rc;  // world 1: only the scale remains
  }
  return FTK_OK;
}

extern "C" int ftk_dp_join(ftk_dp_t* dp, void* stream) {
  FTK_REQUIRE(dp != nullptr, "ftk_dp_join: null handle");
  FTK_CUDA_CHECK(cudaEventRecord(dp->ev_done, dp->comm_stream));
  FTK_CUDA_CHECK(cudaStreamWaitEvent(static_cast<cudaStream_t>(stream), dp->ev_done, 0));
  return FTK_OK;
}

extern "C" int ftk_dp_destroy(ftk_dp_t* dp) {
  if (dp == nullptr) return FTK_OK;
  cudaStreamSynchronize(dp->comm_stream);
  if (dp->comm != nullptr) dp->api->CommDestroy(dp->comm);
  cudaEventDestroy(dp->ev_ready);
  cudaEventDestroy(dp->ev_done);
  cudaStreamDestroy(dp->comm_stream);
  free(dp);
  return FTK_OK;
}
