// Optimizer step on the device, consuming the all-reduced gradient buckets in place.
//
// Replaces the reference's DeepSpeed AdamW (finetuner/ds_config.json:10-18 "optimizer": AdamW with lr / betas / eps /
// weight_decay filled in by HF from TrainingArguments, FT.py:987-998: weight_decay 0.01), which by default runs on the
// CPU behind ZeRO offload (ds_config.json:35-40), and DeepSpeed's gradient clipping (ds_config.json:44
// "gradient_clipping": auto = max_grad_norm).  Arithmetic follows torch.optim.AdamW (decoupled weight decay, bias
// correction), which DeepSpeed's adam_w_mode reproduces.
//
// HBM-bound elementwise work: per parameter 14 B read + 14 B written with bf16 params/grads and fp32 moments/master
// (28 B), or 16 B + 12 B for fp32 params.  16-byte vector accesses, grid-stride, grid = min(needed, 8 blocks / SM).
#include <cuda_bf16.h>
#include <math.h>

#include "ftk_host.h"

namespace {

constexpr int kThreads = 256;
constexpr int kMaxBlocks = 2048;  // partial sums the norm reduction may use

// ---------------------------------------------------------------- squared L2 norm (deterministic two-level)
template <typename T>
__device__ __forceinline__ float sq_sum_vec(const T* g, size_t i);

template <>
__device__ __forceinline__ float sq_sum_vec<float>(const float* g, size_t i) {  // 4 elements at g[4i]
  const float4 v = reinterpret_cast<const float4*>(g)[i];
  return v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
}
template <>
__device__ __forceinline__ float sq_sum_vec<__nv_bfloat16>(const __nv_bfloat16* g, size_t i) {  // 8 elements at g[8i]
  const uint4 v = reinterpret_cast<const uint4*>(g)[i];
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  float s = 0.f;
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&w[e]);
    const float a = __low2float(h), b = __high2float(h);
    s += a * a + b * b;
  }
  return s;
}

template <typename T>
__global__ void __launch_bounds__(kThreads) sqnorm_kernel(const T* __restrict__ g, size_t n, float* __restrict__ partial,
                                                          unsigned int* __restrict__ ticket, float* __restrict__ out,
                                                          int accumulate) {
  constexpr int kVec = 16 / sizeof(T);
  __shared__ float s_warp[kThreads / 32];
  __shared__ bool s_last;
  const size_t nvec = n / kVec;
  float acc = 0.f;
  for (size_t i = blockIdx.x * static_cast<size_t>(kThreads) + threadIdx.x; i < nvec;
       i += static_cast<size_t>(gridDim.x) * kThreads)
    acc += sq_sum_vec<T>(g, i);
  if (blockIdx.x == 0 && threadIdx.x < n - nvec * kVec) {
    const float x = static_cast<float>(g[nvec * kVec + threadIdx.x]);
    acc += x * x;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float b = 0.f;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) b += s_warp[w];
    partial[blockIdx.x] = b;
    __threadfence();
    s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last && threadIdx.x == 0) {  // fixed-order sum of the block partials: same bits every run
    __threadfence();
    double t = 0.0;
    for (unsigned int b = 0; b < gridDim.x; ++b) t += static_cast<double>(reinterpret_cast<volatile float*>(partial)[b]);
    *out = (accumulate ? *out : 0.f) + static_cast<float>(t);
    *ticket = 0u;
  }
}

// clip = min(1, max_norm / (sqrt(sqnorm) + 1e-6))   (torch.nn.utils.clip_grad_norm_)
__global__ void clip_coef_kernel(const float* __restrict__ sqnorm, float max_norm, float* __restrict__ coef) {
  const float nrm = sqrtf(*sqnorm);
  const float c = max_norm / (nrm + 1e-6f);
  *coef = c < 1.f ? c : 1.f;
}

// ---------------------------------------------------------------- AdamW
struct AdamArgs {
  float lr, beta1, beta2, eps, wd, bc1, bc2, gscale;
  const float* gscale_dev;  // optional device scalar multiplied into every gradient (the clip coefficient)
};

__device__ __forceinline__ float adamw_one(float p, float g, float& m, float& v, const AdamArgs& a, float gs) {
  g *= gs;
  p *= (1.f - a.lr * a.wd);  // decoupled weight decay
  m = a.beta1 * m + (1.f - a.beta1) * g;
  v = a.beta2 * v + (1.f - a.beta2) * g * g;
  const float denom = sqrtf(v) / sqrtf(a.bc2) + a.eps;
  return p - (a.lr / a.bc1) * (m / denom);
}

// fp32 parameters updated in place; gradient fp32 or bf16
template <typename GT>
__global__ void __launch_bounds__(kThreads) adamw_f32_kernel(float* __restrict__ p, const GT* __restrict__ g,
                                                             float* __restrict__ m, float* __restrict__ v, size_t n,
                                                             AdamArgs a) {
  const float gs = a.gscale * (a.gscale_dev != nullptr ? __ldg(a.gscale_dev) : 1.f);
  const size_t n4 = n / 4;
  for (size_t i = blockIdx.x * static_cast<size_t>(kThreads) + threadIdx.x; i < n4;
       i += static_cast<size_t>(gridDim.x) * kThreads) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float gg[4];
    if constexpr (sizeof(GT) == 4) {
      const float4 t = reinterpret_cast<const float4*>(g)[i];
      gg[0] = t.x, gg[1] = t.y, gg[2] = t.z, gg[3] = t.w;
    } else {
      const uint2 t = reinterpret_cast<const uint2*>(g)[i];
      const __nv_bfloat162 h0 = *reinterpret_cast<const __nv_bfloat162*>(&t.x);
      const __nv_bfloat162 h1 = *reinterpret_cast<const __nv_bfloat162*>(&t.y);
      gg[0] = __low2float(h0), gg[1] = __high2float(h0), gg[2] = __low2float(h1), gg[3] = __high2float(h1);
    }
    pp.x = adamw_one(pp.x, gg[0], mm.x, vv.x, a, gs);
    pp.y = adamw_one(pp.y, gg[1], mm.y, vv.y, a, gs);
    pp.z = adamw_one(pp.z, gg[2], mm.z, vv.z, a, gs);
    pp.w = adamw_one(pp.w, gg[3], mm.w, vv.w, a, gs);
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
  if (blockIdx.x == 0 && threadIdx.x < n - n4 * 4) {
    const size_t i = n4 * 4 + threadIdx.x;
    p[i] = adamw_one(p[i], static_cast<float>(g[i]), m[i], v[i], a, gs);
  }
}

// bf16 parameters with an fp32 master copy: the master is updated, the bf16 parameter is its rounding.
// The gradient is bf16, or fp32 (main gradients accumulated and all-reduced in fp32, as the reference's
// ds_config.json:45 "communication_data_type": "fp32" and DeepSpeed's fp32 gradient accumulation do).
template <typename GT>
__global__ void __launch_bounds__(kThreads) adamw_bf16_kernel(__nv_bfloat16* __restrict__ p,
                                                              const GT* __restrict__ g,
                                                              float* __restrict__ master, float* __restrict__ m,
                                                              float* __restrict__ v, size_t n, AdamArgs a) {
  const float gs = a.gscale * (a.gscale_dev != nullptr ? __ldg(a.gscale_dev) : 1.f);
  const size_t n8 = n / 8;
  for (size_t i = blockIdx.x * static_cast<size_t>(kThreads) + threadIdx.x; i < n8;
       i += static_cast<size_t>(gridDim.x) * kThreads) {
    // optimizer state is touched once per step: streaming loads and stores keep it from displacing anything in L2
    float gg[8];
    if constexpr (sizeof(GT) == 4) {
      const float4 g0 = __ldcs(reinterpret_cast<const float4*>(g) + 2 * i);
      const float4 g1 = __ldcs(reinterpret_cast<const float4*>(g) + 2 * i + 1);
      gg[0] = g0.x, gg[1] = g0.y, gg[2] = g0.z, gg[3] = g0.w, gg[4] = g1.x, gg[5] = g1.y, gg[6] = g1.z, gg[7] = g1.w;
    } else {
      const uint4 gv = __ldcs(reinterpret_cast<const uint4*>(g) + i);
      const uint32_t gw[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&gw[e]);
        gg[2 * e] = __low2float(h);
        gg[2 * e + 1] = __high2float(h);
      }
    }
    float4 w0 = __ldcs(reinterpret_cast<const float4*>(master) + 2 * i);
    float4 w1 = __ldcs(reinterpret_cast<const float4*>(master) + 2 * i + 1);
    float4 m0 = __ldcs(reinterpret_cast<const float4*>(m) + 2 * i), m1 = __ldcs(reinterpret_cast<const float4*>(m) + 2 * i + 1);
    float4 v0 = __ldcs(reinterpret_cast<const float4*>(v) + 2 * i), v1 = __ldcs(reinterpret_cast<const float4*>(v) + 2 * i + 1);
    w0.x = adamw_one(w0.x, gg[0], m0.x, v0.x, a, gs);
    w0.y = adamw_one(w0.y, gg[1], m0.y, v0.y, a, gs);
    w0.z = adamw_one(w0.z, gg[2], m0.z, v0.z, a, gs);
    w0.w = adamw_one(w0.w, gg[3], m0.w, v0.w, a, gs);
    w1.x = adamw_one(w1.x, gg[4], m1.x, v1.x, a, gs);
    w1.y = adamw_one(w1.y, gg[5], m1.y, v1.y, a, gs);
    w1.z = adamw_one(w1.z, gg[6], m1.z, v1.z, a, gs);
    w1.w = adamw_one(w1.w, gg[7], m1.w, v1.w, a, gs);
    __stcs(reinterpret_cast<float4*>(master) + 2 * i, w0);
    __stcs(reinterpret_cast<float4*>(master) + 2 * i + 1, w1);
    __stcs(reinterpret_cast<float4*>(m) + 2 * i, m0);
    __stcs(reinterpret_cast<float4*>(m) + 2 * i + 1, m1);
    __stcs(reinterpret_cast<float4*>(v) + 2 * i, v0);
    __stcs(reinterpret_cast<float4*>(v) + 2 * i + 1, v1);
    uint4 o;
    __nv_bfloat162 h;
    h = __floats2bfloat162_rn(w0.x, w0.y);
    o.x = *reinterpret_cast<uint32_t*>(&h);
    h = __floats2bfloat162_rn(w0.z, w0.w);
    o.y = *reinterpret_cast<uint32_t*>(&h);
    h = __floats2bfloat162_rn(w1.x, w1.y);
    o.z = *reinterpret_cast<uint32_t*>(&h);
    h = __floats2bfloat162_rn(w1.z, w1.w);
    o.w = *reinterpret_cast<uint32_t*>(&h);
    reinterpret_cast<uint4*>(p)[i] = o;
  }
  if (blockIdx.x == 0 && threadIdx.x < n - n8 * 8) {
    const size_t i = n8 * 8 + threadIdx.x;
    master[i] = adamw_one(master[i], static_cast<float>(g[i]), m[i], v[i], a, gs);
    p[i] = __float2bfloat16(master[i]);
  }
}

int grid_for(size_t nvec, int sm_count) {
  size_t b = (nvec + kThreads - 1) / kThreads;
  const size_t cap = static_cast<size_t>(sm_count) * 8;
  if (b > cap) b = cap;
  if (b > kMaxBlocks) b = kMaxBlocks;
  if (b == 0) b = 1;
  return static_cast<int>(b);
}

}  // namespace

extern "C" size_t ftk_grad_sqnorm_workspace(void) { return sizeof(float) * kMaxBlocks + 256; }

extern "C" int ftk_grad_sqnorm(const void* grad, size_t n, int32_t dtype, float* out, int32_t accumulate,
                               void* workspace, size_t workspace_bytes, void* stream) {
  FTK_REQUIRE(grad && out && workspace, "ftk_grad_sqnorm: null pointer argument");
  FTK_REQUIRE(dtype == FTK_BF16 || dtype == FTK_F32, "ftk_grad_sqnorm: bad dtype %d", dtype);
  FTK_REQUIRE((reinterpret_cast<uintptr_t>(grad) & 15) == 0, "ftk_grad_sqnorm: grad must be 16-byte aligned");
  if (workspace_bytes < ftk_grad_sqnorm_workspace())
    return ftk::fail(FTK_EWORKSPACE, "ftk_grad_sqnorm: workspace %zu bytes < required %zu", workspace_bytes,
                     ftk_grad_sqnorm_workspace());
  ftk::DeviceInfo di;
  if (int rc = ftk::get_device_info(&di)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* partial = static_cast<float*>(workspace);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(static_cast<uint8_t*>(workspace) + sizeof(float) * kMaxBlocks);
  FTK_CUDA_CHECK(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), st));
  ftk::TimelineScope tl_(FTK_TL_SQNORM, st);
  if (dtype == FTK_F32)
    sqnorm_kernel<float><<<grid_for(n / 4, di.sm_count), kThreads, 0, st>>>(static_cast<const float*>(grad), n, partial,
                                                                          ticket, out, accumulate);
  else
    sqnorm_kernel<__nv_bfloat16><<<grid_for(n / 8, di.sm_count), kThreads, 0, st>>>(
        static_cast<const __nv_bfloat16*>(grad), n, partial, ticket, out, accumulate);
  FTK_CUDA_CHECK(cudaGetLastError());
  return FTK_OK;
}

extern "C" int ftk_clip_coef(const float* sqnorm, float max_norm, float* coef, void* stream) {
  FTK_REQUIRE(sqnorm && coef, "ftk_clip_coef: null pointer argument");
  FTK_REQUIRE(max_norm > 0.f, "ftk_clip_coef: max_norm must be positive");
  clip_coef_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(sqnorm, max_norm, coef);
  FTK_CUDA_CHECK(cudaGetLastError());
  return FTK_OK;
}

extern "C" int ftk_adamw_step(void* param, const void* grad, float* master, float* exp_avg, float* exp_avg_sq, size_t n,
                              int32_t param_dtype, int32_t grad_dtype, float lr, float beta1, float beta2, float eps,
                              float weight_decay, int64_t step, float grad_scale, const float* grad_scale_dev,
                              void* stream) {
  FTK_REQUIRE(param && grad && exp_avg && exp_avg_sq, "ftk_adamw_step: null pointer argument");
  FTK_REQUIRE(step >= 1, "ftk_adamw_step: step counts from 1");
  FTK_REQUIRE(param_dtype == FTK_F32 || (param_dtype == FTK_BF16 && master != nullptr),
              "ftk_adamw_step: bf16 parameters need an fp32 master copy");
  FTK_REQUIRE(grad_dtype == FTK_F32 || grad_dtype == FTK_BF16, "ftk_adamw_step: bad grad dtype %d", grad_dtype);
  const uintptr_t al = reinterpret_cast<uintptr_t>(param) | reinterpret_cast<uintptr_t>(grad) |
                       reinterpret_cast<uintptr_t>(exp_avg) | reinterpret_cast<uintptr_t>(exp_avg_sq) |
                       reinterpret_cast<uintptr_t>(master);
  FTK_REQUIRE((al & 15) == 0, "ftk_adamw_step: buffers must be 16-byte aligned");
  if (n == 0) return FTK_OK;
  ftk::DeviceInfo di;
  if (int rc = ftk::get_device_info(&di)) return rc;
  AdamArgs a;
  a.lr = lr;
  a.beta1 = beta1;
  a.beta2 = beta2;
  a.eps = eps;
  a.wd = weight_decay;
  a.bc1 = 1.f - powf(beta1, static_cast<float>(step));
  a.bc2 = 1.f - powf(beta2, static_cast<float>(step));
  a.gscale = grad_scale;
  a.gscale_dev = grad_scale_dev;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ftk::TimelineScope tl_(FTK_TL_ADAMW, st);
  if (param_dtype == FTK_F32) {
    if (grad_dtype == FTK_F32)
      adamw_f32_kernel<float><<<grid_for(n / 4, di.sm_count), kThreads, 0, st>>>(
          static_cast<float*>(param), static_cast<const float*>(grad), exp_avg, exp_avg_sq, n, a);
    else
      adamw_f32_kernel<__nv_bfloat16><<<grid_for(n / 4, di.sm_count), kThreads, 0, st>>>(
          static_cast<float*>(param), static_cast<const __nv_bfloat16*>(grad), exp_avg, exp_avg_sq, n, a);
  } else if (grad_dtype == FTK_F32) {
    adamw_bf16_kernel<float><<<grid_for(n / 8, di.sm_count), kThreads, 0, st>>>(
        static_cast<__nv_bfloat16*>(param), static_cast<const float*>(grad), master, exp_avg, exp_avg_sq, n, a);
  } else {
    adamw_bf16_kernel<__nv_bfloat16><<<grid_for(n / 8, di.sm_count), kThreads, 0, st>>>(
        static_cast<__nv_bfloat16*>(param), static_cast<const __nv_bfloat16*>(grad), master, exp_avg, exp_avg_sq, n, a);
  }
  FTK_CUDA_CHECK(cudaGetLastError());
  return FTK_OK;
}
