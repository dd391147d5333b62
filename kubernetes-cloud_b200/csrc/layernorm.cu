// LayerNorm of the transformer trunk, forward and backward, sm_100a.  HBM bound.
//
// Replaces torch.nn.LayerNorm inside HF's GPT-Neo blocks (ln_1, ln_2, ln_f; reached from finetuner.py's
// model(**inputs), FT.py:479, twice per step under gradient checkpointing, FT.py:884) and the elementwise add with
// which autograd joins the residual branch's gradient to the LayerNorm's at the point where the two fork.
// Arithmetic of torch.nn.functional.layer_norm on bf16 tensors: statistics and normalisation in fp32 (biased
// variance, rstd = 1 / sqrt(var + eps)), one rounding to bf16 at the end.
//
//   forward   one warp per row, the row held in registers (16-byte loads, two shuffle reductions: mean, then the
//             centred sum of squares); writes y, mean, rstd.  Algorithmic bytes: 4 per element (2 read + 2 written).
//   backward  one block per group of rows, D / 8 threads, each owning 8 consecutive columns: dx needs two row sums
//             (block reduction, two rows in flight, one wave of resident blocks), dgamma / dbeta need column sums, which every thread keeps in
//             registers for its own columns over the block's rows; per-block partial sums go to a workspace and a
//             second small kernel adds them in a fixed order (deterministic, no atomics).  dx (+)= the gradient that
//             reached the LayerNorm's input along the residual branch, when given.  Algorithmic bytes: 6 per element
//             (+2 with the residual gradient).
#include <cuda.h>
#include <math.h>

#include "ftk_host.h"
#include "ftk_ptx.cuh"

namespace {

__device__ __forceinline__ void unpack8(const uint4& w, float* f) {
  const uint32_t u[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&u[e]);
    f[2 * e] = __low2float(h);
    f[2 * e + 1] = __high2float(h);
  }
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint32_t u[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * e], f[2 * e + 1]);
    u[e] = *reinterpret_cast<uint32_t*>(&h);
  }
  return make_uint4(u[0], u[1], u[2], u[3]);
}
__device__ __forceinline__ uint4 ld_stream(const void* p) {  // read-once data: do not keep it in L1
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

constexpr int kFwdWarps = 8;

// NC: 256-column chunks of a row a lane keeps in registers (D <= NC * 256, D % 8 == 0).
template <int NC>
__global__ void __launch_bounds__(kFwdWarps * 32)
    ln_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ gamma,
                  const __nv_bfloat16* __restrict__ beta, __nv_bfloat16* __restrict__ y, float* __restrict__ mean,
                  float* __restrict__ rstd, int64_t M, int D, float eps) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t row = static_cast<int64_t>(blockIdx.x) * kFwdWarps + warp;
  if (row >= M) return;
  const __nv_bfloat16* xr = x + row * D;
  uint4 raw[NC];  // the row stays packed (4 registers per 8 values): more warps per SM hide more of the HBM latency
#pragma unroll
  for (int c = 0; c < NC; ++c)
    if (c * 256 + lane * 8 < D) raw[c] = ld_stream(xr + c * 256 + lane * 8);
  float s = 0.f;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    if (c * 256 + lane * 8 < D) {
      float v[8];
      unpack8(raw[c], v);
#pragma unroll
      for (int e = 0; e < 8; ++e) s += v[e];
    }
  }
  const float mu = warp_sum(s) / static_cast<float>(D);
  float q = 0.f;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    if (c * 256 + lane * 8 < D) {
      float v[8];
      unpack8(raw[c], v);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float d = v[e] - mu;
        q += d * d;
      }
    }
  }
  const float rs = rsqrtf(warp_sum(q) / static_cast<float>(D) + eps);
  __nv_bfloat16* yr = y + row * D;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int col = c * 256 + lane * 8;
    if (col < D) {
      float v[8], g[8], bt[8], o[8];
      unpack8(raw[c], v);
      unpack8(__ldg(reinterpret_cast<const uint4*>(gamma + col)), g);
      if (beta != nullptr) {
        unpack8(__ldg(reinterpret_cast<const uint4*>(beta + col)), bt);
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) bt[e] = 0.f;
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = fmaf((v[e] - mu) * rs, g[e], bt[e]);
      *reinterpret_cast<uint4*>(yr + col) = pack8(o);
    }
  }
  if (lane == 0) {
    mean[row] = mu;
    rstd[row] = rs;
  }
}

// R rows in flight per block iteration; blockDim.x = D / 8 (a multiple of 32, at most kMaxT); kMinBlocks resident
// blocks per SM are asked of the register allocator (the grid is sized to exactly that many, see plan_bwd).
template <int R, int kMaxT, int kMinBlocks>
__global__ void __launch_bounds__(kMaxT, kMinBlocks)
    ln_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                  const __nv_bfloat16* __restrict__ gamma, const float* __restrict__ mean, const float* __restrict__ rstd,
                  const __nv_bfloat16* __restrict__ dx_add, __nv_bfloat16* __restrict__ dx, float* __restrict__ part,
                  int64_t M, int D, int rows_per_block) {
  extern __shared__ float red[];  // [2][warps][2 R]: two buffers, so one __syncthreads per iteration is enough
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
  const int col = tid * 8;
  float g[8], ag[8], ab[8];
  unpack8(__ldg(reinterpret_cast<const uint4*>(gamma + col)), g);
#pragma unroll
  for (int e = 0; e < 8; ++e) ag[e] = ab[e] = 0.f;
  const float inv_d = 1.f / static_cast<float>(D);
  const int64_t r0 = static_cast<int64_t>(blockIdx.x) * rows_per_block;
  const int64_t r1 = (r0 + rows_per_block < M) ? r0 + rows_per_block : M;
  int it = 0;
  for (int64_t row = r0; row < r1; row += R, ++it) {
    uint4 xw[R], dw[R], aw[R];  // the residual gradient is fetched with the rest: after the reduction its latency would be exposed
    float mu[R], rs[R], s1[R], s2[R];
#pragma unroll
    for (int rr = 0; rr < R; ++rr) {
      const bool ok = row + rr < r1;
      const size_t off = static_cast<size_t>(row + rr) * D + col;
      xw[rr] = ok ? ld_stream(x + off) : make_uint4(0, 0, 0, 0);
      dw[rr] = ok ? ld_stream(dy + off) : make_uint4(0, 0, 0, 0);
      aw[rr] = (ok && dx_add != nullptr) ? ld_stream(dx_add + off) : make_uint4(0, 0, 0, 0);
      mu[rr] = ok ? mean[row + rr] : 0.f;
      rs[rr] = ok ? rstd[row + rr] : 0.f;
    }
#pragma unroll
    for (int rr = 0; rr < R; ++rr) {  // rows past the end carry dy = 0: they add nothing anywhere
      float xv[8], dv[8];
      unpack8(xw[rr], xv);
      unpack8(dw[rr], dv);
      float a = 0.f, b = 0.f;
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float xh = (xv[e] - mu[rr]) * rs[rr], gg = dv[e] * g[e];
        a += gg;
        b = fmaf(gg, xh, b);
        ab[e] += dv[e];
        ag[e] = fmaf(dv[e], xh, ag[e]);
      }
      s1[rr] = warp_sum(a);
      s2[rr] = warp_sum(b);
    }
    float* rb = red + (it & 1) * nw * 2 * R;
    if (lane == 0) {
#pragma unroll
      for (int rr = 0; rr < R; ++rr) {
        rb[warp * 2 * R + 2 * rr] = s1[rr];
        rb[warp * 2 * R + 2 * rr + 1] = s2[rr];
      }
    }
    __syncthreads();
#pragma unroll
    for (int rr = 0; rr < R; ++rr) {
      if (row + rr >= r1) break;  // block-uniform
      float t1 = 0.f, t2 = 0.f;
      for (int w = 0; w < nw; ++w) {  // same order in every thread: the row sums are identical across the block
        t1 += rb[w * 2 * R + 2 * rr];
        t2 += rb[w * 2 * R + 2 * rr + 1];
      }
      t1 *= inv_d;
      t2 *= inv_d;
      float xv[8], dv[8], o[8];
      unpack8(xw[rr], xv);
      unpack8(dw[rr], dv);
      const size_t off = static_cast<size_t>(row + rr) * D + col;
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float xh = (xv[e] - mu[rr]) * rs[rr];
        o[e] = rs[rr] * (dv[e] * g[e] - t1 - xh * t2);
      }
      if (dx_add != nullptr) {
        float av[8];
        unpack8(aw[rr], av);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] += av[e];
      }
      *reinterpret_cast<uint4*>(dx + off) = pack8(o);
    }
  }
  float* pg = part + static_cast<size_t>(blockIdx.x) * 2 * D + col;
  *reinterpret_cast<float4*>(pg) = make_float4(ag[0], ag[1], ag[2], ag[3]);
  *reinterpret_cast<float4*>(pg + 4) = make_float4(ag[4], ag[5], ag[6], ag[7]);
  *reinterpret_cast<float4*>(pg + D) = make_float4(ab[0], ab[1], ab[2], ab[3]);
  *reinterpret_cast<float4*>(pg + D + 4) = make_float4(ab[4], ab[5], ab[6], ab[7]);
}

// dgamma[c] = sum over blocks of part[b][0][c], dbeta[c] = ... part[b][1][c]; block = 32 columns x 8 row groups.
__global__ void __launch_bounds__(256) ln_bwd_finalize_kernel(const float* __restrict__ part, int n_blocks, int D,
                                                              float* __restrict__ dgamma, float* __restrict__ dbeta) {
  __shared__ float sm[2][8][33];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + cx;
  float a = 0.f, b = 0.f;
  if (col < D) {
#pragma unroll 8
    for (int i = ry; i < n_blocks; i += 8) {  // fixed order per thread; unrolled so that the loads overlap
      a += part[static_cast<size_t>(i) * 2 * D + col];
      b += part[static_cast<size_t>(i) * 2 * D + D + col];
    }
  }
  sm[0][ry][cx] = a;
  sm[1][ry][cx] = b;
  __syncthreads();
  if (ry == 0 && col < D) {
    float ta = 0.f, tb = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      ta += sm[0][i][cx];
      tb += sm[1][i][cx];
    }
    dgamma[col] = ta;
    if (dbeta != nullptr) dbeta[col] = tb;
  }
}

constexpr int kBwdRows = 2;  // rows in flight per block iteration (x, dy and the residual gradient of each)

// Resident blocks per SM the backward is compiled for, by block size (D / 8 threads).
constexpr int bwd_blocks_per_sm(int threads) { return threads <= 256 ? 3 : 1; }

// How the backward cuts the rows: one wave of resident blocks (no tail wave), every block a multiple of kBwdRows rows.
void plan_bwd(int64_t M, int D, int sm_count, int* n_blocks, int* rows_per_block) {
  int64_t blocks = (M + kBwdRows - 1) / kBwdRows;
  const int64_t cap = static_cast<int64_t>(sm_count) * bwd_blocks_per_sm(D / 8);
  if (blocks > cap) blocks = cap;
  int64_t rpb = (M + blocks - 1) / blocks;
  rpb = (rpb + kBwdRows - 1) / kBwdRows * kBwdRows;
  *rows_per_block = static_cast<int>(rpb);
  *n_blocks = static_cast<int>((M + rpb - 1) / rpb);
}

int check_ln(const void* x, const void* gamma, int64_t M, int32_t D) {
  FTK_REQUIRE(x && gamma, "ftk_layernorm: null input");
  FTK_REQUIRE(M > 0 && M < (int64_t(1) << 31), "ftk_layernorm: bad row count %lld", static_cast<long long>(M));
  FTK_REQUIRE(D > 0 && D % 256 == 0 && D <= 8192, "ftk_layernorm: D=%d must be a multiple of 256, at most 8192", D);
  FTK_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(gamma) % 16 == 0,
              "ftk_layernorm: buffers must be 16-byte aligned");
  return FTK_OK;
}

}  // namespace

extern "C" int ftk_layernorm_fwd(const void* X, const void* gamma, const void* beta, int64_t M, int32_t D, float eps,
                                 void* Y, float* mean, float* rstd, void* stream) {
  if (int rc = check_ln(X, gamma, M, D)) return rc;
  FTK_REQUIRE(Y && mean && rstd, "ftk_layernorm_fwd: null output");
  FTK_REQUIRE(reinterpret_cast<uintptr_t>(Y) % 16 == 0 && (beta == nullptr || reinterpret_cast<uintptr_t>(beta) % 16 == 0),
              "ftk_layernorm_fwd: buffers must be 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const auto* x = static_cast<const __nv_bfloat16*>(X);
  const auto* g = static_cast<const __nv_bfloat16*>(gamma);
  const auto* b = static_cast<const __nv_bfloat16*>(beta);
  auto* y = static_cast<__nv_bfloat16*>(Y);
  const unsigned grid = static_cast<unsigned>((M + kFwdWarps - 1) / kFwdWarps);
  ftk::TimelineScope tl_(FTK_TL_LAYERNORM_FWD, st);
  const int nc = D / 256;
  if (nc <= 4)
    ln_fwd_kernel<4><<<grid, kFwdWarps * 32, 0, st>>>(x, g, b, y, mean, rstd, M, D, eps);
  else if (nc <= 8)
    ln_fwd_kernel<8><<<grid, kFwdWarps * 32, 0, st>>>(x, g, b, y, mean, rstd, M, D, eps);
  else if (nc <= 16)
    ln_fwd_kernel<16><<<grid, kFwdWarps * 32, 0, st>>>(x, g, b, y, mean, rstd, M, D, eps);
  else
    ln_fwd_kernel<32><<<grid, kFwdWarps * 32, 0, st>>>(x, g, b, y, mean, rstd, M, D, eps);
  FTK_CUDA_CHECK(cudaGetLastError());
  return FTK_OK;
}

extern "C" size_t ftk_layernorm_bwd_workspace(int64_t M, int32_t D) {
  if (M <= 0 || D <= 0) return 0;
  ftk::DeviceInfo di;
  if (ftk::get_device_info(&di) != FTK_OK) return 0;
  int n_blocks, rpb;
  plan_bwd(M, D, di.sm_count, &n_blocks, &rpb);
  return static_cast<size_t>(n_blocks) * 2 * D * sizeof(float);
}

extern "C" int ftk_layernorm_bwd(const void* dY, const void* X, const void* gamma, const float* mean, const float* rstd,
                                 const void* dx_add, int64_t M, int32_t D, void* dX, float* dgamma, float* dbeta,
                                 void* workspace, size_t workspace_bytes, void* stream) {
  if (int rc = check_ln(X, gamma, M, D)) return rc;
  FTK_REQUIRE(dY && mean && rstd && dX && dgamma && workspace, "ftk_layernorm_bwd: null pointer argument");
  FTK_REQUIRE(reinterpret_cast<uintptr_t>(dY) % 16 == 0 && reinterpret_cast<uintptr_t>(dX) % 16 == 0 &&
                  reinterpret_cast<uintptr_t>(dx_add) % 16 == 0 && reinterpret_cast<uintptr_t>(workspace) % 16 == 0,
              "ftk_layernorm_bwd: buffers must be 16-byte aligned");
  ftk::DeviceInfo di;
  if (int rc = ftk::get_device_info(&di)) return rc;
  int n_blocks, rpb;
  plan_bwd(M, D, di.sm_count, &n_blocks, &rpb);
  const size_t need = static_cast<size_t>(n_blocks) * 2 * D * sizeof(float);
  if (workspace_bytes < need)
    return ftk::fail(FTK_EWORKSPACE, "ftk_layernorm_bwd: workspace %zu bytes < required %zu", workspace_bytes, need);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* part = static_cast<float*>(workspace);
  const int threads = D / 8;
  const size_t smem = static_cast<size_t>(2) * (threads / 32) * 2 * kBwdRows * sizeof(float);
  ftk::TimelineScope tl_(FTK_TL_LAYERNORM_BWD, st);
  const auto* dy_ = static_cast<const __nv_bfloat16*>(dY);
  const auto* x_ = static_cast<const __nv_bfloat16*>(X);
  const auto* g_ = static_cast<const __nv_bfloat16*>(gamma);
  const auto* add_ = static_cast<const __nv_bfloat16*>(dx_add);
  auto* dx_ = static_cast<__nv_bfloat16*>(dX);
  if (threads <= 256)
    ln_bwd_kernel<kBwdRows, 256, bwd_blocks_per_sm(256)><<<n_blocks, threads, smem, st>>>(dy_, x_, g_, mean, rstd, add_, dx_,
                                                                                          part, M, D, rpb);
  else if (threads <= 512)
    ln_bwd_kernel<kBwdRows, 512, bwd_blocks_per_sm(512)><<<n_blocks, threads, smem, st>>>(dy_, x_, g_, mean, rstd, add_, dx_,
                                                                                          part, M, D, rpb);
  else
    ln_bwd_kernel<kBwdRows, 1024, 1><<<n_blocks, threads, smem, st>>>(dy_, x_, g_, mean, rstd, add_, dx_, part, M, D, rpb);
  FTK_CUDA_CHECK(cudaGetLastError());
  ln_bwd_finalize_kernel<<<(D + 31) / 32, 256, 0, st>>>(part, n_blocks, D, dgamma, dbeta);
  FTK_CUDA_CHECK(cudaGetLastError());
  return FTK_OK;
}
