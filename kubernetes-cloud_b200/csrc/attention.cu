// Causal / local-window self-attention for the transformer trunk, forward and backward, sm_100a.
//
// Replaces HF GPT-Neo's eager attention (GPTNeoSelfAttention._attn: fp32 matmul -> where(causal / sliding-window mask)
// -> + attention_mask -> softmax -> matmul, reached from finetuner.py's model(**inputs), FT.py:479, and recomputed by
// gradient checkpointing, FT.py:884), which materialises a [batch, heads, T, T] fp32 score tensor three times per
// layer and step.  Here the scores live in TMEM only.
//
// Layout: Q, K, V, O, dO, dQ, dK, dV are the projections' own [batch * T, heads * head_dim] bf16 matrices (no head
// split / merge copies); one CTA owns a 128-row block of one (batch, head).
//
// Forward, two passes over the key blocks so that the output accumulator never needs rescaling:
//   pass 0   S = Q.K^T (tcgen05.mma, accumulator in TMEM) -> masked row max / sum-exp -> lse
//   pass 1   S again -> P = exp(S - lse) as bf16 into swizzled shared memory -> O += P.V (second tcgen05.mma)
// Backward, two kernels, no atomics:
//   kv-stationary   for a key block: over the query blocks that see it, S and dP = dO.V^T -> P, dS = P*(dP - D)
//                   -> dV += P^T.dO, dK += dS^T.Q      (both accumulators stay in TMEM)
//   q-stationary    for a query block: S, dP again -> dS -> dQ += dS.K
// The extra S / dP recomputations cost tensor flops that this workload has to spare (attention is ~4 % of a step's
// tensor work); what mattered was removing the fp32 score tensors.
//
// One warp (lane 0) issues TMA and MMA, four warps own one accumulator row each (softmax, tile conversion); the phases
// of a tile are strictly ordered by mbarriers and __syncthreads, with the next tile's operands prefetched by TMA.
#include <cuda.h>
#include <math.h>

#include "ftk_host.h"
#include "ftk_ptx.cuh"

namespace {
using namespace ftk;

constexpr int QB = 128;            // rows of a block (queries or keys) = TMEM lanes
// Warps 0-7: row threads.  Thread t owns accumulator row t % 128 (warp w reads TMEM lanes 32 * (w % 4)) and the column
// half t / 128 of every tile, so each row is worked on by two threads on the same SM sub-partition: with a single warp
// per scheduler nothing hid the TMEM-load and MUFU latencies (ncu: issue slots 22 % busy, tensor pipe 15 %).
// Warp 8: TMA + MMA issue (lane 0), TMEM allocation.
constexpr int kAttnThreads = 288;
constexpr int kIssuerWarp = 8;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr int kMaxKeyBlocks = 256;  // per-key-block padding flags kept in shared memory: T <= 32768 with a key mask

struct AttnParams {
  const uint8_t* key_mask;  // [batch * T] (1 = may be attended) or null
  float* lse;               // [batch * heads * T]
  const float* delta;       // backward: rowsum(dO * O) [batch * heads * T]
  __nv_bfloat16* out0;      // forward: O; kv kernel: dK; q kernel: dQ     ([batch * T, heads * head_dim])
  __nv_bfloat16* out1;      // kv kernel: dV
  int T, H, window;         // window > 0: key j is visible from query i iff i - window < j <= i; 0: plain causal
  float scale;              // S = scale * Q.K^T
};

__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

// Is key `kj` visible from query `qi`?
__device__ __forceinline__ bool visible(int qi, int kj, int window) {
  return kj <= qi && (window <= 0 || kj > qi - window);
}

// Does every query of block qb see every key of block kb (so that no per-element test is needed)?
__device__ __forceinline__ bool tile_all_visible(int qb, int kb, int window, const uint8_t* mrow) {
  return mrow == nullptr && kb * QB + QB - 1 <= qb * QB && (window <= 0 || kb * QB > qb * QB + QB - 1 - window);
}

// First / last key block a query block needs, and first / last query block a key block is seen by (T % 128 == 0).
__device__ __forceinline__ void kv_range(int qb, int window, int& lo, int& hi) {
  hi = qb;
  lo = 0;
  if (window > 0) {
    const int first_key = qb * QB - window + 1;  // lowest key the block's first query sees
    lo = first_key > 0 ? first_key / QB : 0;
  }
}
__device__ __forceinline__ void q_range(int kb, int window, int n_blk, int& lo, int& hi) {
  lo = kb;
  hi = n_blk - 1;
  if (window > 0) {
    const int last_q = kb * QB + QB - 1 + window - 1;  // highest query that still sees the block's last key
    hi = min(n_blk - 1, last_q / QB);
  }
}

// TMA: a [128 rows x DH cols] tile as 64 x 64 boxes: column half kk at +kk * 16 KB, row r at r * 128 B inside it
// (SWIZZLE_128B).  That is the K-major operand layout (rows = M or N, head_dim = K).  The same bytes serve as an
// MN-major operand (rows = K, head_dim = M or N): K block kb (64 rows) starts at +kb * 8 KB, 16 K rows are 2 KB apart,
// and the two 64-column halves are 16 KB apart (the descriptor's leading byte offset).
template <int DH>
__device__ __forceinline__ void load_tile(uint8_t* dst, const CUtensorMap* tm, uint64_t* bar, int row0, int col0) {
#pragma unroll
  for (int kk = 0; kk < DH / 64; ++kk)
#pragma unroll
    for (int rr = 0; rr < 2; ++rr) tma_load_2d(dst + kk * 16384 + rr * 8192, tm, bar, col0 + kk * 64, row0 + rr * 64);
}

// D[tmem] (+)= A.B^T with both operands K-major tiles of [128 rows x KD] (KD = 64 or 128): S = Q.K^T, dP = dO.V^T.
template <int KD>
__device__ __forceinline__ void mma_kk(uint32_t tmem_d, const uint8_t* a, const uint8_t* b, uint32_t n, bool accumulate) {
  const uint32_t idesc = make_idesc_bf16(128, n, false, false);
  const uint32_t aa = smem_u32(a), bb = smem_u32(b);
#pragma unroll
  for (int kk = 0; kk < KD / 64; ++kk)
#pragma unroll
    for (int k = 0; k < 4; ++k)
      umma_bf16<1>(tmem_d, make_sw128_desc(aa + kk * 16384 + k * 32, 0, 1024), make_sw128_desc(bb + kk * 16384 + k * 32, 0, 1024),
                   idesc, (accumulate || kk != 0 || k != 0) ? 1u : 0u);
}
// D[tmem] (+)= A.B, A K-major [128 x 128] (a P / dS tile written by the row threads), B a TMA-loaded [128 K rows x ND]
// tile in its MN-major view.
template <int ND>
__device__ __forceinline__ void mma_k_mn(uint32_t tmem_d, const uint8_t* a, const uint8_t* b, bool accumulate) {
  const uint32_t idesc = make_idesc_bf16(128, ND, false, true);
  const uint32_t aa = smem_u32(a), bb = smem_u32(b);
#pragma unroll
  for (int kb = 0; kb < 2; ++kb)
#pragma unroll
    for (int k = 0; k < 4; ++k)
      umma_bf16<1>(tmem_d, make_sw128_desc(aa + kb * 16384 + k * 32, 0, 1024),
                   make_sw128_desc(bb + kb * 8192 + k * 2048, 16384, 1024), idesc,
                   (accumulate || kb != 0 || k != 0) ? 1u : 0u);
}
// D[tmem] (+)= A^T.B, A stored as [128 K rows x 128 M] (MN-major: a P / dS tile used transposed), B a TMA-loaded
// [128 K rows x ND] tile in its MN-major view.
template <int ND>
__device__ __forceinline__ void mma_mn_mn(uint32_t tmem_d, const uint8_t* a, const uint8_t* b, bool accumulate) {
  const uint32_t idesc = make_idesc_bf16(128, ND, true, true);
  const uint32_t aa = smem_u32(a), bb = smem_u32(b);
#pragma unroll
  for (int kb = 0; kb < 2; ++kb)
#pragma unroll
    for (int k = 0; k < 4; ++k)
      umma_bf16<1>(tmem_d, make_sw128_desc(aa + kb * 16384 + k * 2048, 8192, 1024),
                   make_sw128_desc(bb + kb * 8192 + k * 2048, 16384, 1024), idesc,
                   (accumulate || kb != 0 || k != 0) ? 1u : 0u);
}

// Row thread `r` writes 8 consecutive values (columns c0 .. c0+7 of its row) of a [128 x 128] bf16 tile.
//   K-major form  (tile used as A with rows = M):        half (c0 / 64) at +16 KB, row r at r * 128 B
//   MN-major form (tile used transposed, rows = K):      K block (r / 64) at +16 KB, column half at +8 KB, row r % 64
// Both put 16-byte chunk c at position c ^ (row & 7) of its 128-byte line: the SWIZZLE_128B pattern TMA and UMMA use.
__device__ __forceinline__ void put8_kmajor(uint8_t* tile, int r, int c0, const uint32_t w[4]) {
  uint8_t* line = tile + (c0 >> 6) * 16384 + r * 128;
  *reinterpret_cast<uint4*>(line + ((((c0 & 63) >> 3) ^ (r & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
}
__device__ __forceinline__ void put8_mnmajor(uint8_t* tile, int r, int c0, const uint32_t w[4]) {
  uint8_t* line = tile + (r >> 6) * 16384 + (c0 >> 6) * 8192 + (r & 63) * 128;
  *reinterpret_cast<uint4*>(line + ((((c0 & 63) >> 3) ^ (r & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
}

// The visible-and-unmasked test for 32 consecutive keys of one query row, folded into the scores:
// v[j] = scale * log2(e) * s  (so that exp2 can be used), or -inf where the key must not be attended.
__device__ __forceinline__ void mask_scores(float* v, int qi, int key0, const AttnParams& p, const uint8_t* mrow, float sl2,
                                            bool all_visible) {
  if (all_visible) {  // a tile wholly inside the causal / window band of every row of the block, no padding mask
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] *= sl2;
    return;
  }
  uint32_t mbits = 0xffffffffu;
  if (mrow != nullptr) {
    const uint4 m0 = *reinterpret_cast<const uint4*>(mrow + key0), m1 = *reinterpret_cast<const uint4*>(mrow + key0 + 16);
    const uint32_t mw[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
    mbits = 0;
#pragma unroll
    for (int j = 0; j < 32; ++j) mbits |= ((mw[j >> 2] >> ((j & 3) * 8)) & 0xffu) ? (1u << j) : 0u;
  }
#pragma unroll
  for (int j = 0; j < 32; ++j)
    v[j] = (visible(qi, key0 + j, p.window) && ((mbits >> j) & 1u)) ? v[j] * sl2 : -INFINITY;
}

// ------------------------------------------------------------------------------------------------ forward
// Which 128-key blocks of one sequence hold no padding (every mask byte non-zero)?  With those known, a tile only pays
// for the per-element padding test where padding really is; HF always hands the layers an attention mask, and in a
// training batch it is all ones except in the last context of the token file.  Called by the 256 row threads.
__device__ __forceinline__ void key_block_flags(const uint8_t* mrow, int T, uint8_t* blk_ok, int tid, int lane) {
  for (int base = 0; base < T; base += 256 * 16) {
    const int off = base + tid * 16;
    bool ok = true;
    if (off < T) {
      const uint4 mk = *reinterpret_cast<const uint4*>(mrow + off);
      const uint32_t w[4] = {mk.x, mk.y, mk.z, mk.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) ok = ok && (((w[e] - 0x01010101u) & ~w[e] & 0x80808080u) == 0u);  // no zero byte
    }
    const uint32_t bal = __ballot_sync(0xffffffffu, ok);
    if ((lane & 7) == 0 && off < T) blk_ok[off >> 7] = (((bal >> lane) & 0xffu) == 0xffu) ? 1 : 0;
  }
}

// 32 raw scores of one query row -> -inf where the key must not be attended (the forward scales afterwards).
__device__ __forceinline__ void mask_raw(float* v, int qi, int key0, int window, const uint8_t* mrow) {
  uint32_t mbits = 0xffffffffu;
  if (mrow != nullptr) {
    const uint4 m0 = *reinterpret_cast<const uint4*>(mrow + key0), m1 = *reinterpret_cast<const uint4*>(mrow + key0 + 16);
    const uint32_t mw[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
    mbits = 0;
#pragma unroll
    for (int j = 0; j < 32; ++j) mbits |= ((mw[j >> 2] >> ((j & 3) * 8)) & 0xffu) ? (1u << j) : 0u;
  }
#pragma unroll
  for (int j = 0; j < 32; ++j)
    if (!(visible(qi, key0 + j, window) && ((mbits >> j) & 1u))) v[j] = -INFINITY;
}

template <int DH>
struct FwdSmem {
  static constexpr uint32_t TILE = QB * DH * 2;  // one [128 x DH] bf16 operand tile
  static constexpr uint32_t Q_OFF = 0, K_OFF = TILE, V_OFF = 3 * TILE, P_OFF = 5 * TILE, BAR_OFF = P_OFF + 32768;
  static constexpr uint32_t STAT_OFF = BAR_OFF + 128;    // [256] floats: the two column halves' row statistics
  static constexpr uint32_t FLAG_OFF = STAT_OFF + 1024;  // [kMaxKeyBlocks] bytes: key block holds no padding
  static constexpr uint32_t BYTES = 1024 + FLAG_OFF + kMaxKeyBlocks;
};

// Schedule.  Tiles t = 0 .. 2 nj - 1: pass 0 (row maxima only: no exponentials) over the nj key blocks the query block
// sees, then pass 1 (P = exp2(s - max) as bf16, row sums in fp32, O += P.V; O / sum at the end).  The scores have two
// TMEM buffers: while the row threads work on S(t) the tensor pipe already forms S(t+1), and P.V of tile t runs while
// they work on tile t+1 (they wait for it only at the point where they overwrite the P tile).  K and V each have two
// shared-memory buffers with their own barriers; a buffer is refilled as soon as the MMA that read it has completed.
template <int DH>
__global__ void __launch_bounds__(kAttnThreads, 1)
    attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                    const __grid_constant__ CUtensorMap tmV, const AttnParams p) {
  using L = FwdSmem<DH>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem + L::Q_OFF;
  uint8_t* sK = smem + L::K_OFF;  // two buffers
  uint8_t* sV = smem + L::V_OFF;  // two buffers
  uint8_t* sP = smem + L::P_OFF;
  uint64_t* bar_q = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
  uint64_t* bar_k = bar_q + 1;  // [2]
  uint64_t* bar_v = bar_q + 3;  // [2]
  uint64_t* bar_s = bar_q + 5;  // [2]
  uint64_t* bar_o = bar_q + 7;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_q + 8);
  float* stat = reinterpret_cast<float*>(smem + L::STAT_OFF);
  uint8_t* blk_ok = smem + L::FLAG_OFF;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // the query blocks that see the most keys start first, so the short ones fill the tail of the grid
  const int qb = static_cast<int>(gridDim.x) - 1 - static_cast<int>(blockIdx.x);
  const int bh = blockIdx.y, b = bh / p.H, h = bh - b * p.H;
  const int row0 = b * p.T + qb * QB;  // first row of this block in the [batch * T] matrices
  const int col0 = h * DH;
  const bool issuer = (warp == kIssuerWarp && lane == 0);
  const uint8_t* mrow = p.key_mask != nullptr ? p.key_mask + static_cast<size_t>(b) * p.T : nullptr;

  if (issuer) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    for (int i = 0; i < 8; ++i) mbar_init(bar_q + i, 1);
    fence_mbar_init();
  }
  if (warp == kIssuerWarp) {
    tmem_alloc<1>(tmem_slot, 512);
    tmem_relinquish<1>();
  } else if (mrow != nullptr) {
    key_block_flags(mrow, p.T, blk_ok, threadIdx.x, lane);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_s = *tmem_slot, tmem_o = tmem_s + 256;  // S buffers at +0 and +128

  int j_lo, j_hi;
  kv_range(qb, p.window, j_lo, j_hi);
  const int nj = j_hi - j_lo + 1, n_tiles = 2 * nj;
  auto load_k = [&](int t) {  // K block of tile t -> K buffer t & 1
    const int bi = t & 1, u = t >= nj ? t - nj : t;
    mbar_arrive_expect_tx(&bar_k[bi], L::TILE);
    load_tile<DH>(sK + bi * L::TILE, &tmK, &bar_k[bi], b * p.T + (j_lo + u) * QB, col0);
  };
  auto load_v = [&](int u) {  // V block of pass-1 tile u -> V buffer u & 1
    const int bi = u & 1;
    mbar_arrive_expect_tx(&bar_v[bi], L::TILE);
    load_tile<DH>(sV + bi * L::TILE, &tmV, &bar_v[bi], b * p.T + (j_lo + u) * QB, col0);
  };
  auto issue_s = [&](int t) {  // S(t) = Q.K(t)^T into S buffer t & 1
    const int bi = t & 1;
    mbar_wait(&bar_k[bi], (t >> 1) & 1);
    tc_fence_after();
    mma_kk<DH>(tmem_s + bi * 128, sQ, sK + bi * L::TILE, 128, false);
    umma_commit<1>(&bar_s[bi]);
  };
  if (issuer) {
    mbar_arrive_expect_tx(bar_q, L::TILE);
    load_tile<DH>(sQ, &tmQ, bar_q, row0, col0);
    load_k(0);
    load_k(1);  // n_tiles >= 2
    load_v(0);
    if (nj > 1) load_v(1);
    mbar_wait(bar_q, 0);
    issue_s(0);
  }

  const int r = threadIdx.x & 127, half = (threadIdx.x >> 7) & 1;  // row threads: accumulator row and column half
  const int qi = qb * QB + r;                                      // position of this thread's query in its sequence
  const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
  const float sl2 = p.scale * kLog2e;  // scale > 0 (checked on the host): max(scale * s) = scale * max(s)
  float m_raw = -INFINITY;             // pass 0: running maximum of this thread's half of the row, unscaled
  float m2 = 0.f, l = 0.f;             // pass 1: the row's maximum in the exp2 domain (0 if it sees no key), partial sum
  bool row_empty = false;

  for (int t = 0; t < n_tiles; ++t) {
    const int pass = t >= nj, u = pass ? t - nj : t, j = j_lo + u, bi = t & 1;
    const uint32_t par = (t >> 1) & 1;
    if (issuer) {
      mbar_wait(&bar_s[bi], par);           // S(t) is complete: its K buffer is free
      if (t + 2 < n_tiles) load_k(t + 2);
      if (t + 1 < n_tiles) issue_s(t + 1);  // S buffer (t+1) & 1 was last read for tile t-1, before that tile's sync
    } else if (warp < kIssuerWarp) {
      mbar_wait_poll(&bar_s[bi], par);
      tc_fence_after();
      const uint8_t* mr = (mrow != nullptr && !blk_ok[j]) ? mrow : nullptr;
      const bool all_vis = tile_all_visible(qb, j, p.window, mr);
      const int key0 = j * QB + half * 64;
      float v[64];
      tmem_ld64(tmem_s + bi * 128 + lane_off + half * 64, v);
      tmem_ld_wait();
      if (!all_vis) {
        mask_raw(v, qi, key0, p.window, mr);
        mask_raw(v + 32, qi, key0 + 32, p.window, mr);
      }
      if (!pass) {
        float c0 = fmaxf(v[0], v[1]), c1 = fmaxf(v[2], v[3]);
#pragma unroll
        for (int e = 4; e < 64; e += 4) {
          c0 = fmaxf(c0, fmaxf(v[e], v[e + 1]));
          c1 = fmaxf(c1, fmaxf(v[e + 2], v[e + 3]));
        }
        m_raw = fmaxf(m_raw, fmaxf(c0, c1));
        if (t == nj - 1) stat[threadIdx.x] = m_raw;
      } else {
        uint32_t w[32];
        float l0 = 0.f, l1 = 0.f;
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          const float p0 = fast_exp2(fmaf(v[2 * e], sl2, -m2)), p1 = fast_exp2(fmaf(v[2 * e + 1], sl2, -m2));
          l0 += p0;
          l1 += p1;
          w[e] = pack2(p0, p1);
        }
        l += l0 + l1;
        if (u > 0) mbar_wait_poll(bar_o, (u - 1) & 1);  // P.V of the previous tile has read the P tile
#pragma unroll
        for (int q8 = 0; q8 < 8; ++q8) put8_kmajor(sP, r, half * 64 + q8 * 8, &w[q8 * 4]);
        fence_async_shared();  // P was written through the generic proxy; the MMA reads it through the async one
      }
    }
    tc_fence_before();
    __syncthreads();
    if (t == nj - 1 && warp < kIssuerWarp) {  // end of pass 0: the row's maximum from its two halves
      const float mm = fmaxf(stat[r], stat[r + 128]);
      row_empty = (mm == -INFINITY);  // a row that sees no key: every p becomes exp2(-inf - 0) = 0
      m2 = row_empty ? 0.f : mm * sl2;
    }
    if (pass && issuer) {
      if (u > 0) mbar_wait(bar_o, (u - 1) & 1);  // (already complete: the row threads waited for it) V buffer (u+1) & 1 is free
      if (u >= 1 && u + 1 < nj) load_v(u + 1);
      mbar_wait(&bar_v[u & 1], (u >> 1) & 1);
      tc_fence_after();
      mma_k_mn<DH>(tmem_o, sP, sV + (u & 1) * L::TILE, u != 0);  // first tile of pass 1 overwrites
      umma_commit<1>(bar_o);
    }
  }

  if (warp < kIssuerWarp) {
    stat[threadIdx.x] = l;  // (the maxima were read a whole pass ago)
    mbar_wait_poll(bar_o, (nj - 1) & 1);  // the last P.V: O is complete
    tc_fence_after();
  }
  __syncthreads();
  if (warp < kIssuerWarp) {  // O row / sum -> bf16 -> [batch * T, heads * DH]
    const float lsum = stat[r] + stat[r + 128];
    const float inv = lsum > 0.f ? 1.f / lsum : 0.f;
    if (half == 0)
      p.lse[static_cast<size_t>(bh) * p.T + qi] = (row_empty || !(lsum > 0.f)) ? INFINITY : (m2 + log2f(lsum)) * kLn2;
    const uint32_t taddr = tmem_o + lane_off;
    __nv_bfloat16* dst = p.out0 + static_cast<size_t>(row0 + r) * (p.H * DH) + col0;
#pragma unroll 1
    for (int c = half; c < DH / 32; c += 2) {
      float v[32];
      tmem_ld32(taddr + c * 32, v);
      tmem_ld_wait();
#pragma unroll
      for (int q8 = 0; q8 < 4; ++q8)
        *reinterpret_cast<uint4*>(dst + c * 32 + q8 * 8) =
            make_uint4(pack2(v[q8 * 8] * inv, v[q8 * 8 + 1] * inv), pack2(v[q8 * 8 + 2] * inv, v[q8 * 8 + 3] * inv),
                       pack2(v[q8 * 8 + 4] * inv, v[q8 * 8 + 5] * inv), pack2(v[q8 * 8 + 6] * inv, v[q8 * 8 + 7] * inv));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kIssuerWarp) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_s, 512);
  }
}

// ------------------------------------------------------------------------------------------------ backward
// delta[b, h, t] = sum_c dO[row, h * DH + c] * O[row, h * DH + c]: one thread per (row, head).
template <int DH>
__global__ void __launch_bounds__(256) attn_delta_kernel(const __nv_bfloat16* __restrict__ o,
                                                         const __nv_bfloat16* __restrict__ d_o, float* __restrict__ delta,
                                                         int n_rows, int T, int H) {
  const int idx = blockIdx.x * 256 + threadIdx.x;
  if (idx >= n_rows * H) return;
  const int row = idx / H, h = idx - row * H;
  const uint4* a = reinterpret_cast<const uint4*>(o + static_cast<size_t>(row) * H * DH + h * DH);
  const uint4* b = reinterpret_cast<const uint4*>(d_o + static_cast<size_t>(row) * H * DH + h * DH);
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < DH / 8; ++i) {
    const uint4 x = a[i], y = b[i];
    const uint32_t xw[4] = {x.x, x.y, x.z, x.w}, yw[4] = {y.x, y.y, y.z, y.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const __nv_bfloat162 xh = *reinterpret_cast<const __nv_bfloat162*>(&xw[e]);
      const __nv_bfloat162 yh = *reinterpret_cast<const __nv_bfloat162*>(&yw[e]);
      acc += __low2float(xh) * __low2float(yh) + __high2float(xh) * __high2float(yh);
    }
  }
  const int bb = row / T, t = row - bb * T;
  delta[(static_cast<size_t>(bb) * H + h) * T + t] = acc;
}

template <int DH>
struct BwdSmem {
  static constexpr uint32_t TILE = QB * DH * 2;
  // stationary pair, streamed pair, then the two [128 x 128] tiles the row threads write (P and dS)
  static constexpr uint32_t A_OFF = 0, B_OFF = TILE, C_OFF = 2 * TILE, D_OFF = 3 * TILE, P_OFF = 4 * TILE,
                            DS_OFF = P_OFF + 32768, BAR_OFF = DS_OFF + 32768;
  static constexpr uint32_t FLAG_OFF = BAR_OFF + 8 * 8 + 16;  // [kMaxKeyBlocks] bytes: key block holds no padding
  static constexpr uint32_t BYTES = 1024 + FLAG_OFF + kMaxKeyBlocks;
};

// One thread's 128 scores and 128 dP values of a tile -> P and dS (8 columns at a time, straight into shared memory).
// lse2 = log2-sum-exp of the row, dl = delta of the row.  kKMajor: layout of the dS tile (see put8_*).
template <bool kWriteP, bool kKMajor>
__device__ __forceinline__ void tile_p_ds(uint32_t taddr_s, uint32_t taddr_dp, int r, int qi, int key0, const AttnParams& p,
                                          const uint8_t* mrow, float sl2, float lse2, float dl, uint8_t* sP, uint8_t* sDS,
                                          bool all_visible, int half) {
#pragma unroll 1
  for (int c = 2 * half; c < 2 * half + 2; ++c) {
    float v[32], g[32];
    tmem_ld32(taddr_s + c * 32, v);
    tmem_ld32(taddr_dp + c * 32, g);
    tmem_ld_wait();
    mask_scores(v, qi, key0 + c * 32, p, mrow, sl2, all_visible);
#pragma unroll
    for (int q8 = 0; q8 < 4; ++q8) {
      uint32_t wp[4], wd[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int j = q8 * 8 + 2 * e;
        const float p0 = fast_exp2(v[j] - lse2), p1 = fast_exp2(v[j + 1] - lse2);
        wp[e] = pack2(p0, p1);
        wd[e] = pack2(p.scale * p0 * (g[j] - dl), p.scale * p1 * (g[j + 1] - dl));
      }
      if (kWriteP) put8_mnmajor(sP, r, c * 32 + q8 * 8, wp);
      if (kKMajor)
        put8_kmajor(sDS, r, c * 32 + q8 * 8, wd);
      else
        put8_mnmajor(sDS, r, c * 32 + q8 * 8, wd);
    }
  }
  fence_async_shared();
}

// Key-block stationary: dK and dV of one 128-key block of one (batch, head).
template <int DH>
__global__ void __launch_bounds__(kAttnThreads, 1)
    attn_bwd_kv_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                       const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                       const AttnParams p) {
  using L = BwdSmem<DH>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem + L::A_OFF;
  uint8_t* sV = smem + L::B_OFF;
  uint8_t* sQ = smem + L::C_OFF;
  uint8_t* sDO = smem + L::D_OFF;
  uint8_t* sP = smem + L::P_OFF;
  uint8_t* sDS = smem + L::DS_OFF;
  uint64_t* bar_kv = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
  uint64_t* bar_q = bar_kv + 1;
  uint64_t* bar_s = bar_kv + 2;
  uint64_t* bar_o = bar_kv + 3;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_kv + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kb = blockIdx.x, bh = blockIdx.y, b = bh / p.H, h = bh - b * p.H;
  const int col0 = h * DH, n_blk = p.T / QB;
  const bool issuer = (warp == kIssuerWarp && lane == 0);
  // Lanes 1-31 of the issuing warp take no part in mbarrier waits: spinning beside the one lane that has to issue
  // the work the barrier is waiting for starves it.  They only join the __syncthreads.
  const bool waits = (warp < kIssuerWarp || lane == 0);
  if (issuer) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmDO);
    mbar_init(bar_kv, 1);
    mbar_init(bar_q, 1);
    mbar_init(bar_s, 1);
    mbar_init(bar_o, 1);
    fence_mbar_init();
  }
  const uint8_t* mrow = p.key_mask != nullptr ? p.key_mask + static_cast<size_t>(b) * p.T : nullptr;
  uint8_t* blk_ok = smem + L::FLAG_OFF;
  if (warp == kIssuerWarp) {
    tmem_alloc<1>(tmem_slot, 512);
    tmem_relinquish<1>();
  } else if (mrow != nullptr && warp == 0) {  // does this key block hold any padding?  (8 lanes x 16 mask bytes)
    bool ok = true;
    if (lane < 8) {
      const uint4 mk = *reinterpret_cast<const uint4*>(mrow + kb * QB + lane * 16);
      const uint32_t w[4] = {mk.x, mk.y, mk.z, mk.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) ok = ok && (((w[e] - 0x01010101u) & ~w[e] & 0x80808080u) == 0u);
    }
    ok = __all_sync(0xffffffffu, ok);
    if (lane == 0) blk_ok[0] = ok ? 1 : 0;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_s = *tmem_slot, tmem_dp = tmem_s + 128, tmem_dv = tmem_s + 256, tmem_dk = tmem_s + 384;
  if (mrow != nullptr && blk_ok[0]) mrow = nullptr;  // no padding among this block's keys: geometric tests only

  int i_lo, i_hi;
  q_range(kb, p.window, n_blk, i_lo, i_hi);
  if (issuer) {
    mbar_arrive_expect_tx(bar_kv, 2 * L::TILE);
    load_tile<DH>(sK, &tmK, bar_kv, b * p.T + kb * QB, col0);
    load_tile<DH>(sV, &tmV, bar_kv, b * p.T + kb * QB, col0);
  }
  const int r = threadIdx.x & 127, half = threadIdx.x >> 7;
  const float sl2 = p.scale * kLog2e;
  uint32_t ph = 0;
  if (waits) mbar_wait(bar_kv, 0);

  for (int i = i_lo; i <= i_hi; ++i) {
    if (issuer) {
      mbar_arrive_expect_tx(bar_q, 2 * L::TILE);  // the previous tile's MMAs have completed: both buffers are free
      load_tile<DH>(sQ, &tmQ, bar_q, b * p.T + i * QB, col0);
      load_tile<DH>(sDO, &tmDO, bar_q, b * p.T + i * QB, col0);
      mbar_wait(bar_q, ph);
      tc_fence_after();
      mma_kk<DH>(tmem_s, sQ, sK, 128, false);     // S  = Q_i . K^T     [queries x keys]
      mma_kk<DH>(tmem_dp, sDO, sV, 128, false);   // dP = dO_i . V^T
      umma_commit<1>(bar_s);
    }
    if (waits) mbar_wait_poll(bar_s, ph);
    tc_fence_after();
    if (warp < kIssuerWarp) {
      const int qi = i * QB + r;
      const size_t sidx = static_cast<size_t>(bh) * p.T + qi;
      const float lse = p.lse[sidx];
      const float lse2 = lse * kLog2e;  // +inf for a row that saw no key: every p is 0
      const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
      tile_p_ds<true, false>(tmem_s + lane_off, tmem_dp + lane_off, r, qi, kb * QB, p, mrow, sl2, lse2, p.delta[sidx], sP,
                             sDS, tile_all_visible(i, kb, p.window, mrow), half);
    }
    tc_fence_before();
    __syncthreads();
    if (issuer) {
      tc_fence_after();
      mma_mn_mn<DH>(tmem_dv, sP, sDO, i != i_lo);  // dV += P^T . dO_i
      mma_mn_mn<DH>(tmem_dk, sDS, sQ, i != i_lo);  // dK += dS^T . Q_i
      umma_commit<1>(bar_o);
    }
    if (waits) mbar_wait_poll(bar_o, ph);
    tc_fence_after();
    ph ^= 1;
  }

  if (warp < kIssuerWarp) {  // accumulator lane = key row
    const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const size_t row = static_cast<size_t>(b) * p.T + kb * QB + r;
#pragma unroll 1
    for (int which = 0; which < 2; ++which) {
      __nv_bfloat16* dst = (which == 0 ? p.out0 : p.out1) + row * (p.H * DH) + col0;
      const uint32_t taddr = (which == 0 ? tmem_dk : tmem_dv) + lane_off;
#pragma unroll 1
      for (int c = half; c < DH / 32; c += 2) {
        float v[32];
        tmem_ld32(taddr + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int q8 = 0; q8 < 4; ++q8)
          *reinterpret_cast<uint4*>(dst + c * 32 + q8 * 8) =
              make_uint4(pack2(v[q8 * 8], v[q8 * 8 + 1]), pack2(v[q8 * 8 + 2], v[q8 * 8 + 3]),
                         pack2(v[q8 * 8 + 4], v[q8 * 8 + 5]), pack2(v[q8 * 8 + 6], v[q8 * 8 + 7]));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kIssuerWarp) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_s, 512);
  }
}

// Query-block stationary: dQ of one 128-query block of one (batch, head).
template <int DH>
__global__ void __launch_bounds__(kAttnThreads, 1)
    attn_bwd_q_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                      const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                      const AttnParams p) {
  using L = BwdSmem<DH>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem + L::A_OFF;
  uint8_t* sDO = smem + L::B_OFF;
  uint8_t* sK = smem + L::C_OFF;
  uint8_t* sV = smem + L::D_OFF;
  uint8_t* sDS = smem + L::DS_OFF;
  uint64_t* bar_q = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
  uint64_t* bar_kv = bar_q + 1;
  uint64_t* bar_s = bar_q + 2;
  uint64_t* bar_o = bar_q + 3;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_q + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qb = static_cast<int>(gridDim.x) - 1 - static_cast<int>(blockIdx.x);  // longest key ranges first
  const int bh = blockIdx.y, b = bh / p.H, h = bh - b * p.H;
  const int col0 = h * DH;
  const bool issuer = (warp == kIssuerWarp && lane == 0);
  // Lanes 1-31 of the issuing warp take no part in mbarrier waits: spinning beside the one lane that has to issue
  // the work the barrier is waiting for starves it.  They only join the __syncthreads.
  const bool waits = (warp < kIssuerWarp || lane == 0);
  if (issuer) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmDO);
    mbar_init(bar_q, 1);
    mbar_init(bar_kv, 1);
    mbar_init(bar_s, 1);
    mbar_init(bar_o, 1);
    fence_mbar_init();
  }
  const uint8_t* mrow = p.key_mask != nullptr ? p.key_mask + static_cast<size_t>(b) * p.T : nullptr;
  uint8_t* blk_ok = smem + L::FLAG_OFF;
  if (warp == kIssuerWarp) {
    tmem_alloc<1>(tmem_slot, 512);
    tmem_relinquish<1>();
  } else if (mrow != nullptr) {
    key_block_flags(mrow, p.T, blk_ok, threadIdx.x, lane);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_s = *tmem_slot, tmem_dp = tmem_s + 128, tmem_dq = tmem_s + 256;

  int j_lo, j_hi;
  kv_range(qb, p.window, j_lo, j_hi);
  if (issuer) {
    mbar_arrive_expect_tx(bar_q, 2 * L::TILE);
    load_tile<DH>(sQ, &tmQ, bar_q, b * p.T + qb * QB, col0);
    load_tile<DH>(sDO, &tmDO, bar_q, b * p.T + qb * QB, col0);
  }
  const int r = threadIdx.x & 127, half = threadIdx.x >> 7, qi = qb * QB + r;
  const float sl2 = p.scale * kLog2e;
  float lse2 = 0.f, dl = 0.f;
  if (warp < kIssuerWarp) {
    const size_t sidx = static_cast<size_t>(bh) * p.T + qi;
    lse2 = p.lse[sidx] * kLog2e;
    dl = p.delta[sidx];
  }
  uint32_t ph = 0;
  if (waits) mbar_wait(bar_q, 0);

  for (int j = j_lo; j <= j_hi; ++j) {
    if (issuer) {
      mbar_arrive_expect_tx(bar_kv, 2 * L::TILE);
      load_tile<DH>(sK, &tmK, bar_kv, b * p.T + j * QB, col0);
      load_tile<DH>(sV, &tmV, bar_kv, b * p.T + j * QB, col0);
      mbar_wait(bar_kv, ph);
      tc_fence_after();
      mma_kk<DH>(tmem_s, sQ, sK, 128, false);    // S  = Q . K_j^T
      mma_kk<DH>(tmem_dp, sDO, sV, 128, false);  // dP = dO . V_j^T
      umma_commit<1>(bar_s);
    }
    if (waits) mbar_wait_poll(bar_s, ph);
    tc_fence_after();
    if (warp < kIssuerWarp) {
      const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
      const uint8_t* mr = (mrow != nullptr && !blk_ok[j]) ? mrow : nullptr;
      tile_p_ds<false, true>(tmem_s + lane_off, tmem_dp + lane_off, r, qi, j * QB, p, mr, sl2, lse2, dl, nullptr, sDS,
                             tile_all_visible(qb, j, p.window, mr), half);
    }
    tc_fence_before();
    __syncthreads();
    if (issuer) {
      tc_fence_after();
      mma_k_mn<DH>(tmem_dq, sDS, sK, j != j_lo);  // dQ += dS . K_j
      umma_commit<1>(bar_o);
    }
    if (waits) mbar_wait_poll(bar_o, ph);
    tc_fence_after();
    ph ^= 1;
  }

  if (warp < kIssuerWarp) {
    const uint32_t taddr = tmem_dq + (static_cast<uint32_t>((warp & 3) * 32) << 16);
    __nv_bfloat16* dst = p.out0 + (static_cast<size_t>(b) * p.T + qi) * (p.H * DH) + col0;
#pragma unroll 1
    for (int c = half; c < DH / 32; c += 2) {
      float v[32];
      tmem_ld32(taddr + c * 32, v);
      tmem_ld_wait();
#pragma unroll
      for (int q8 = 0; q8 < 4; ++q8)
        *reinterpret_cast<uint4*>(dst + c * 32 + q8 * 8) =
            make_uint4(pack2(v[q8 * 8], v[q8 * 8 + 1]), pack2(v[q8 * 8 + 2], v[q8 * 8 + 3]),
                       pack2(v[q8 * 8 + 4], v[q8 * 8 + 5]), pack2(v[q8 * 8 + 6], v[q8 * 8 + 7]));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kIssuerWarp) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_s, 512);
  }
}

// ------------------------------------------------------------------------------------------------ host side
int check_attn(const void* q, const void* k, const void* v, int B, int T, int H, int DH, float scale,
               ftk::DeviceInfo* di) {
  FTK_REQUIRE(q && k && v, "ftk_attention: null Q, K or V");
  FTK_REQUIRE(B > 0 && H > 0 && T > 0 && T % QB == 0, "ftk_attention: T=%d must be a positive multiple of %d (B=%d, H=%d)", T,
              QB, B, H);
  FTK_REQUIRE(DH == 64 || DH == 128, "ftk_attention: head_dim %d is not 64 or 128", DH);
  FTK_REQUIRE(static_cast<int64_t>(B) * T < (int64_t(1) << 31), "ftk_attention: batch * T too large");
  FTK_REQUIRE(T <= kMaxKeyBlocks * QB, "ftk_attention: T=%d exceeds %d", T, kMaxKeyBlocks * QB);
  FTK_REQUIRE(scale > 0.f, "ftk_attention: scale must be positive, not %g", static_cast<double>(scale));
  if (int rc = ftk::get_device_info(di)) return rc;
  if (di->cc_major != 10)
    return ftk::fail(FTK_EUNSUPPORTED, "ftk_attention needs an sm_100 device (tcgen05/TMEM); found sm_%d%d", di->cc_major,
                     di->cc_minor);
  return FTK_OK;
}

int attn_tmap(CUtensorMap* tm, const void* base, int B, int T, int H, int DH) {
  return ftk::make_tmap(tm, base, static_cast<uint64_t>(H) * DH, static_cast<uint64_t>(B) * T,
                        static_cast<uint64_t>(H) * DH * 2, 64, 64);
}

template <class Kern>
int set_smem(Kern kern, uint32_t bytes) {
  FTK_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes)));
  return FTK_OK;
}

template <int DH>
int run_attn_fwd(const void* q, const void* k, const void* v, const AttnParams& p, int B, cudaStream_t st) {
  CUtensorMap tq, tk, tv;
  if (int rc = attn_tmap(&tq, q, B, p.T, p.H, DH)) return rc;
  if (int rc = attn_tmap(&tk, k, B, p.T, p.H, DH)) return rc;
  if (int rc = attn_tmap(&tv, v, B, p.T, p.H, DH)) return rc;
  if (int rc = set_smem(attn_fwd_kernel<DH>, FwdSmem<DH>::BYTES)) return rc;
  ftk::TimelineScope tl_(FTK_TL_ATTN_FWD, st);
  attn_fwd_kernel<DH><<<dim3(p.T / QB, B * p.H), kAttnThreads, FwdSmem<DH>::BYTES, st>>>(tq, tk, tv, p);
  FTK_CUDA_CHECK(cudaGetLastError());
  return FTK_OK;
}

template <int DH>
int run_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* d_o, AttnParams p, int B,
                 float* delta, void* dq, void* dk, void* dv, cudaStream_t st) {
  CUtensorMap tq, tk, tv, tdo;
  if (int rc = attn_tmap(&tq, q, B, p.T, p.H, DH)) return rc;
  if (int rc = attn_tmap(&tk, k, B, p.T, p.H, DH)) return rc;
  if (int rc = attn_tmap(&tv, v, B, p.T, p.H, DH)) return rc;
  if (int rc = attn_tmap(&tdo, d_o, B, p.T, p.H, DH)) return rc;
  if (int rc = set_smem(attn_bwd_kv_kernel<DH>, BwdSmem<DH>::BYTES)) return rc;
  if (int rc = set_smem(attn_bwd_q_kernel<DH>, BwdSmem<DH>::BYTES)) return rc;
  ftk::TimelineScope tl_(FTK_TL_ATTN_BWD, st);
  const int n = B * p.T * p.H;
  attn_delta_kernel<DH><<<(n + 255) / 256, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(o),
                                                        static_cast<const __nv_bfloat16*>(d_o), delta, B * p.T, p.T, p.H);
  FTK_CUDA_CHECK(cudaGetLastError());
  p.delta = delta;
  const dim3 grid(p.T / QB, B * p.H);
  p.out0 = static_cast<__nv_bfloat16*>(dk);
  p.out1 = static_cast<__nv_bfloat16*>(dv);
  attn_bwd_kv_kernel<DH><<<grid, kAttnThreads, BwdSmem<DH>::BYTES, st>>>(tq, tk, tv, tdo, p);
  FTK_CUDA_CHECK(cudaGetLastError());
  p.out0 = static_cast<__nv_bfloat16*>(dq);
  p.out1 = nullptr;
  attn_bwd_q_kernel<DH><<<grid, kAttnThreads, BwdSmem<DH>::BYTES, st>>>(tq, tk, tv, tdo, p);
  FTK_CUDA_CHECK(cudaGetLastError());
  return FTK_OK;
}

}  // namespace

extern "C" int ftk_attention_fwd(const void* Q, const void* K, const void* V, const uint8_t* key_mask, int32_t B, int32_t T,
                                 int32_t H, int32_t head_dim, int32_t window, float scale, void* O, float* lse,
                                 void* stream) {
  ftk::DeviceInfo di;
  if (int rc = check_attn(Q, K, V, B, T, H, head_dim, scale, &di)) return rc;
  FTK_REQUIRE(O && lse, "ftk_attention_fwd: null output");
  FTK_REQUIRE(window >= 0, "ftk_attention_fwd: negative window");
  AttnParams p = {};
  p.key_mask = key_mask;
  p.lse = lse;
  p.out0 = static_cast<__nv_bfloat16*>(O);
  p.T = T;
  p.H = H;
  p.window = window;
  p.scale = scale;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return head_dim == 128 ? run_attn_fwd<128>(Q, K, V, p, B, st) : run_attn_fwd<64>(Q, K, V, p, B, st);
}

extern "C" size_t ftk_attention_bwd_workspace(int32_t B, int32_t T, int32_t H) {
  if (B <= 0 || T <= 0 || H <= 0) return 0;
  return static_cast<size_t>(B) * T * H * sizeof(float);
}

extern "C" int ftk_attention_bwd(const void* Q, const void* K, const void* V, const void* O, const void* dO,
                                 const float* lse, const uint8_t* key_mask, int32_t B, int32_t T, int32_t H,
                                 int32_t head_dim, int32_t window, float scale, void* dQ, void* dK, void* dV,
                                 void* workspace, size_t workspace_bytes, void* stream) {
  ftk::DeviceInfo di;
  if (int rc = check_attn(Q, K, V, B, T, H, head_dim, scale, &di)) return rc;
  FTK_REQUIRE(O && dO && lse && dQ && dK && dV && workspace, "ftk_attention_bwd: null pointer argument");
  FTK_REQUIRE(window >= 0, "ftk_attention_bwd: negative window");
  if (workspace_bytes < ftk_attention_bwd_workspace(B, T, H))
    return ftk::fail(FTK_EWORKSPACE, "ftk_attention_bwd: workspace %zu bytes < required %zu", workspace_bytes,
                     ftk_attention_bwd_workspace(B, T, H));
  AttnParams p = {};
  p.key_mask = key_mask;
  p.lse = const_cast<float*>(lse);
  p.T = T;
  p.H = H;
  p.window = window;
  p.scale = scale;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* delta = static_cast<float*>(workspace);
  return head_dim == 128 ? run_attn_bwd<128>(Q, K, V, O, dO, p, B, delta, dQ, dK, dV, st)
                         : run_attn_bwd<64>(Q, K, V, O, dO, p, B, delta, dQ, dK, dV, st);
}
