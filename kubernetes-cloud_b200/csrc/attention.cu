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
//   pass 0   S = Q.K^T (tcgen05.mma, accumulator in TMEM) -> masked row maximum (no exponentials)
//   pass 1   S again -> P = exp2(S - max) as bf16 into swizzled shared memory, row sums in fp32 -> O += P.V (second
//            tcgen05.mma); O / sum and lse at the end
// Backward, two kernels, no atomics, over 64-wide sub-tiles of the streamed operand:
//   kv-stationary   for a key block, transposed (accumulator lanes = keys): over the query sub-blocks that see it,
//                   S^T = K.Q_i^T and dP^T = V.dO_i^T -> P^T, dS^T = P^T * (dP^T - D) -> dV += P^T.dO_i, dK += dS^T.Q_i
//   q-stationary    for a query block: S = Q.K_j^T, dP = dO.V_j^T again -> dS -> dQ += dS.K_j
// The extra S / dP recomputations cost tensor flops that this workload has to spare (attention is ~4 % of a step's
// tensor work); what mattered was removing the fp32 score tensors.
//
// Every mbarrier wait here is a pure poll (mbarrier.test_wait): the suspending form (try_wait + NANOSLEEP.SYNCS) wakes
// up late enough to stall these pipelines on their issuing warps.
//
// Two warps issue TMA and MMA (an elected lane each); eight warps own the accumulator rows, two threads per row
// (softmax, tile conversion).  All three kernels are software pipelines: S (and dP) have two TMEM buffers, so the tensor pipe forms
// the scores of tile t+1 while the row threads work on tile t and applies tile t's P / dS to the accumulators while
// they work on tile t+1; operands arrive through multi-stage TMA rings.  A tile that lies wholly inside the causal /
// window band and holds no padding skips every per-element test; the others use one visibility bit mask per row.
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
// Warp 8: TMA loads + the score MMAs (S, dP), TMEM allocation.  Warp 9: the MMAs that consume what the row threads wrote
// (P.V, the gradient accumulations).  Two issuing warps because a tcgen05.mma issue is paced by the tensor pipe: one
// warp doing both spent a tile's MMA time twice in series, with its barrier waits and TMA issue on top.
constexpr int kAttnThreads = 320;
constexpr int kRowThreads = 256;
constexpr int kIssuerWarp = 8;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr int kMaxKeyBlocks = 256;  // per-key-block padding flags kept in shared memory: T <= 32768 with a key mask

struct AttnParams {
  const uint8_t* key_mask;  // [batch * T] (1 = may be attended) or null
  float* lse;               // [batch * heads * T]
  const float2* stats;      // backward: (lse * log2(e), scale * rowsum(dO * O)) [batch * heads * T]
  __nv_bfloat16* out0;      // forward: O; kv kernel: dK; q kernel: dQ     ([batch * T, heads * head_dim])
  __nv_bfloat16* out1;      // kv kernel: dV
  int T, H, window;         // window > 0: key j is visible from query i iff i - window < j <= i; 0: plain causal
  float scale;              // S = scale * Q.K^T
};

// Explicit shared-space accesses for the hot loops: the tile pointers are carved out of an aligned-up generic pointer,
// so plain dereferences compile to generic LD.E / ST.E with 64-bit addresses instead of LDS / STS.
__device__ __forceinline__ void sts128(uint32_t saddr, const uint32_t w[4]) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
}
__device__ __forceinline__ float4 lds128f(uint32_t saddr) {
  float4 r;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(saddr));
  return r;
}

__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

// Does every query of block qb see every key of block kb (so that no per-element test is needed)?
__device__ __forceinline__ bool tile_all_visible(int qb, int kb, int window, const uint8_t* mrow) {
  return mrow == nullptr && kb * QB + QB - 1 <= qb * QB && (window <= 0 || kb * QB > qb * QB + QB - 1 - window);
}

// First / last key block a query block needs (T % 128 == 0).
__device__ __forceinline__ void kv_range(int qb, int window, int& lo, int& hi) {
  hi = qb;
  lo = 0;
  if (window > 0) {
    const int first_key = qb * QB - window + 1;  // lowest key the block's first query sees
    lo = first_key > 0 ? first_key / QB : 0;
  }
}

// The MMA-issuing code is written for the compiler's uniform datapath: the whole issuing warp runs it in lockstep on
// warp-uniform values and one elected lane executes the tcgen05 / TMA instructions.  Issued from a divergent
// `if (lane == 0)` region with per-thread values, every tcgen05.mma cost ~90 cycles of operand shuffling
// (ELECT / R2UR.BROADCAST loops), more than the MMA itself takes to execute.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P1;\n\telect.sync _|P1, 0xffffffff;\n\tselp.u32 %0, 1, 0, P1;\n\t}" : "=r"(pred));
  return pred != 0;
}
// The instructions one lane issues for the warp, each predicated on elect.sync inside its own asm block: the form the
// compiler turns into ELECT + one predicated UTCHMMA / UTMALDG / UTCBAR (inside an ordinary `if (lane == 0)` region it
// wraps every one of them in a loop over the active lanes).  Call them from the whole, converged issuing warp.
__device__ __forceinline__ void e_umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\telect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void e_umma_commit(uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred q;\n\telect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void e_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile(
      "{\n\t.reg .pred q;\n\telect.sync _|q, 0xffffffff;\n\t"
      "@q mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n\t}" ::"r"(smem_u32(bar)),
      "r"(bytes)
      : "memory");
}
__device__ __forceinline__ void e_tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "{\n\t.reg .pred q;\n\telect.sync _|q, 0xffffffff;\n\t"
      "@q cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n\t}" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void e_bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred q;\n\telect.sync _|q, 0xffffffff;\n\t"
      "@q cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n\t}" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(gsrc)), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// A SWIZZLE_128B shared-memory descriptor (ftk_ptx.cuh: make_sw128_desc) as its two words; stepping through a tile
// only adds (byte offset >> 4) to the low word (no carry: shared-memory addresses stay below 2^18).
__device__ __forceinline__ uint32_t desc_lo(uint32_t smem_addr, uint32_t lbo_bytes) {
  return ((smem_addr >> 4) & 0x3FFFu) | (((lbo_bytes >> 4) & 0x3FFFu) << 16);
}
constexpr uint32_t kDescHi = ((1024u >> 4) & 0x3FFFu) | (1u << 14) | (2u << 29);  // SBO 1024, version 1, SWIZZLE_128B
__device__ __forceinline__ uint64_t desc_at(uint32_t lo, uint32_t byte_off) {
  return (static_cast<uint64_t>(kDescHi) << 32) | (lo + (byte_off >> 4));
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
    for (int rr = 0; rr < 2; ++rr) e_tma_load_2d(dst + kk * 16384 + rr * 8192, tm, bar, col0 + kk * 64, row0 + rr * 64);
}

// D[tmem] (+)= A.B^T with both operands K-major tiles of [128 rows x KD] (KD = 64 or 128): S = Q.K^T, dP = dO.V^T.
template <int KD>
__device__ __forceinline__ void mma_kk(uint32_t tmem_d, const uint8_t* a, const uint8_t* b, uint32_t n, bool accumulate) {
  const uint32_t idesc = make_idesc_bf16(128, n, false, false);
  const uint32_t aa = desc_lo(smem_u32(a), 0), bb = desc_lo(smem_u32(b), 0);
#pragma unroll
  for (int kk = 0; kk < KD / 64; ++kk)
#pragma unroll
    for (int k = 0; k < 4; ++k)
      e_umma_bf16(tmem_d, desc_at(aa, kk * 16384 + k * 32), desc_at(bb, kk * 16384 + k * 32), idesc,
                   (accumulate || kk != 0 || k != 0) ? 1u : 0u);
}
// D[tmem] (+)= A.B, A K-major [128 x 128] (a P / dS tile written by the row threads), B a TMA-loaded [128 K rows x ND]
// tile in its MN-major view.
template <int ND>
__device__ __forceinline__ void mma_k_mn(uint32_t tmem_d, const uint8_t* a, const uint8_t* b, bool accumulate) {
  const uint32_t idesc = make_idesc_bf16(128, ND, false, true);
  const uint32_t aa = desc_lo(smem_u32(a), 0), bb = desc_lo(smem_u32(b), 16384);
#pragma unroll
  for (int kb = 0; kb < 2; ++kb)
#pragma unroll
    for (int k = 0; k < 4; ++k)
      e_umma_bf16(tmem_d, desc_at(aa, kb * 16384 + k * 32), desc_at(bb, kb * 8192 + k * 2048), idesc,
                   (accumulate || kb != 0 || k != 0) ? 1u : 0u);
}
// Row thread `r` writes 8 consecutive values (columns c0 .. c0+7 of its row) of a [128 x 128] bf16 tile in K-major
// form (the tile is an A operand with rows = M): half (c0 / 64) at +16 KB, row r at r * 128 B, 16-byte chunk c at
// position c ^ (row & 7) of its 128-byte line: the SWIZZLE_128B pattern TMA and UMMA use.
__device__ __forceinline__ void put8_kmajor(uint32_t tile_saddr, int r, int c0, const uint32_t w[4]) {
  sts128(tile_saddr + (c0 >> 6) * 16384 + r * 128 + ((((c0 & 63) >> 3) ^ (r & 7)) << 4), w);
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

// Bits j (0 .. 31) with lo <= j <= hi.
__device__ __forceinline__ uint32_t range_bits(int lo, int hi) {
  if (hi < 0 || lo > 31 || lo > hi) return 0u;
  lo = max(lo, 0);
  hi = min(hi, 31);
  return (0xffffffffu >> (31 - hi)) & (0xffffffffu << lo);
}
// Which of the 32 keys key0 .. key0 + 31 may query qi attend?  Causal: key <= qi; window > 0: also key > qi - window;
// mrow != null: and the key is not padding.  One bit mask per row and chunk instead of two compares per element.
__device__ __forceinline__ uint32_t visible_bits(int qi, int key0, int window, const uint8_t* mrow) {
  uint32_t bits = range_bits(window > 0 ? qi - window + 1 - key0 : 0, qi - key0);
  if (mrow != nullptr) {
    const uint4 m0 = *reinterpret_cast<const uint4*>(mrow + key0), m1 = *reinterpret_cast<const uint4*>(mrow + key0 + 16);
    const uint32_t mw[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
    uint32_t mbits = 0;
#pragma unroll
    for (int j = 0; j < 32; ++j) mbits |= ((mw[j >> 2] >> ((j & 3) * 8)) & 0xffu) ? (1u << j) : 0u;
    bits &= mbits;
  }
  return bits;
}
// 32 raw scores of one query row -> -inf where the key must not be attended (the caller scales afterwards).
__device__ __forceinline__ void mask_raw(float* v, int qi, int key0, int window, const uint8_t* mrow) {
  const uint32_t bits = visible_bits(qi, key0, window, mrow);
#pragma unroll
  for (int j = 0; j < 32; ++j)
    if (!((bits >> j) & 1u)) v[j] = -INFINITY;
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
// they work on tile t+1 (they wait for it only at the point where they overwrite the P tile).  In pass 1 K and V each
// have two shared-memory buffers with their own barriers, refilled as soon as the MMA that read them has completed; in
// pass 0 all four serve K as one ring.
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
  uint64_t* bar_done = bar_q + 8;  // [2]: all 256 row threads are through with S buffer b (and, in pass 1, wrote P)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_q + 10);
  float* stat = reinterpret_cast<float*>(smem + L::STAT_OFF);
  uint8_t* blk_ok = smem + L::FLAG_OFF;

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;  // warp-uniform for the compiler
  // the query blocks that see the most keys start first, so the short ones fill the tail of the grid
  const int qb = static_cast<int>(gridDim.x) - 1 - static_cast<int>(blockIdx.x);
  const int bh = blockIdx.y, b = bh / p.H, h = bh - b * p.H;
  const int row0 = b * p.T + qb * QB;  // first row of this block in the [batch * T] matrices
  const int col0 = h * DH;
  const bool issue_warp = (warp == kIssuerWarp);  // runs the issuing code in lockstep (the e_* instructions elect a lane)
  const bool issuer = issue_warp && lane == 0;    // one-time set-up only
  const uint8_t* mrow = p.key_mask != nullptr ? p.key_mask + static_cast<size_t>(b) * p.T : nullptr;

  if (issuer) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    for (int i = 0; i < 8; ++i) mbar_init(bar_q + i, 1);
    mbar_init(&bar_done[0], kRowThreads);
    mbar_init(&bar_done[1], kRowThreads);
    fence_mbar_init();
  }
  if (warp == kIssuerWarp) {
    tmem_alloc<1>(tmem_slot, 512);
    tmem_relinquish<1>();
  } else if (mrow != nullptr && warp < kIssuerWarp) {
    key_block_flags(mrow, p.T, blk_ok, threadIdx.x, lane);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_s = __shfl_sync(0xffffffffu, *tmem_slot, 0), tmem_o = tmem_s + 256;  // S buffers at +0 and +128

  int j_lo, j_hi;
  kv_range(qb, p.window, j_lo, j_hi);
  const int nj = j_hi - j_lo + 1, n_tiles = 2 * nj;
  // (the lambdas below are called by the whole, converged issuing warp: the e_* instructions elect their own lane)
  // Pass 0 has no use for the V buffers, so K streams through all four tile buffers [K0 K1 V0 V1] as one ring: a K block
  // is requested four tiles ahead and the TMA latency stays hidden behind the 512 cycles of MMA a tile takes.  The four
  // barriers bar_k[0..1], bar_v[0..1] are contiguous and serve as the ring's; pass 1 goes on using them per buffer, its
  // phase counts starting where pass 0 left them (ring_uses).
  uint64_t* ring_bar = bar_k;
  auto ring_uses = [&](int s) { return nj > s ? ((nj - 1 - s) >> 2) + 1 : 0; };  // pass-0 tiles that used ring stage s
  auto load_ring = [&](int t) {  // pass 0: K block of tile t -> ring stage t & 3
    const int s = t & 3;
    e_arrive_expect_tx(&ring_bar[s], L::TILE);
    load_tile<DH>(sK + s * L::TILE, &tmK, &ring_bar[s], b * p.T + (j_lo + t) * QB, col0);
  };
  auto issue_s0 = [&](int t) {  // pass 0: S(t) = Q.K(t)^T into S buffer t & 1
    const int s = t & 3;
    mbar_wait_poll(&ring_bar[s], (t >> 2) & 1);
    tc_fence_after();
    mma_kk<DH>(tmem_s + (t & 1) * 128, sQ, sK + s * L::TILE, 128, false);
    e_umma_commit(&bar_s[t & 1]);
  };
  auto load_k1 = [&](int u) {  // pass 1: K block u -> K buffer u & 1
    const int bi = u & 1;
    e_arrive_expect_tx(&bar_k[bi], L::TILE);
    load_tile<DH>(sK + bi * L::TILE, &tmK, &bar_k[bi], b * p.T + (j_lo + u) * QB, col0);
  };
  auto load_v1 = [&](int u) {  // pass 1: V block u -> V buffer u & 1
    const int bi = u & 1;
    e_arrive_expect_tx(&bar_v[bi], L::TILE);
    load_tile<DH>(sV + bi * L::TILE, &tmV, &bar_v[bi], b * p.T + (j_lo + u) * QB, col0);
  };
  auto issue_s1 = [&](int u) {  // pass 1: S(nj + u) into S buffer (nj + u) & 1
    const int bi = u & 1, t = nj + u;
    mbar_wait_poll(&bar_k[bi], (ring_uses(bi) + (u >> 1)) & 1);
    tc_fence_after();
    mma_kk<DH>(tmem_s + (t & 1) * 128, sQ, sK + bi * L::TILE, 128, false);
    e_umma_commit(&bar_s[t & 1]);
  };
  if (issue_warp) {
    e_arrive_expect_tx(bar_q, L::TILE);
    load_tile<DH>(sQ, &tmQ, bar_q, row0, col0);
    for (int t = 0; t < 4 && t < nj; ++t) load_ring(t);
    mbar_wait_poll(bar_q, 0);
    issue_s0(0);
    if (nj > 1) issue_s0(1);
  }

  const int r = threadIdx.x & 127, half = (threadIdx.x >> 7) & 1;  // row threads: accumulator row and column half
  const int qi = qb * QB + r;                                      // position of this thread's query in its sequence
  const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
  const float sl2 = p.scale * kLog2e;  // scale > 0 (checked on the host): max(scale * s) = scale * max(s)
  const uint32_t sP_addr = smem_u32(sP);
  float m_raw = -INFINITY;             // pass 0: running maximum of this thread's half of the row, unscaled
  float m2 = 0.f, l = 0.f;             // pass 1: the row's maximum in the exp2 domain (0 if it sees no key), partial sum
  bool row_empty = false;

  // No block-wide barrier inside the loop: the row threads report each tile on bar_done and move on, so the eight
  // warps drift apart by up to a tile and one warp's TMEM-load / MUFU latencies hide behind another's arithmetic.
  if (issue_warp) {
    // pass 0: two S tiles in flight; S(t+2) goes into the TMEM buffer the row threads have just left
    for (int t = 0; t < nj; ++t) {
      mbar_wait_poll(&bar_s[t & 1], (t >> 1) & 1);  // S(t) is complete: its ring stage is free
      if (t + 4 < nj) load_ring(t + 4);
      if (t + 2 < nj) {
        mbar_wait_poll(&bar_done[t & 1], (t >> 1) & 1);  // tile t has left its S buffer
        issue_s0(t + 2);
      }
    }
    // every S of pass 0 is complete: the K buffers are free (the other warp refills the V buffers).  Pass 1: K through
    // two buffers, S one tile ahead of the row threads.
    load_k1(0);
    if (nj > 1) {
      load_k1(1);
      mbar_wait_poll(&bar_done[nj & 1], ((nj - 2) >> 1) & 1);  // tile nj-2 has left the S buffer S(nj) goes into
    }
    issue_s1(0);
    for (int u = 0; u + 1 < nj; ++u) {
      const int t = nj + u;
      mbar_wait_poll(&bar_s[t & 1], (t >> 1) & 1);  // S(t) is complete: K buffer u & 1 is free
      if (u + 2 < nj) load_k1(u + 2);
      mbar_wait_poll(&bar_done[(t + 1) & 1], ((t - 1) >> 1) & 1);  // tile t-1 has left that S buffer
      issue_s1(u + 1);
    }
  } else if (warp == kIssuerWarp + 1) {  // V ring + the P.V MMAs of pass 1
    // Pass 0: only follow the barriers.  A parity wait is valid only for the phase right after the last one this warp
    // saw complete (a wait for phase k on a barrier still in phase k-2 returns at once), so every phase is awaited.
    for (int t = 0; t < nj; ++t) {
      mbar_wait_poll(&bar_s[t & 1], (t >> 1) & 1);
      mbar_wait_poll(&bar_done[t & 1], (t >> 1) & 1);
    }
    // every S of pass 0 is complete and read: the V buffers (ring stages 2, 3) are free
    load_v1(0);
    if (nj > 1) load_v1(1);
    for (int u = 0; u < nj; ++u) {
      const int t = nj + u;
      mbar_wait_poll(&bar_done[t & 1], (t >> 1) & 1);  // P(u) is in shared memory
      if (u > 0) mbar_wait_poll(bar_o, (u - 1) & 1);   // P.V(u-1) done: V buffer (u + 1) & 1 is free
      if (u >= 1 && u + 1 < nj) load_v1(u + 1);
      mbar_wait_poll(&bar_v[u & 1], (ring_uses(2 + (u & 1)) + (u >> 1)) & 1);
      tc_fence_after();
      mma_k_mn<DH>(tmem_o, sP, sV + (u & 1) * L::TILE, u != 0);  // first tile of pass 1 overwrites
      e_umma_commit(bar_o);
    }
  } else {
    for (int t = 0; t < n_tiles; ++t) {
      const int pass = t >= nj, u = pass ? t - nj : t, j = j_lo + u, bi = t & 1;
      mbar_wait_poll(&bar_s[bi], (t >> 1) & 1);
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
        tc_fence_before();
        mbar_arrive(&bar_done[bi]);
        if (t == nj - 1) {  // end of pass 0: the row's maximum from its two halves
          stat[threadIdx.x] = m_raw;
          named_bar_sync(1, kRowThreads);
          const float mm = fmaxf(stat[r], stat[r + 128]);
          row_empty = (mm == -INFINITY);  // a row that sees no key: every p becomes exp2(-inf - 0) = 0
          m2 = row_empty ? 0.f : mm * sl2;
        }
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
        for (int q8 = 0; q8 < 8; ++q8) put8_kmajor(sP_addr, r, half * 64 + q8 * 8, &w[q8 * 4]);
        fence_async_shared();  // P was written through the generic proxy; the MMA reads it through the async one
        tc_fence_before();
        mbar_arrive(&bar_done[bi]);
      }
    }
    named_bar_sync(1, kRowThreads);  // every thread has read the maxima (a pass ago, but nothing else orders it)
    stat[threadIdx.x] = l;
    mbar_wait_poll(bar_o, (nj - 1) & 1);  // the last P.V: O is complete
    tc_fence_after();
    named_bar_sync(1, kRowThreads);
  }
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
      uint32_t w[16];
#pragma unroll
      for (int e = 0; e < 16; ++e) w[e] = pack2(v[2 * e] * inv, v[2 * e + 1] * inv);
      st_global_32B(dst + c * 32, w);  // a thread owns a row: 32-byte stores write whole sectors
      st_global_32B(dst + c * 32 + 16, w + 8);
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
// stats[b, h, t] = (lse * log2(e), scale * delta) with delta = sum_c dO[row, h * DH + c] * O[row, h * DH + c]: what the
// two backward kernels need per query (dS = P * (scale * dP - scale * delta)), packed so that one 512-byte bulk copy
// brings a sub-tile's 64 pairs.
template <int DH>
__global__ void __launch_bounds__(256) attn_delta_kernel(const __nv_bfloat16* __restrict__ o,
                                                         const __nv_bfloat16* __restrict__ d_o,
                                                         const float* __restrict__ lse, float2* __restrict__ stats,
                                                         int n_rows, int T, int H, float scale) {
  // DH / 8 lanes share one (row, head): a warp reads 512 contiguous bytes of O and of dO per instruction
  constexpr int kLanes = DH / 8, kItems = 32 / kLanes;
  const int lane = threadIdx.x & 31, sub = lane % kLanes;
  const int64_t warp_id = (static_cast<int64_t>(blockIdx.x) * 256 + threadIdx.x) >> 5;
  const int64_t idx = warp_id * kItems + lane / kLanes;  // (row, head) pairs are consecutive in memory: row * H + head
  const bool live = idx < static_cast<int64_t>(n_rows) * H;
  float acc = 0.f;
  if (live) {
    const uint4 x = __ldg(reinterpret_cast<const uint4*>(o + idx * DH) + sub);
    const uint4 y = __ldg(reinterpret_cast<const uint4*>(d_o + idx * DH) + sub);
    const uint32_t xw[4] = {x.x, x.y, x.z, x.w}, yw[4] = {y.x, y.y, y.z, y.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const __nv_bfloat162 xh = *reinterpret_cast<const __nv_bfloat162*>(&xw[e]);
      const __nv_bfloat162 yh = *reinterpret_cast<const __nv_bfloat162*>(&yw[e]);
      acc += __low2float(xh) * __low2float(yh) + __high2float(xh) * __high2float(yh);
    }
  }
#pragma unroll
  for (int off = kLanes / 2; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
  if (live && sub == 0) {
    const int row = static_cast<int>(idx / H), h = static_cast<int>(idx - static_cast<int64_t>(row) * H);
    const int bb = row / T, t = row - bb * T;
    const size_t i = (static_cast<size_t>(bb) * H + h) * T + t;
    stats[i] = make_float2(lse[i] * kLog2e, acc * scale);  // lse = +inf for a row that saw no key: every p is 0
  }
}

// The backward kernels work on 64-wide sub-tiles of the streamed operand so that everything can be double-buffered
// inside 227 KB of shared memory and 512 TMEM columns: while the row threads turn S / dP of sub-tile t into P / dS,
// the tensor pipe forms S / dP of sub-tile t+1 and applies sub-tile t-1's P / dS to the gradient accumulators, and TMA
// fetches the operands of sub-tile t+2 (three-stage ring).  M = 128, N = 64 MMAs run at the full tensor rate.
constexpr int SB = 64;  // rows of a streamed sub-tile

// TMA: a [64 rows x DH cols] sub-tile as 64 x 64 boxes, column half kk at +kk * 8 KB, row r at r * 128 B (SWIZZLE_128B).
template <int DH>
__device__ __forceinline__ void load_sub(uint8_t* dst, const CUtensorMap* tm, uint64_t* bar, int row0, int col0) {
#pragma unroll
  for (int kk = 0; kk < DH / 64; ++kk) e_tma_load_2d(dst + kk * 8192, tm, bar, col0 + kk * 64, row0);
}
// D[128 x 64] = A.B^T: A a [128 x KD] K-major tile (64-column halves 16 KB apart), B a [64 x KD] K-major sub-tile
// (halves 8 KB apart).
template <int KD>
__device__ __forceinline__ void mma_full_sub(uint32_t tmem_d, const uint8_t* a, const uint8_t* b) {
  const uint32_t idesc = make_idesc_bf16(128, SB, false, false);
  const uint32_t aa = desc_lo(smem_u32(a), 0), bb = desc_lo(smem_u32(b), 0);
#pragma unroll
  for (int kk = 0; kk < KD / 64; ++kk)
#pragma unroll
    for (int k = 0; k < 4; ++k)
      e_umma_bf16(tmem_d, desc_at(aa, kk * 16384 + k * 32), desc_at(bb, kk * 8192 + k * 32), idesc,
                   (kk != 0 || k != 0) ? 1u : 0u);
}
// D[128 x ND] (+)= A.B: A a [128 x 64] K-major tile the row threads wrote (one 128-byte line per row), B a TMA-loaded
// [64 K rows x ND] sub-tile in its MN-major view (16 K rows are 2 KB apart, the 64-column halves 8 KB).
template <int ND>
__device__ __forceinline__ void mma_k64_mn(uint32_t tmem_d, const uint8_t* a, const uint8_t* b, bool accumulate) {
  const uint32_t idesc = make_idesc_bf16(128, ND, false, true);
  const uint32_t aa = desc_lo(smem_u32(a), 0), bb = desc_lo(smem_u32(b), 8192);
#pragma unroll
  for (int k = 0; k < 4; ++k)
    e_umma_bf16(tmem_d, desc_at(aa, k * 32), desc_at(bb, k * 2048), idesc, (accumulate || k != 0) ? 1u : 0u);
}
// Row `r` of a [128 x 64] K-major tile: 8 consecutive bf16 values at columns c0 .. c0 + 7 (c0 < 64).
__device__ __forceinline__ void put8_row64(uint32_t tile_saddr, int r, int c0, const uint32_t w[4]) {
  sts128(tile_saddr + r * 128 + (((c0 >> 3) ^ (r & 7)) << 4), w);
}
// Per-64-key flags: the sub-block holds no padding.  Called by the 256 row threads.
__device__ __forceinline__ void key_sub_flags(const uint8_t* mrow, int T, uint8_t* sub_ok, int tid, int lane) {
  for (int base = 0; base < T; base += 256 * 16) {
    const int off = base + tid * 16;
    bool ok = true;
    if (off < T) {
      const uint4 mk = *reinterpret_cast<const uint4*>(mrow + off);
      const uint32_t w[4] = {mk.x, mk.y, mk.z, mk.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) ok = ok && (((w[e] - 0x01010101u) & ~w[e] & 0x80808080u) == 0u);
    }
    const uint32_t bal = __ballot_sync(0xffffffffu, ok);
    if ((lane & 3) == 0 && off < T) sub_ok[off >> 6] = (((bal >> lane) & 0xfu) == 0xfu) ? 1 : 0;
  }
}
// O / dO / dQ-style epilogue: this thread's accumulator row (columns c = half, half + 2, ... of 32) -> bf16 -> global.
template <int DH>
__device__ __forceinline__ void store_acc_row(uint32_t taddr, __nv_bfloat16* dst, int half) {
#pragma unroll 1
  for (int c = half; c < DH / 32; c += 2) {
    float v[32];
    tmem_ld32(taddr + c * 32, v);
    tmem_ld_wait();
    uint32_t w[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) w[e] = pack2(v[2 * e], v[2 * e + 1]);
    st_global_32B(dst + c * 32, w);  // a thread owns a row: 32-byte stores write whole sectors
    st_global_32B(dst + c * 32 + 16, w + 8);
  }
}

template <int DH>
struct BwdKvSmem {
  static constexpr uint32_t TILE = QB * DH * 2, SUB = SB * DH * 2;
  // stationary K and V, a three-stage ring of (Q_i, dO_i) sub-tiles, two buffers each of the P^T and dS^T tiles
  static constexpr uint32_t K_OFF = 0, V_OFF = TILE, Q_OFF = 2 * TILE, DO_OFF = Q_OFF + 3 * SUB, PT_OFF = DO_OFF + 3 * SUB,
                            DST_OFF = PT_OFF + 2 * 16384, BAR_OFF = DST_OFF + 2 * 16384;
  static constexpr uint32_t STAT_OFF = BAR_OFF + 128;  // [3][64] (log2-sum-exp, delta) of the queries of a sub-tile
  static constexpr uint32_t BYTES = 1024 + STAT_OFF + 3 * SB * 8;
  static_assert(BYTES <= 232448, "exceeds the 227 KB of shared memory a CTA can have");
};

// Key-block stationary: dK and dV of one 128-key block of one (batch, head).  Transposed formulation: the accumulator
// lanes are the KEYS (S^T = K.Q_i^T, dP^T = V.dO_i^T, both [128 keys x 64 queries]), so P^T and dS^T leave the row
// threads already in the K-major layout the two gradient MMAs want as their A operand (dV += P^T.dO_i,
// dK += dS^T.Q_i), and the per-query statistics are read from shared memory as broadcasts.
template <int DH>
__global__ void __launch_bounds__(kAttnThreads, 1)
    attn_bwd_kv_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                       const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                       const AttnParams p) {
  using L = BwdKvSmem<DH>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem + L::K_OFF;
  uint8_t* sV = smem + L::V_OFF;
  uint8_t* sQ = smem + L::Q_OFF;    // three stages
  uint8_t* sDO = smem + L::DO_OFF;  // three stages
  uint8_t* sPT = smem + L::PT_OFF;  // two buffers
  uint8_t* sDST = smem + L::DST_OFF;
  uint64_t* bar_kv = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
  uint64_t* bar_qd = bar_kv + 1;  // [3]
  uint64_t* bar_s = bar_kv + 4;   // [2]
  uint64_t* bar_o = bar_kv + 6;   // [2]
  uint64_t* bar_done = bar_kv + 8;  // [2]: all 256 row threads have written P^T / dS^T of buffer b
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_kv + 10);
  float2* stat = reinterpret_cast<float2*>(smem + L::STAT_OFF);  // [3][64], filled by TMA with the (Q_i, dO_i) stage

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;  // warp-uniform for the compiler
  const int kb = blockIdx.x, bh = blockIdx.y, b = bh / p.H, h = bh - b * p.H;  // key block 0 sees the most queries
  const int col0 = h * DH;
  const bool issue_warp = (warp == kIssuerWarp);  // runs the issuing code in lockstep (the e_* instructions elect a lane)
  const bool issuer = issue_warp && lane == 0;    // one-time set-up only
  if (issuer) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmDO);
    for (int i = 0; i < 8; ++i) mbar_init(bar_kv + i, 1);
    mbar_init(&bar_done[0], kRowThreads);
    mbar_init(&bar_done[1], kRowThreads);
    fence_mbar_init();
  }
  // query sub-blocks that see this key block
  const int is_lo = 2 * kb;
  int last_q = p.T - 1;
  if (p.window > 0) last_q = min(last_q, kb * QB + QB - 1 + p.window - 1);
  const int n = last_q / SB - is_lo + 1;
  const float2* gstat = p.stats + static_cast<size_t>(bh) * p.T + static_cast<size_t>(is_lo) * SB;
  if (warp == kIssuerWarp) {
    tmem_alloc<1>(tmem_slot, 512);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0), tmem_dv = tmem_base + 256, tmem_dk = tmem_base + 384;

  auto load_qd = [&](int t) {  // (Q_i, dO_i) of sub-tile t -> ring stage t % 3
    const int st = t % 3, row = b * p.T + (is_lo + t) * SB;
    e_arrive_expect_tx(&bar_qd[st], 2 * L::SUB + SB * 8);
    load_sub<DH>(sQ + st * L::SUB, &tmQ, &bar_qd[st], row, col0);
    load_sub<DH>(sDO + st * L::SUB, &tmDO, &bar_qd[st], row, col0);
    e_bulk_load_1d(stat + st * SB, gstat + t * SB, SB * 8, &bar_qd[st]);  // the 64 queries' (log2-sum-exp, delta)
  };
  auto issue_s = [&](int t) {  // S^T(t) and dP^T(t) into TMEM buffer t & 1
    const int st = t % 3;
    mbar_wait_poll(&bar_qd[st], (t / 3) & 1);
    tc_fence_after();
    mma_full_sub<DH>(tmem_base + (t & 1) * 128, sK, sQ + st * L::SUB);
    mma_full_sub<DH>(tmem_base + (t & 1) * 128 + SB, sV, sDO + st * L::SUB);
    e_umma_commit(&bar_s[t & 1]);
  };
  if (issue_warp) {  // (the lambdas are called by the whole, converged issuing warp: the e_* instructions elect their own lane)
    e_arrive_expect_tx(bar_kv, 2 * L::TILE);
    load_tile<DH>(sK, &tmK, bar_kv, b * p.T + kb * QB, col0);
    load_tile<DH>(sV, &tmV, bar_kv, b * p.T + kb * QB, col0);
    load_qd(0);
    if (n > 1) load_qd(1);
    mbar_wait_poll(bar_kv, 0);
    issue_s(0);
  }

  const int r = threadIdx.x & 127, half = (threadIdx.x >> 7) & 1;  // row threads: key row, half of the 64 query columns
  const int kj = kb * QB + r;                                      // this thread's key
  const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
  const float sl2 = p.scale * kLog2e;
  bool key_ok = true;  // a padded key is never attended: its whole row of P is 0
  if (warp < kIssuerWarp && p.key_mask != nullptr) key_ok = p.key_mask[static_cast<size_t>(b) * p.T + kj] != 0;

  // No block-wide barrier inside the loop (see the forward kernel): the row threads report on bar_done and move on.
  if (issue_warp) {  // operand ring + score MMAs, one sub-tile ahead of the row threads
    for (int t = 0; t + 1 < n; ++t) {
      if (t >= 1) {
        mbar_wait_poll(&bar_o[(t - 1) & 1], ((t - 1) >> 1) & 1);     // gradient MMAs of t-1 done: ring stage (t + 2) % 3 is free
        mbar_wait_poll(&bar_done[(t + 1) & 1], ((t - 1) >> 1) & 1);  // row threads have left TMEM buffer (t + 1) & 1
      }
      if (t + 2 < n) load_qd(t + 2);
      issue_s(t + 1);
    }
  } else if (warp == kIssuerWarp + 1) {  // gradient MMAs
    for (int t = 0; t < n; ++t) {
      const int sb = t & 1, st = t % 3;
      mbar_wait_poll(&bar_qd[st], (t / 3) & 1);     // (complete) this warp's own acquire of the stage's TMA writes
      mbar_wait_poll(&bar_done[sb], (t >> 1) & 1);  // P^T(t) and dS^T(t) are in shared memory
      tc_fence_after();
      mma_k64_mn<DH>(tmem_dv, sPT + sb * 16384, sDO + st * L::SUB, t != 0);  // dV += P^T . dO_i
      mma_k64_mn<DH>(tmem_dk, sDST + sb * 16384, sQ + st * L::SUB, t != 0);  // dK += dS^T . Q_i
      e_umma_commit(&bar_o[sb]);
    }
  } else {
    for (int t = 0; t < n; ++t) {
      const int sb = t & 1, st = t % 3;
      mbar_wait_poll(&bar_s[sb], (t >> 1) & 1);
      mbar_wait_poll(&bar_qd[st], (t / 3) & 1);  // (complete) acquire of this stage's TMA writes: the queries' statistics
      tc_fence_after();
      float v[32], g[32];
      tmem_ld32(tmem_base + sb * 128 + lane_off + half * 32, v);
      tmem_ld32(tmem_base + sb * 128 + SB + lane_off + half * 32, g);
      tmem_ld_wait();
      const int q0 = (is_lo + t) * SB + half * 32;  // first query of this thread's 32 columns
      // every query of the sub-tile sees every key of the block?  (then only the padding flag of the key matters)
      const bool all_vis = (is_lo + t) * SB >= kb * QB + QB - 1 &&
                           (p.window <= 0 || (is_lo + t) * SB + SB - 1 - p.window < kb * QB);
      // (log2-sum-exp, scale * delta) of this thread's 32 queries, two queries per 16-byte broadcast read
      const uint32_t sq = smem_u32(stat + st * SB + half * 32);
      uint32_t wp[16], wd[16];
      if (all_vis && key_ok) {
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const float4 sd = lds128f(sq + e * 16);
          const float p0 = fast_exp2(fmaf(v[2 * e], sl2, -sd.x)), p1 = fast_exp2(fmaf(v[2 * e + 1], sl2, -sd.z));
          wp[e] = pack2(p0, p1);
          wd[e] = pack2(p0 * fmaf(g[2 * e], p.scale, -sd.y), p1 * fmaf(g[2 * e + 1], p.scale, -sd.w));
        }
      } else {
        // queries q0 + c that see key kj: q >= kj, and q < kj + window for a sliding window; none if the key is padding
        uint32_t bits = all_vis ? 0xffffffffu : range_bits(kj - q0, p.window > 0 ? kj + p.window - 1 - q0 : 31);
        if (!key_ok) bits = 0u;
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const float4 sd = lds128f(sq + e * 16);
          float p0 = fast_exp2(fmaf(v[2 * e], sl2, -sd.x)), p1 = fast_exp2(fmaf(v[2 * e + 1], sl2, -sd.z));
          if (!((bits >> (2 * e)) & 1u)) p0 = 0.f;
          if (!((bits >> (2 * e + 1)) & 1u)) p1 = 0.f;
          wp[e] = pack2(p0, p1);
          wd[e] = pack2(p0 * fmaf(g[2 * e], p.scale, -sd.y), p1 * fmaf(g[2 * e + 1], p.scale, -sd.w));
        }
      }
      if (t >= 2) mbar_wait_poll(&bar_o[sb], ((t - 2) >> 1) & 1);  // the MMAs that read these two buffers are done
      const uint32_t pt = smem_u32(sPT + sb * 16384), dst = smem_u32(sDST + sb * 16384);
#pragma unroll
      for (int q8 = 0; q8 < 4; ++q8) {
        put8_row64(pt, r, half * 32 + q8 * 8, &wp[q8 * 4]);
        put8_row64(dst, r, half * 32 + q8 * 8, &wd[q8 * 4]);
      }
      fence_async_shared();
      tc_fence_before();
      mbar_arrive(&bar_done[sb]);
    }
  }

  if (warp < kIssuerWarp) {  // accumulator lane = key row
    mbar_wait_poll(&bar_o[(n - 1) & 1], ((n - 1) >> 1) & 1);  // the last pair of MMAs (they complete in order)
    tc_fence_after();
    const size_t row = static_cast<size_t>(b) * p.T + kj;
    store_acc_row<DH>(tmem_dk + lane_off, p.out0 + row * (p.H * DH) + col0, half);
    store_acc_row<DH>(tmem_dv + lane_off, p.out1 + row * (p.H * DH) + col0, half);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kIssuerWarp) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 512);
  }
}

template <int DH>
struct BwdQSmem {
  static constexpr uint32_t TILE = QB * DH * 2, SUB = SB * DH * 2;
  // stationary Q and dO, a three-stage ring of (K_j, V_j) sub-tiles, two buffers of the dS tile
  static constexpr uint32_t Q_OFF = 0, DO_OFF = TILE, K_OFF = 2 * TILE, V_OFF = K_OFF + 3 * SUB, DS_OFF = V_OFF + 3 * SUB,
                            BAR_OFF = DS_OFF + 2 * 16384;
  static constexpr uint32_t FLAG_OFF = BAR_OFF + 128;  // [2 * kMaxKeyBlocks] bytes: 64-key sub-block holds no padding
  static constexpr uint32_t BYTES = 1024 + FLAG_OFF + 2 * kMaxKeyBlocks;
  static_assert(BYTES <= 232448, "exceeds the 227 KB of shared memory a CTA can have");
};

// Query-block stationary: dQ of one 128-query block of one (batch, head), over 64-key sub-tiles.
template <int DH>
__global__ void __launch_bounds__(kAttnThreads, 1)
    attn_bwd_q_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                      const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                      const AttnParams p) {
  using L = BwdQSmem<DH>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem + L::Q_OFF;
  uint8_t* sDO = smem + L::DO_OFF;
  uint8_t* sK = smem + L::K_OFF;  // three stages
  uint8_t* sV = smem + L::V_OFF;  // three stages
  uint8_t* sDS = smem + L::DS_OFF;  // two buffers
  uint64_t* bar_q = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
  uint64_t* bar_kv = bar_q + 1;  // [3]
  uint64_t* bar_s = bar_q + 4;   // [2]
  uint64_t* bar_o = bar_q + 6;   // [2]
  uint64_t* bar_done = bar_q + 8;  // [2]: all 256 row threads have written dS of buffer b
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_q + 10);
  uint8_t* sub_ok = smem + L::FLAG_OFF;

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;  // warp-uniform for the compiler
  const int qb = static_cast<int>(gridDim.x) - 1 - static_cast<int>(blockIdx.x);  // longest key ranges first
  const int bh = blockIdx.y, b = bh / p.H, h = bh - b * p.H;
  const int col0 = h * DH;
  const bool issue_warp = (warp == kIssuerWarp);  // runs the issuing code in lockstep (the e_* instructions elect a lane)
  const bool issuer = issue_warp && lane == 0;    // one-time set-up only
  const uint8_t* mrow = p.key_mask != nullptr ? p.key_mask + static_cast<size_t>(b) * p.T : nullptr;
  if (issuer) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmDO);
    for (int i = 0; i < 8; ++i) mbar_init(bar_q + i, 1);
    mbar_init(&bar_done[0], kRowThreads);
    mbar_init(&bar_done[1], kRowThreads);
    fence_mbar_init();
  }
  if (warp == kIssuerWarp) {
    tmem_alloc<1>(tmem_slot, 512);
    tmem_relinquish<1>();
  } else if (mrow != nullptr && warp < kIssuerWarp) {
    key_sub_flags(mrow, p.T, sub_ok, threadIdx.x, lane);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0), tmem_dq = tmem_base + 256;

  // key sub-blocks this query block sees
  int first_key = 0;
  if (p.window > 0) first_key = max(0, qb * QB - p.window + 1);
  const int js_lo = first_key / SB, n = 2 * qb + 1 - js_lo + 1;
  auto load_kv = [&](int t) {  // (K_j, V_j) of sub-tile t -> ring stage t % 3
    const int st = t % 3, row = b * p.T + (js_lo + t) * SB;
    e_arrive_expect_tx(&bar_kv[st], 2 * L::SUB);
    load_sub<DH>(sK + st * L::SUB, &tmK, &bar_kv[st], row, col0);
    load_sub<DH>(sV + st * L::SUB, &tmV, &bar_kv[st], row, col0);
  };
  auto issue_s = [&](int t) {  // S(t) = Q.K_j^T and dP(t) = dO.V_j^T into TMEM buffer t & 1
    const int st = t % 3;
    mbar_wait_poll(&bar_kv[st], (t / 3) & 1);
    tc_fence_after();
    mma_full_sub<DH>(tmem_base + (t & 1) * 128, sQ, sK + st * L::SUB);
    mma_full_sub<DH>(tmem_base + (t & 1) * 128 + SB, sDO, sV + st * L::SUB);
    e_umma_commit(&bar_s[t & 1]);
  };
  if (issue_warp) {  // (the lambdas are called by the whole, converged issuing warp: the e_* instructions elect their own lane)
    e_arrive_expect_tx(bar_q, 2 * L::TILE);
    load_tile<DH>(sQ, &tmQ, bar_q, b * p.T + qb * QB, col0);
    load_tile<DH>(sDO, &tmDO, bar_q, b * p.T + qb * QB, col0);
    load_kv(0);
    if (n > 1) load_kv(1);
    mbar_wait_poll(bar_q, 0);
    issue_s(0);
  }

  const int r = threadIdx.x & 127, half = (threadIdx.x >> 7) & 1, qi = qb * QB + r;
  const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
  const float sl2 = p.scale * kLog2e;
  float lse2 = 0.f, dl = 0.f;
  if (warp < kIssuerWarp) {
    const float2 sd = p.stats[static_cast<size_t>(bh) * p.T + qi];
    lse2 = sd.x;  // +inf for a row that saw no key: every p is 0
    dl = sd.y;
  }

  if (issue_warp) {  // operand ring + score MMAs, one sub-tile ahead of the row threads
    for (int t = 0; t + 1 < n; ++t) {
      if (t >= 1) {
        mbar_wait_poll(&bar_o[(t - 1) & 1], ((t - 1) >> 1) & 1);     // dQ MMA of t-1 done: ring stage (t + 2) % 3 is free
        mbar_wait_poll(&bar_done[(t + 1) & 1], ((t - 1) >> 1) & 1);  // row threads have left TMEM buffer (t + 1) & 1
      }
      if (t + 2 < n) load_kv(t + 2);
      issue_s(t + 1);
    }
  } else if (warp == kIssuerWarp + 1) {  // dQ MMAs
    for (int t = 0; t < n; ++t) {
      const int sb = t & 1, st = t % 3;
      mbar_wait_poll(&bar_kv[st], (t / 3) & 1);     // (complete) this warp's own acquire of the stage's TMA writes
      mbar_wait_poll(&bar_done[sb], (t >> 1) & 1);  // dS(t) is in shared memory
      tc_fence_after();
      mma_k64_mn<DH>(tmem_dq, sDS + sb * 16384, sK + st * L::SUB, t != 0);  // dQ += dS . K_j
      e_umma_commit(&bar_o[sb]);
    }
  } else {
    for (int t = 0; t < n; ++t) {
      const int sb = t & 1, js = js_lo + t;
      mbar_wait_poll(&bar_s[sb], (t >> 1) & 1);
      tc_fence_after();
      float v[32], g[32];
      tmem_ld32(tmem_base + sb * 128 + lane_off + half * 32, v);
      tmem_ld32(tmem_base + sb * 128 + SB + lane_off + half * 32, g);
      tmem_ld_wait();
      const int key0 = js * SB + half * 32;
      const uint8_t* mr = (mrow != nullptr && !sub_ok[js]) ? mrow : nullptr;
      const bool all_vis = mr == nullptr && js * SB + SB - 1 <= qb * QB &&
                           (p.window <= 0 || js * SB > qb * QB + QB - 1 - p.window);
      if (!all_vis) mask_raw(v, qi, key0, p.window, mr);  // -inf where the key must not be attended -> p = 0
      uint32_t wd[16];
#pragma unroll
      for (int e = 0; e < 16; ++e) {
        const float p0 = fast_exp2(fmaf(v[2 * e], sl2, -lse2)), p1 = fast_exp2(fmaf(v[2 * e + 1], sl2, -lse2));
        wd[e] = pack2(p0 * fmaf(g[2 * e], p.scale, -dl), p1 * fmaf(g[2 * e + 1], p.scale, -dl));  // dl = scale * delta
      }
      if (t >= 2) mbar_wait_poll(&bar_o[sb], ((t - 2) >> 1) & 1);  // the MMA that read this dS buffer is done
      const uint32_t ds = smem_u32(sDS + sb * 16384);
#pragma unroll
      for (int q8 = 0; q8 < 4; ++q8) put8_row64(ds, r, half * 32 + q8 * 8, &wd[q8 * 4]);
      fence_async_shared();
      tc_fence_before();
      mbar_arrive(&bar_done[sb]);
    }
  }

  if (warp < kIssuerWarp) {
    mbar_wait_poll(&bar_o[(n - 1) & 1], ((n - 1) >> 1) & 1);
    tc_fence_after();
    store_acc_row<DH>(tmem_dq + lane_off, p.out0 + (static_cast<size_t>(b) * p.T + qi) * (p.H * DH) + col0, half);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kIssuerWarp) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 512);
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
                 float2* stats, void* dq, void* dk, void* dv, cudaStream_t st) {
  CUtensorMap tq, tk, tv, tdo;
  if (int rc = attn_tmap(&tq, q, B, p.T, p.H, DH)) return rc;
  if (int rc = attn_tmap(&tk, k, B, p.T, p.H, DH)) return rc;
  if (int rc = attn_tmap(&tv, v, B, p.T, p.H, DH)) return rc;
  if (int rc = attn_tmap(&tdo, d_o, B, p.T, p.H, DH)) return rc;
  if (int rc = set_smem(attn_bwd_kv_kernel<DH>, BwdKvSmem<DH>::BYTES)) return rc;
  if (int rc = set_smem(attn_bwd_q_kernel<DH>, BwdQSmem<DH>::BYTES)) return rc;
  ftk::TimelineScope tl_(FTK_TL_ATTN_BWD, st);
  const int64_t n = static_cast<int64_t>(B) * p.T * p.H * (DH / 8);  // threads: DH / 8 lanes per (row, head)
  attn_delta_kernel<DH><<<static_cast<unsigned>((n + 255) / 256), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(o),
                                                        static_cast<const __nv_bfloat16*>(d_o), p.lse, stats, B * p.T, p.T,
                                                        p.H, p.scale);
  FTK_CUDA_CHECK(cudaGetLastError());
  p.stats = stats;
  const dim3 grid(p.T / QB, B * p.H);
  p.out0 = static_cast<__nv_bfloat16*>(dk);
  p.out1 = static_cast<__nv_bfloat16*>(dv);
  attn_bwd_kv_kernel<DH><<<grid, kAttnThreads, BwdKvSmem<DH>::BYTES, st>>>(tq, tk, tv, tdo, p);
  FTK_CUDA_CHECK(cudaGetLastError());
  p.out0 = static_cast<__nv_bfloat16*>(dq);
  p.out1 = nullptr;
  attn_bwd_q_kernel<DH><<<grid, kAttnThreads, BwdQSmem<DH>::BYTES, st>>>(tq, tk, tv, tdo, p);
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
  FTK_REQUIRE(reinterpret_cast<uintptr_t>(O) % 32 == 0, "ftk_attention_fwd: O must be 32-byte aligned");
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
  return static_cast<size_t>(B) * T * H * sizeof(float2);  // (lse * log2(e), rowsum(dO * O)) per query and head
}

extern "C" int ftk_attention_bwd(const void* Q, const void* K, const void* V, const void* O, const void* dO,
                                 const float* lse, const uint8_t* key_mask, int32_t B, int32_t T, int32_t H,
                                 int32_t head_dim, int32_t window, float scale, void* dQ, void* dK, void* dV,
                                 void* workspace, size_t workspace_bytes, void* stream) {
  ftk::DeviceInfo di;
  if (int rc = check_attn(Q, K, V, B, T, H, head_dim, scale, &di)) return rc;
  FTK_REQUIRE(O && dO && lse && dQ && dK && dV && workspace, "ftk_attention_bwd: null pointer argument");
  FTK_REQUIRE((reinterpret_cast<uintptr_t>(dQ) | reinterpret_cast<uintptr_t>(dK) | reinterpret_cast<uintptr_t>(dV)) % 32 == 0,
              "ftk_attention_bwd: dQ, dK and dV must be 32-byte aligned");
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
  FTK_REQUIRE(reinterpret_cast<uintptr_t>(workspace) % 16 == 0, "ftk_attention_bwd: workspace must be 16-byte aligned");
  float2* stats = static_cast<float2*>(workspace);
  return head_dim == 128 ? run_attn_bwd<128>(Q, K, V, O, dO, p, B, stats, dQ, dK, dV, st)
                         : run_attn_bwd<64>(Q, K, V, O, dO, p, B, stats, dQ, dK, dV, st);
}
