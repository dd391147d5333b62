// G2/G3: fused lm_head projection + cross-entropy for sm_100a.
//
// One warp-specialised tcgen05 kernel template (TMA -> shared memory -> tcgen05.mma -> TMEM -> epilogue warps)
// is instantiated four times:
//   fwd   S = H.W^T, online-softmax epilogue: per-row running max / sum-exp / target logit; no logits in HBM
//   dz    S recomputed, epilogue forms dZ = (exp(S - lse) - onehot) * scale in bf16 into the slab workspace
//   dH    dH = dZ . W        (A K-major, B MN-major)
//   dW    dW (+)= dZ^T . H   (A MN-major, B MN-major)
// Reference arithmetic being replaced: HF lm_head Linear + loss_utils.ForCausalLMLoss and their autograd,
// reached from FT.py:479 (see include/ftk.h).
//
// Warp roles per CTA (256 threads): warp 0 = TMA producer (one lane), warp 1 = MMA issuer (one lane, leader CTA
// of a pair only), warp 2 = TMEM allocator, warps 4-7 = epilogue (one accumulator row per thread).
// Three mbarrier pipelines: smem full/empty (TMA <-> MMA), TMEM full/empty (MMA <-> epilogue).
#include <cuda.h>
#include <math.h>

#include <mutex>
#include <stdlib.h>
#include <string.h>

#include "ftk_host.h"
#include "ftk_ptx.cuh"

namespace {
using namespace ftk;

constexpr int BM = 128;  // accumulator rows per CTA (= TMEM lanes)
constexpr int BK = 64;   // K elements per smem stage: one 128-byte swizzle atom of bf16
constexpr int UK = 16;   // K per tcgen05.mma (bf16)
constexpr int kThreads = 256;
constexpr int kAccStages = 2;
constexpr int kTmemCols = 512;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr int kMaxChunks = 64;  // upper bound on vocabulary chunks (sizes the forward workspace)

// Work decomposition. A unit is (m block, chunk of consecutive N tiles); units are dealt round-robin to CTAs
// (or CTA pairs), so the units running at any moment are ~one wave of consecutive unit ids. The ids are ordered so
// that such a wave covers a group of `m_group` m blocks times a few consecutive chunks: the group's A blocks
// (m_group x rows x K) stay L2-resident while every chunk sweeps over them, and each B chunk is shared by the
// m_group clusters that sweep it together. m_group = 1 gives "all N tiles of one m block first".
struct Sched {
  int n_units;
  int n_mblk;           // m blocks (of BM * kCG rows)
  int n_chunks;         // chunks per m block
  int m_group;          // m blocks per group
  int tiles_per_chunk;  // N tiles per unit
  int n_ntile;          // N tiles in total
  int n_kblk;           // K blocks of BK
  int k_split;          // >= 1: every (m block, chunk) is cut into k_split units over disjoint K ranges (split-K);
                        // n_chunks then counts chunk x split pairs and the epilogue is told which split it has
  uint64_t hint_a;      // L2 eviction policy of the A / B operand loads
  uint64_t hint_b;
};

__device__ __forceinline__ void decode_unit(const Sched& sc, int u, int& m_blk, int& chunk) {
  const int per_group = sc.m_group * sc.n_chunks;
  const int grp = u / per_group;
  const int r = u - grp * per_group;
  const int m_first = grp * sc.m_group;
  const int gm = min(sc.m_group, sc.n_mblk - m_first);  // the last group may be smaller
  chunk = r / gm;
  m_blk = m_first + (r - chunk * gm);
}

// K range [kb0, kb1) of a unit whose decoded chunk index carries the split (chunk = real chunk * k_split + split).
__device__ __forceinline__ void split_of(const Sched& sc, int& chunk, int& split, int& kb0, int& kb1) {
  split = 0;
  kb0 = 0;
  kb1 = sc.n_kblk;
  if (sc.k_split > 1) {
    split = chunk % sc.k_split;
    chunk = chunk / sc.k_split;
    const int per = (sc.n_kblk + sc.k_split - 1) / sc.k_split;
    kb0 = split * per;
    kb1 = min(sc.n_kblk, kb0 + per);
  }
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

// ------------------------------------------------------------------------------------------------ epilogues
// Each epilogue thread owns one accumulator row. tile() reads the BN fp32 columns of that row from TMEM in
// groups of 32 and must finish all TMEM loads before returning (the caller then frees the accumulator stage).

// Per-thread context the kernel hands to an epilogue: this warp's TMEM lane quarter, lane id and a pointer to the
// epilogue's own shared-memory staging area (Epi::kSmemBytes bytes, 1024-byte aligned).
struct EpiCtx {
  uint8_t* smem;
  int q, lane;
  int half;  // epilogues with kColSplit: which half of every tile's columns this warp handles (else 0)
};

// Threads per CTA: warp 0 TMA, 1 MMA, 2 TMEM allocation, 3 idle, 4-7 epilogue (one accumulator row per thread), and
// for epilogues that declare kColSplit four more epilogue warps (8-11) on the other half of every tile's columns.
template <class Epi, class = void>
struct EpiColSplit { static constexpr bool value = false; };
template <class Epi>
struct EpiColSplit<Epi, decltype(void(Epi::kColSplit))> { static constexpr bool value = Epi::kColSplit; };
template <class Epi>
constexpr int threads_of() { return EpiColSplit<Epi>::value ? 384 : 256; }

// Forward: online softmax statistics over this unit's vocabulary chunk.
struct FwdEpi {
  static constexpr uint32_t kSmemBytes = 0;
  struct Params {
    const int64_t* targets;     // [M]
    const __nv_bfloat16* bias;  // [V] or null
    float* pmax;                // [n_chunks][Mpad] running max (natural units)
    float* psum;                // [n_chunks][Mpad] sum exp(z - max)
    float* ztgt;                // [Mpad] logit at the target column
    int M, V, Mpad;
  };
  struct State {
    float m, s, zt;
    int tgt, row, chunk;
    bool has;
  };
  template <int BN>
  static __device__ __forceinline__ void begin_unit(State& st, const Params& p, int row, int chunk) {
    st.row = row;
    st.chunk = chunk;
    st.m = -INFINITY;
    st.s = 0.f;
    st.zt = 0.f;
    st.has = false;
    st.tgt = (row < p.M) ? static_cast<int>(p.targets[row]) : -1;
  }
  template <int BN>
  static __device__ __forceinline__ void tile(State& st, const Params& p, uint32_t taddr, int n0, const EpiCtx&) {
#pragma unroll 1
    for (int c = 0; c < BN / 32; ++c) {
      const int col0 = n0 + c * 32;
      if (col0 >= p.V) break;  // warp-uniform
      float v[32];
      tmem_ld32(taddr + c * 32, v);
      tmem_ld_wait();
      if (p.bias != nullptr) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (col0 + j < p.V) v[j] += __bfloat162float(__ldg(p.bias + col0 + j));
      }
      if (col0 + 32 > p.V) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (col0 + j >= p.V) v[j] = -INFINITY;  // TMA zero-fills phantom rows of W: mask, do not count
      }
      if (st.tgt >= col0 && st.tgt < col0 + 32) {
        const int jj = st.tgt - col0;
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (j == jj) st.zt = v[j];
        st.has = true;
      }
      float cmax = v[0];
#pragma unroll
      for (int j = 1; j < 32; ++j) cmax = fmaxf(cmax, v[j]);
      const float m_new = fmaxf(st.m, cmax);  // finite: column col0 < V is real
      const float neg = -m_new * kLog2e;
      float acc = 0.f;
#pragma unroll
      for (int j = 0; j < 32; ++j) acc += fast_exp2(fmaf(v[j], kLog2e, neg));
      st.s = st.s * fast_exp2(fmaf(st.m, kLog2e, neg)) + acc;  // exp2(-inf) = 0 on the first group
      st.m = m_new;
    }
  }
  template <int BN>
  static __device__ __forceinline__ void end_unit(State& st, const Params& p) {
    if (st.row < p.M) {
      p.pmax[static_cast<size_t>(st.chunk) * p.Mpad + st.row] = st.m;
      p.psum[static_cast<size_t>(st.chunk) * p.Mpad + st.row] = st.s;
      if (st.has) p.ztgt[st.row] = st.zt;
    }
  }
  static __device__ __forceinline__ void set_split(State&, const Params&, int) {}
  static __device__ __forceinline__ void finish(const EpiCtx&) {}
};

// Backward stage 1: dZ tile from the recomputed logits.
struct DzEpi {
  // Staging for the TMA stores: per epilogue warp two [32 rows x 128 B] buffers in the 128-byte swizzle layout.
  static constexpr uint32_t kSmemBytes = 4 * 2 * 4096;
  struct Params {
    CUtensorMap tm_out;         // dZ slab as bf16 [rows][ld], box 64 columns x 32 rows, 128-byte swizzle
    const int64_t* targets;     // [rows] (slab-relative)
    const __nv_bfloat16* bias;  // [V] or null
    const float* lse;           // [rows]
    const float* scale_dev;     // optional device scalar multiplied into scale
    const int32_t* nvalid_dev;  // optional device count divided out of scale
    uint64_t store_hint;        // L2 eviction policy of the dZ stores
    float scale;
    int rows, V, ld;
  };
  struct State {
    float l2, sc;
    int tgt, row;
    uint32_t nstore;
  };
  template <int BN>
  static __device__ __forceinline__ void begin_unit(State& st, const Params& p, int row, int) {
    st.row = row;
    st.tgt = -1;
    st.l2 = 0.f;
    st.sc = 0.f;
    if (row < p.rows) {
      st.tgt = static_cast<int>(p.targets[row]);
      st.l2 = -p.lse[row] * kLog2e;
      float sc = p.scale;
      if (p.scale_dev != nullptr) sc *= __ldg(p.scale_dev);
      if (p.nvalid_dev != nullptr) sc /= static_cast<float>(__ldg(p.nvalid_dev));
      st.sc = (st.tgt >= 0) ? sc : 0.f;  // ignored rows contribute no gradient
    }
  }
  template <int BN>
  static __device__ __forceinline__ void tile(State& st, const Params& p, uint32_t taddr, int n0, const EpiCtx& cx) {
    uint8_t* warp_buf = cx.smem + cx.q * 8192;
    const int row0 = st.row - cx.lane;  // first row of this warp's 32-row band
#pragma unroll 1
    for (int g = 0; g < BN / 64; ++g) {
      const int col0 = n0 + g * 64;
      if (col0 >= p.ld) break;  // warp-uniform
      uint8_t* buf = warp_buf + (st.nstore & 1) * 4096;
      // the store issued two groups ago out of this buffer must have finished reading it
      if (cx.lane == 0) tma_store_wait_read<1>();
      __syncwarp();
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int c0 = col0 + h * 32;
        float v[32];
        tmem_ld32(taddr + g * 64 + h * 32, v);
        tmem_ld_wait();
        if (p.bias != nullptr) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (c0 + j < p.V) v[j] += __bfloat162float(__ldg(p.bias + c0 + j));
        }
        uint32_t w[16];
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          float g0 = fast_exp2(fmaf(v[j], kLog2e, st.l2));
          float g1 = fast_exp2(fmaf(v[j + 1], kLog2e, st.l2));
          g0 = (c0 + j == st.tgt) ? g0 - 1.f : g0;
          g1 = (c0 + j + 1 == st.tgt) ? g1 - 1.f : g1;
          g0 = (c0 + j < p.V) ? g0 * st.sc : 0.f;  // pad columns [V, ld) are written as zeros
          g1 = (c0 + j + 1 < p.V) ? g1 * st.sc : 0.f;
          w[j >> 1] = pack_bf16x2(g0, g1);
        }
        // row `lane` of the box, 16-byte chunk c at position c ^ (lane & 7): the layout TMA expects for
        // SWIZZLE_128B, and conflict-free for the 8 lanes of each store phase
#pragma unroll
        for (int qq = 0; qq < 4; ++qq) {
          const int c = h * 4 + qq;
          uint4* dst = reinterpret_cast<uint4*>(buf + cx.lane * 128 + ((c ^ (cx.lane & 7)) << 4));
          *dst = make_uint4(w[4 * qq], w[4 * qq + 1], w[4 * qq + 2], w[4 * qq + 3]);
        }
      }
      fence_async_shared();  // make the generic-proxy smem writes visible to the TMA engine
      __syncwarp();
      if (cx.lane == 0) {
        // rows >= p.rows / columns >= ld are clipped by the tensor map
        tma_store_2d_hint(&p.tm_out, buf, col0, row0, p.store_hint);
        tma_store_commit();
      }
      ++st.nstore;
    }
  }
  template <int BN>
  static __device__ __forceinline__ void end_unit(State&, const Params&) {}
  static __device__ __forceinline__ void set_split(State&, const Params&, int) {}
  static __device__ __forceinline__ void finish(const EpiCtx& cx) {
    if (cx.lane == 0) tma_store_wait_all<0>();  // all dZ writes complete before the CTA (and its smem) goes away
  }
};

// Forward that also keeps the softmax numerators (opt-in; see ftk_lmhead_ce_fwd_keep): same statistics as FwdEpi,
// and every tile additionally leaves exp(z - c) in bf16 in a [rows][ld] slab, c = the row's running maximum after
// that tile (written to cref[tile][row]).  c is a true maximum, so the exponent is <= 0 and nothing can overflow.
// The backward then needs no logit recompute: dZ = numerators * exp(c - lse) * scale - onehot * scale, elementwise.
struct FwdKeepEpi {
  static constexpr uint32_t kSmemBytes = 4 * 2 * 4096;  // staging for the TMA stores, as in DzEpi
  struct Params {
    CUtensorMap tm_out;         // numerators slab, bf16 [rows][ld], box 64 columns x 32 rows, 128-byte swizzle
    const int64_t* targets;     // [M]
    const __nv_bfloat16* bias;  // [V] or null
    float* pmax;                // [n_chunks][Mpad]
    float* psum;                // [n_chunks][Mpad]
    float* ztgt;                // [Mpad]
    float* cref;                // [n_ntile][Mpad] reference maximum of each (tile, row)
    uint64_t store_hint;
    int M, V, Mpad, ld;
  };
  struct State {
    float m, s, zt;
    int tgt, row, chunk;
    bool has;
    uint32_t nstore;  // survives across units: parity of the staging buffers
  };
  template <int BN>
  static __device__ __forceinline__ void begin_unit(State& st, const Params& p, int row, int chunk) {
    st.row = row;
    st.chunk = chunk;
    st.m = -INFINITY;
    st.s = 0.f;
    st.zt = 0.f;
    st.has = false;
    st.tgt = (row < p.M) ? static_cast<int>(p.targets[row]) : -1;
  }
  template <int BN>
  static __device__ __forceinline__ void tile(State& st, const Params& p, uint32_t taddr, int n0, const EpiCtx& cx) {
    // ---- pass 1 over the accumulator: row maximum of this tile (and the target logit)
    float tmax = -INFINITY;
#pragma unroll 1
    for (int c = 0; c < BN / 64; ++c) {
      const int col0 = n0 + c * 64;
      if (col0 >= p.V) break;  // warp-uniform
      float v[64];
      tmem_ld64(taddr + c * 64, v);
      tmem_ld_wait();
      if (p.bias != nullptr) {
#pragma unroll
        for (int j = 0; j < 64; ++j)
          if (col0 + j < p.V) v[j] += __bfloat162float(__ldg(p.bias + col0 + j));
      }
      if (col0 + 64 > p.V) {
#pragma unroll
        for (int j = 0; j < 64; ++j)
          if (col0 + j >= p.V) v[j] = -INFINITY;
      }
      if (st.tgt >= col0 && st.tgt < col0 + 64) {
        const int jj = st.tgt - col0;
#pragma unroll
        for (int j = 0; j < 64; ++j)
          if (j == jj) st.zt = v[j];
        st.has = true;
      }
#pragma unroll
      for (int j = 0; j < 64; ++j) tmax = fmaxf(tmax, v[j]);
    }
    const float m_new = fmaxf(st.m, tmax);  // finite: column n0 < V is real
    const float neg = -m_new * kLog2e;
    st.s *= fast_exp2(fmaf(st.m, kLog2e, neg));  // exp2(-inf) = 0 on the first tile of a unit
    st.m = m_new;
    if (st.row < p.M) p.cref[static_cast<size_t>(n0 / BN) * p.Mpad + st.row] = m_new;
    // ---- pass 2: numerators relative to m_new, summed in fp32 and stored in bf16
    uint8_t* warp_buf = cx.smem + cx.q * 8192;
    const int row0 = st.row - cx.lane;
    float acc = 0.f;
#pragma unroll 1
    for (int g = 0; g < BN / 64; ++g) {
      const int col0 = n0 + g * 64;
      if (col0 >= p.ld) break;  // warp-uniform
      uint8_t* buf = warp_buf + (st.nstore & 1) * 4096;
      if (cx.lane == 0) tma_store_wait_read<1>();
      __syncwarp();
      float v[64];
      tmem_ld64(taddr + g * 64, v);
      tmem_ld_wait();
      if (p.bias != nullptr) {
#pragma unroll
        for (int j = 0; j < 64; ++j)
          if (col0 + j < p.V) v[j] += __bfloat162float(__ldg(p.bias + col0 + j));
      }
      const bool tail = col0 + 64 > p.V;  // warp-uniform: only the last real group needs the column predicate
#pragma unroll
      for (int qq = 0; qq < 8; ++qq) {
        uint32_t w[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int j = qq * 8 + 2 * e;
          float e0 = fast_exp2(fmaf(v[j], kLog2e, neg));
          float e1 = fast_exp2(fmaf(v[j + 1], kLog2e, neg));
          if (tail) {
            e0 = (col0 + j < p.V) ? e0 : 0.f;
            e1 = (col0 + j + 1 < p.V) ? e1 : 0.f;
          }
          acc += e0 + e1;
          w[e] = pack_bf16x2(e0, e1);
        }
        uint4* dst = reinterpret_cast<uint4*>(buf + cx.lane * 128 + ((qq ^ (cx.lane & 7)) << 4));
        *dst = make_uint4(w[0], w[1], w[2], w[3]);
      }
      fence_async_shared();
      __syncwarp();
      if (cx.lane == 0) {
        tma_store_2d_hint(&p.tm_out, buf, col0, row0, p.store_hint);
        tma_store_commit();
      }
      ++st.nstore;
    }
    st.s += acc;
  }
  template <int BN>
  static __device__ __forceinline__ void end_unit(State& st, const Params& p) {
    if (st.row < p.M) {
      p.pmax[static_cast<size_t>(st.chunk) * p.Mpad + st.row] = st.m;
      p.psum[static_cast<size_t>(st.chunk) * p.Mpad + st.row] = st.s;
      if (st.has) p.ztgt[st.row] = st.zt;
    }
  }
  static __device__ __forceinline__ void set_split(State&, const Params&, int) {}
  static __device__ __forceinline__ void finish(const EpiCtx& cx) {
    if (cx.lane == 0) tma_store_wait_all<0>();
  }
};

// 32-byte global accesses (STG.E.ENL2.256 / LDG.E.ENL2.256).  An epilogue thread owns one accumulator row, so a warp's
// store instruction touches 32 different rows: with 16-byte accesses every one of them is half a 32-byte sector and the
// L2 has to merge two partial writes per sector (the c_fc GEMM that also stores its pre-activation ran at 868 TF/s
// with 16-byte stores and at 1326 TF/s, the speed of the plain GEMM, with 32-byte ones).
// Plain GEMM epilogue: out[row, col] (+)= acc, bf16 or fp32 row-major.  Under split-K each split writes its partial
// sums to its own copy of the output, split_stride elements apart (summed afterwards in a fixed order).
template <bool kF32>
struct StoreEpi {
  static constexpr uint32_t kSmemBytes = 0;
  struct Params {
    void* out;
    int rows, cols, ld, accumulate;
    size_t split_stride;
  };
  struct State {
    int row;
    size_t base;  // element offset of this unit's split
  };
  template <int BN>
  static __device__ __forceinline__ void begin_unit(State& st, const Params&, int row, int) {
    st.row = row;
    st.base = 0;
  }
  static __device__ __forceinline__ void set_split(State& st, const Params& p, int split) {
    st.base = static_cast<size_t>(split) * p.split_stride;
  }
  template <int BN>
  static __device__ __forceinline__ void tile(State& st, const Params& p, uint32_t taddr, int n0, const EpiCtx&) {
#pragma unroll 1
    for (int c = 0; c < BN / 32; ++c) {
      const int col0 = n0 + c * 32;
      if (col0 >= p.cols) break;  // warp-uniform
      float v[32];
      tmem_ld32(taddr + c * 32, v);
      tmem_ld_wait();
      // 32-byte accesses when the chunk is whole and the output is 32-byte aligned (always, for the callers in this
      // library: d is a multiple of 64 and the buffers come from an allocator); otherwise the 16-byte path below
      const bool wide = col0 + 32 <= p.cols && (p.ld & 15) == 0 && (reinterpret_cast<uintptr_t>(p.out) & 31) == 0 &&
                        (st.base & 15) == 0;
      if (st.row >= p.rows) {
        // rows past the end: nothing to store (no `continue`: the next tcgen05.ld needs a converged warp)
      } else if (wide) {
        if constexpr (kF32) {
          float* dst = static_cast<float*>(p.out) + st.base + static_cast<size_t>(st.row) * p.ld + col0;
          if (p.accumulate) {
#pragma unroll
            for (int g8 = 0; g8 < 4; ++g8) {
              uint32_t old[8];
              ld_global_32B(dst + g8 * 8, old);
#pragma unroll
              for (int e = 0; e < 8; ++e) v[g8 * 8 + e] += __uint_as_float(old[e]);
            }
          }
#pragma unroll
          for (int g8 = 0; g8 < 4; ++g8) st_global_32B(dst + g8 * 8, reinterpret_cast<const uint32_t*>(v + g8 * 8));
        } else {
          __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(p.out) + st.base + static_cast<size_t>(st.row) * p.ld + col0;
          if (p.accumulate) {
#pragma unroll
            for (int g16 = 0; g16 < 2; ++g16) {
              uint32_t old[8];
              ld_global_32B(dst + g16 * 16, old);
#pragma unroll
              for (int e = 0; e < 8; ++e) {
                const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&old[e]);
                v[g16 * 16 + 2 * e] += __low2float(h);
                v[g16 * 16 + 2 * e + 1] += __high2float(h);
              }
            }
          }
#pragma unroll
          for (int g16 = 0; g16 < 2; ++g16) {
            uint32_t w[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) w[e] = pack_bf16x2(v[g16 * 16 + 2 * e], v[g16 * 16 + 2 * e + 1]);
            st_global_32B(dst + g16 * 16, w);
          }
        }
      } else if constexpr (kF32) {
        float4* dst = reinterpret_cast<float4*>(static_cast<float*>(p.out) + st.base + static_cast<size_t>(st.row) * p.ld + col0);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          if (col0 + q * 4 >= p.cols) break;
          float4 o = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
          if (p.accumulate) {
            const float4 old = dst[q];
            o.x += old.x;
            o.y += old.y;
            o.z += old.z;
            o.w += old.w;
          }
          dst[q] = o;
        }
      } else {
        uint4* dst = reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(p.out) + st.base +
                                              static_cast<size_t>(st.row) * p.ld + col0);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          if (col0 + q * 8 >= p.cols) break;
          float* x = v + 8 * q;
          if (p.accumulate) {
            const uint4 old = dst[q];
            const uint32_t ow[4] = {old.x, old.y, old.z, old.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&ow[e]);
              x[2 * e] += __low2float(h);
              x[2 * e + 1] += __high2float(h);
            }
          }
          dst[q] = make_uint4(pack_bf16x2(x[0], x[1]), pack_bf16x2(x[2], x[3]), pack_bf16x2(x[4], x[5]),
                              pack_bf16x2(x[6], x[7]));
        }
      }
    }
  }
  template <int BN>
  static __device__ __forceinline__ void end_unit(State&, const Params&) {}
  static __device__ __forceinline__ void finish(const EpiCtx&) {}
};

// Linear-layer epilogue (the trunk's projections): bf16 out[row, col] = f(acc + bias[col]).
//   act == 1: f = gelu_new (the tanh form HF's GPT-Neo / GPT-J MLPs use); `pre` (optional) receives
//             gelu_new'(acc + bias), the factor the backward needs (same tanh, a few more FMAs here, none there);
//   mul_pre != null: out = acc * mul_pre[row, col] -- the dX GEMM of the layer AFTER a GELU multiplies by that saved
//             derivative and hands back the gradient with respect to the pre-activation directly, so no elementwise
//             pass ever touches [M, 4d].
//   residual != null: out += residual[row, col] (last, in fp32 before the one rounding): the block's residual
//             connection in the forward, the sum of several layers' input gradients in the backward.  May alias out.
struct LinearEpi {
  static constexpr uint32_t kSmemBytes = 0;
  // Two threads per accumulator row, each on half of the tile's columns: with K = 2048 a 256 x 256 tile's MMAs take
  // ~12 us and the gelu_new + two-store epilogue of c_fc took longer than that on four warps (868 TF/s against 1271
  // for the plain GEMM).
  static constexpr bool kColSplit = true;
  struct Params {
    __nv_bfloat16* out;
    __nv_bfloat16* pre;
    const __nv_bfloat16* bias;
    const __nv_bfloat16* mul_pre;
    const __nv_bfloat16* residual;
    int rows, cols, ld, act;
  };
  struct State {
    int row;
  };
  static __device__ __forceinline__ float tanh_fast(float x) {
    float r;
    asm("tanh.approx.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
  }
  // y = gelu_new(x) = 0.5 x (1 + tanh(u)), u = sqrt(2/pi) (x + 0.044715 x^3), and dy/dx = 0.5 (1 + t) + 0.5 x (1 - t^2) u'
  static __device__ __forceinline__ void gelu_new_both(float x, float& y, float& dy) {
    const float x2 = x * x;
    const float t = tanh_fast(0.7978845608028654f * fmaf(0.044715f * x, x2, x));
    const float du = 0.7978845608028654f * fmaf(3.f * 0.044715f, x2, 1.f);
    const float h = 0.5f * (1.f + t);
    y = x * h;
    dy = fmaf(0.5f * x * du, fmaf(-t, t, 1.f), h);
  }
  template <int BN>
  static __device__ __forceinline__ void begin_unit(State& st, const Params&, int row, int) {
    st.row = row;
  }
  static __device__ __forceinline__ void set_split(State&, const Params&, int) {}
  // The side inputs of one 32-column chunk of this thread's row (mul_pre and residual, 64 bytes each).  They are
  // fetched one chunk ahead: a thread reads its own row, so these loads are latency-, not bandwidth-bound, and left in
  // the chunk's dependency chain they stretch the epilogue past the MMAs of the next tile.
  struct Side {
    uint32_t m[16], r[16];
  };
  static __device__ __forceinline__ void fetch(Side& sd, const State& st, const Params& p, int col0) {
    if (st.row >= p.rows || col0 >= p.cols) return;
    const size_t off = static_cast<size_t>(st.row) * p.ld + col0;  // 64-byte aligned: two 32-byte loads per input
    if (p.mul_pre != nullptr) {
      ld_global_32B(p.mul_pre + off, sd.m);
      ld_global_32B(p.mul_pre + off + 16, sd.m + 8);
    }
    if (p.residual != nullptr) {
      ld_global_32B(p.residual + off, sd.r);
      ld_global_32B(p.residual + off + 16, sd.r + 8);
    }
  }
  template <int BN>
  static __device__ __forceinline__ void tile(State& st, const Params& p, uint32_t taddr, int n0, const EpiCtx& cx) {
    constexpr int kChunks = BN / 32 / 2;  // this warp's half of the tile
    const int c_begin = cx.half * kChunks, c_end = c_begin + kChunks;
    Side nxt = {};
    fetch(nxt, st, p, n0 + c_begin * 32);
#pragma unroll 1
    for (int c = c_begin; c < c_end; ++c) {
      const int col0 = n0 + c * 32;
      if (col0 >= p.cols) break;  // warp-uniform
      float v[32];
      tmem_ld32(taddr + c * 32, v);
      const Side cur = nxt;
      if (c + 1 < c_end) fetch(nxt, st, p, col0 + 32);
      tmem_ld_wait();
      if (st.row < p.rows) {  // no `continue`: the next tcgen05.ld needs a converged warp
        // cols is a multiple of 64 (checked on the host), so a 32-column chunk is wholly inside or was skipped above
        const size_t off = static_cast<size_t>(st.row) * p.ld + col0;
        uint32_t wo[16], wp[16];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float* x = v + 8 * q;
          if (p.bias != nullptr) {
            const uint4 bw = __ldg(reinterpret_cast<const uint4*>(p.bias + col0) + q);
            const uint32_t b4[4] = {bw.x, bw.y, bw.z, bw.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&b4[e]);
              x[2 * e] += __low2float(h);
              x[2 * e + 1] += __high2float(h);
            }
          }
          if (p.mul_pre != nullptr) {  // the saved gelu_new'(pre) of the layer that produced X
            const uint32_t* p4 = cur.m + 4 * q;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&p4[e]);
              x[2 * e] *= __low2float(h);
              x[2 * e + 1] *= __high2float(h);
            }
          }
          if (p.act == 1) {  // gelu_new and its derivative share the tanh
            float dg[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) gelu_new_both(x[e], x[e], dg[e]);
#pragma unroll
            for (int e = 0; e < 4; ++e) wp[4 * q + e] = pack_bf16x2(dg[2 * e], dg[2 * e + 1]);
          }
          if (p.residual != nullptr) {
            const uint32_t* r4 = cur.r + 4 * q;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&r4[e]);
              x[2 * e] += __low2float(h);
              x[2 * e + 1] += __high2float(h);
            }
          }
#pragma unroll
          for (int e = 0; e < 4; ++e) wo[4 * q + e] = pack_bf16x2(x[2 * e], x[2 * e + 1]);
        }
        // 32-byte stores: every instruction writes whole 32-byte sectors of this thread's row
        if (p.act == 1 && p.pre != nullptr) {
          st_global_32B(p.pre + off, wp);
          st_global_32B(p.pre + off + 16, wp + 8);
        }
        st_global_32B(p.out + off, wo);
        st_global_32B(p.out + off + 16, wo + 8);
      }
    }
  }
  template <int BN>
  static __device__ __forceinline__ void end_unit(State&, const Params&) {}
  static __device__ __forceinline__ void finish(const EpiCtx&) {}
};

// ------------------------------------------------------------------------------------------------ the kernel
template <int kCG, int BN, int kStages, uint32_t kEpiBytes>
struct SmemLayout {
  static constexpr int LOAD_BN = BN / kCG;
  static constexpr uint32_t A_BYTES = BM * BK * 2;
  static constexpr uint32_t B_BYTES = LOAD_BN * BK * 2;
  static constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr uint32_t EPI_OFF = kStages * STAGE_BYTES;  // 1024-byte aligned (stages are multiples of 1024)
  static constexpr uint32_t BAR_OFF = EPI_OFF + kEpiBytes;
  static constexpr uint32_t BYTES = 1024 /*alignment slack*/ + BAR_OFF + (2 * kStages + 2 * kAccStages) * 8 + 16;
  static_assert(BYTES <= 232448, "exceeds the 227 KB of shared memory a CTA can have");
};

template <int kCG, int BN, int kStages, bool kAMN, bool kBMN, class Epi>
__global__ void __launch_bounds__(threads_of<Epi>(), 1)
    umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const Sched sc,
                const __grid_constant__ typename Epi::Params ep) {
  using L = SmemLayout<kCG, BN, kStages, Epi::kSmemBytes>;
  constexpr int LOAD_BN = L::LOAD_BN;
  static_assert(BN % 32 == 0 && BN <= 256 && kAccStages * BN <= kTmemCols, "accumulator does not fit TMEM");
  static_assert(LOAD_BN % 64 == 0, "B tile must be whole 64-wide swizzle atoms");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + kStages * L::A_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
  uint64_t* empty = full + kStages;
  uint64_t* tfull = empty + kStages;
  uint64_t* tempty = tfull + kAccStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + kAccStages);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = (kCG == 2) ? cluster_ctarank() : 0u;
  const int cl = blockIdx.x / kCG, n_cl = gridDim.x / kCG;

  if constexpr (kCG == 2) cluster_sync_all();  // both CTAs resident before the paired TMEM allocation

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full[i], kCG);  // leader: own arrive.expect_tx + the peer's arrive
      mbar_init(&empty[i], 1);   // tcgen05.commit (multicast to both CTAs of a pair)
    }
    for (int i = 0; i < kAccStages; ++i) {
      mbar_init(&tfull[i], 1);         // tcgen05.commit
      mbar_init(&tempty[i], kCG * (threads_of<Epi>() / 32 - 4));  // one arrive per epilogue warp of each CTA
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc<kCG>(tmem_slot, kTmemCols);
    tmem_relinquish<kCG>();
  }
  tc_fence_before();
  if constexpr (kCG == 2)
    cluster_sync_all();
  else
    __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ============================== TMA producer ==============================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int u = cl; u < sc.n_units; u += n_cl) {
        int m_blk, chunk, split, kb0, kb1;
        decode_unit(sc, u, m_blk, chunk);
        split_of(sc, chunk, split, kb0, kb1);
        const int m0 = (m_blk * kCG + static_cast<int>(rank)) * BM;
        const int nt_begin = chunk * sc.tiles_per_chunk;
        const int nt_end = min(nt_begin + sc.tiles_per_chunk, sc.n_ntile);
        for (int nt = nt_begin; nt < nt_end; ++nt) {
          const int n0 = nt * BN + static_cast<int>(rank) * LOAD_BN;
          for (int kb = kb0; kb < kb1; ++kb) {
            mbar_wait(&empty[stage], phase ^ 1);
            uint64_t* bar = &full[stage];
            if constexpr (kCG == 1) {
              mbar_arrive_expect_tx(bar, L::STAGE_BYTES);
            } else {
              if (rank == 0)
                mbar_arrive_expect_tx(bar, 2 * L::STAGE_BYTES);
              else
                mbar_arrive_cluster(bar, 0);
            }
            uint8_t* a_dst = sA + stage * L::A_BYTES;
            uint8_t* b_dst = sB + stage * L::B_BYTES;
            const int k0 = kb * BK;
            if constexpr (!kAMN) {
              if constexpr (kCG == 1)
                tma_load_2d_hint(a_dst, &tmA, bar, k0, m0, sc.hint_a);
              else
                tma_load_2d_pair_hint(a_dst, &tmA, bar, k0, m0, sc.hint_a);
            } else {
#pragma unroll
              for (int j = 0; j < BM / 64; ++j) {
                if constexpr (kCG == 1)
                  tma_load_2d_hint(a_dst + j * (BK * 128), &tmA, bar, m0 + 64 * j, k0, sc.hint_a);
                else
                  tma_load_2d_pair_hint(a_dst + j * (BK * 128), &tmA, bar, m0 + 64 * j, k0, sc.hint_a);
              }
            }
            if constexpr (!kBMN) {
              if constexpr (kCG == 1)
                tma_load_2d_hint(b_dst, &tmB, bar, k0, n0, sc.hint_b);
              else
                tma_load_2d_pair_hint(b_dst, &tmB, bar, k0, n0, sc.hint_b);
            } else {
#pragma unroll
              for (int j = 0; j < LOAD_BN / 64; ++j) {
                if constexpr (kCG == 1)
                  tma_load_2d_hint(b_dst + j * (BK * 128), &tmB, bar, n0 + 64 * j, k0, sc.hint_b);
                else
                  tma_load_2d_pair_hint(b_dst + j * (BK * 128), &tmB, bar, n0 + 64 * j, k0, sc.hint_b);
              }
            }
            if (++stage == kStages) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ============================== MMA issuer (leader CTA) ==============================
    if (lane == 0 && rank == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM * kCG, BN, kAMN, kBMN);
      // K-major operand: rows of 128 B, 8-row groups 1024 B apart (SBO); LBO unused.
      // MN-major operand: 64-element (128 B) runs along MN, 8 k-rows per 1024 B group (SBO), next 64 MN
      // elements one whole [BK x 128 B] box further (LBO).
      constexpr uint32_t kLboMN = BK * 128, kSbo = 1024;
      constexpr uint32_t a_kstep = kAMN ? (UK * 128) : (UK * 2);
      constexpr uint32_t b_kstep = kBMN ? (UK * 128) : (UK * 2);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t it = 0;
      for (int u = cl; u < sc.n_units; u += n_cl) {
        int m_blk, chunk, split, kb0, kb1;
        decode_unit(sc, u, m_blk, chunk);
        split_of(sc, chunk, split, kb0, kb1);
        const int nt_begin = chunk * sc.tiles_per_chunk;
        const int nt_end = min(nt_begin + sc.tiles_per_chunk, sc.n_ntile);
        for (int nt = nt_begin; nt < nt_end; ++nt, ++it) {
          const uint32_t as = it & 1, aphase = (it >> 1) & 1;
          mbar_wait(&tempty[as], aphase ^ 1);
          tc_fence_after();
          const uint32_t tmem_d = tmem_base + as * BN;
          for (int kb = kb0; kb < kb1; ++kb) {
            mbar_wait(&full[stage], phase);
            tc_fence_after();
            const uint32_t a_addr = smem_u32(sA + stage * L::A_BYTES);
            const uint32_t b_addr = smem_u32(sB + stage * L::B_BYTES);
#pragma unroll
            for (int k = 0; k < BK / UK; ++k) {
              const uint64_t adesc = make_sw128_desc(a_addr + k * a_kstep, kAMN ? kLboMN : 0u, kSbo);
              const uint64_t bdesc = make_sw128_desc(b_addr + k * b_kstep, kBMN ? kLboMN : 0u, kSbo);
              umma_bf16<kCG>(tmem_d, adesc, bdesc, idesc, (kb != kb0 || k != 0) ? 1u : 0u);
            }
            umma_commit<kCG>(&empty[stage]);  // frees the smem stage once these MMAs have read it
            if (kb == kb1 - 1) umma_commit<kCG>(&tfull[as]);
            if (++stage == kStages) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
      }
    }
  } else if (warp >= 4) {
    // ============================== epilogue ==============================
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    const int row_in_tile = q * 32 + lane;
    const EpiCtx cx{smem + L::EPI_OFF, q, lane, (warp - 4) >> 2};
    typename Epi::State st = {};
    uint32_t it = 0;
    for (int u = cl; u < sc.n_units; u += n_cl) {
      int m_blk, chunk, split, kb0, kb1;
      decode_unit(sc, u, m_blk, chunk);
      split_of(sc, chunk, split, kb0, kb1);
      const int row = (m_blk * kCG + static_cast<int>(rank)) * BM + row_in_tile;
      Epi::template begin_unit<BN>(st, ep, row, chunk);
      Epi::set_split(st, ep, split);
      const int nt_begin = chunk * sc.tiles_per_chunk;
      const int nt_end = min(nt_begin + sc.tiles_per_chunk, sc.n_ntile);
      for (int nt = nt_begin; nt < nt_end; ++nt, ++it) {
        const uint32_t as = it & 1, aphase = (it >> 1) & 1;
        mbar_wait(&tfull[as], aphase);
        tc_fence_after();
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN;
        Epi::template tile<BN>(st, ep, taddr, nt * BN, cx);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if constexpr (kCG == 1)
            mbar_arrive(&tempty[as]);
          else
            mbar_arrive_cluster(&tempty[as], 0);
        }
      }
      Epi::template end_unit<BN>(st, ep);
    }
    Epi::finish(cx);
  }

  // ---- teardown: nobody may leave while a peer can still touch its smem / barriers / TMEM
  __syncwarp();
  tc_fence_before();
  if constexpr (kCG == 2)
    cluster_sync_all();
  else
    __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<kCG>(tmem_base, kTmemCols);
  }
}

// ------------------------------------------------------------------------------------------------ small kernels
// Merge the per-chunk softmax statistics: lse per row, then a deterministic two-level reduction of the loss.
__global__ void __launch_bounds__(256) fwd_finalize_kernel(const float* __restrict__ pmax,
                                                           const float* __restrict__ psum,
                                                           const float* __restrict__ ztgt,
                                                           const int64_t* __restrict__ targets, int M, int Mpad,
                                                           int V, int n_chunks, float* __restrict__ lse,
                                                           float* __restrict__ blk_loss, int* __restrict__ blk_cnt,
                                                           unsigned int* __restrict__ ticket,
                                                           float* __restrict__ loss_sum, int* __restrict__ n_valid,
                                                           int add_to_loss) {
  __shared__ float s_loss[8];
  __shared__ int s_cnt[8];
  __shared__ bool s_last;
  const int tid = threadIdx.x;
  const int row = blockIdx.x * 256 + tid;
  float loss = 0.f;
  int cnt = 0;
  if (row < M) {
    float gm = -INFINITY;
    for (int c = 0; c < n_chunks; ++c) gm = fmaxf(gm, pmax[static_cast<size_t>(c) * Mpad + row]);
    float s = 0.f;
    for (int c = 0; c < n_chunks; ++c) {
      const size_t o = static_cast<size_t>(c) * Mpad + row;
      s += psum[o] * exp2f((pmax[o] - gm) * kLog2e);
    }
    const float l = gm + log2f(s) * kLn2;
    lse[row] = l;
    const int64_t tg = targets[row];
    if (tg >= 0) {
      // a target beyond the vocabulary matched no column, so ztgt[row] was never written: poison the loss
      // (torch's cross_entropy raises a device assert here) instead of reading stale scratch
      loss = (tg < V) ? l - ztgt[row] : __int_as_float(0x7fc00000);
      cnt = 1;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    loss += __shfl_xor_sync(0xffffffffu, loss, o);
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  }
  if ((tid & 31) == 0) {
    s_loss[tid >> 5] = loss;
    s_cnt[tid >> 5] = cnt;
  }
  __syncthreads();
  if (tid == 0) {
    float bl = 0.f;
    int bc = 0;
    for (int w = 0; w < 8; ++w) {
      bl += s_loss[w];
      bc += s_cnt[w];
    }
    blk_loss[blockIdx.x] = bl;
    blk_cnt[blockIdx.x] = bc;
    __threadfence();
    const unsigned int t = atomicAdd(ticket, 1u);
    s_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last && tid == 0) {  // last block to finish: fixed-order sum of the block partials
    __threadfence();
    double tl = 0.0;
    int tc = 0;
    for (unsigned int b = 0; b < gridDim.x; ++b) {
      tl += static_cast<double>(reinterpret_cast<volatile float*>(blk_loss)[b]);
      tc += reinterpret_cast<volatile int*>(blk_cnt)[b];
    }
    // add_to_loss: this launch covers one slab of rows of a larger call; slabs run one after another on the stream
    *loss_sum = add_to_loss ? *loss_sum + static_cast<float>(tl) : static_cast<float>(tl);
    if (n_valid != nullptr) *n_valid = tc;
    *ticket = 0u;
  }
}

// n_valid = number of targets >= 0 (out must be zero on entry; integer atomics, so the result is exact).
__global__ void __launch_bounds__(256) count_valid_kernel(const int64_t* __restrict__ targets, int64_t M,
                                                          int32_t* __restrict__ out) {
  int c = 0;
  for (int64_t i = blockIdx.x * 256ll + threadIdx.x; i < M; i += gridDim.x * 256ll) c += (targets[i] >= 0) ? 1 : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0 && c != 0) atomicAdd(out, c);
}

// buf[i] *= *factor unless the factor is exactly 1 (then every block leaves at once): the fix-up the fused
// forward+backward needs when the upstream gradient turns out not to be the 1.0 it assumed.
template <typename T>
__global__ void __launch_bounds__(256) scale_unless_one_kernel(T* __restrict__ buf, size_t n_vec,
                                                               const float* __restrict__ factor, int poison) {
  float f = __ldg(factor);
  if (f == 1.0f) return;
  if (poison) f = __int_as_float(0x7fc00000);
  uint4* b = reinterpret_cast<uint4*>(buf);
  for (size_t i = blockIdx.x * static_cast<size_t>(256) + threadIdx.x; i < n_vec; i += gridDim.x * static_cast<size_t>(256)) {
    uint4 v = b[i];
    uint32_t* w = reinterpret_cast<uint32_t*>(&v);
    if constexpr (sizeof(T) == 4) {
#pragma unroll
      for (int e = 0; e < 4; ++e) w[e] = __float_as_uint(__uint_as_float(w[e]) * f);
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&w[e]);
        w[e] = pack_bf16x2(__low2float(h) * f, __high2float(h) * f);
      }
    }
    b[i] = v;
  }
}

// dH = sum over the split-K copies, in split order (fixed, so the result does not depend on scheduling), as bf16.
// rows * d is a multiple of 8 (d is a multiple of 64).
__global__ void __launch_bounds__(256) dh_reduce_kernel(const float* __restrict__ partials, int k_split, size_t stride,
                                                        size_t n_vec, __nv_bfloat16* __restrict__ out) {
  for (size_t i = blockIdx.x * static_cast<size_t>(256) + threadIdx.x; i < n_vec; i += gridDim.x * static_cast<size_t>(256)) {
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
    for (int sp = 0; sp < k_split; ++sp) {
      const float4* src = reinterpret_cast<const float4*>(partials + static_cast<size_t>(sp) * stride) + 2 * i;
      const float4 a = __ldcs(src), b = __ldcs(src + 1);
      acc[0] += a.x, acc[1] += a.y, acc[2] += a.z, acc[3] += a.w;
      acc[4] += b.x, acc[5] += b.y, acc[6] += b.z, acc[7] += b.w;
    }
    reinterpret_cast<uint4*>(out)[i] = make_uint4(pack_bf16x2(acc[0], acc[1]), pack_bf16x2(acc[2], acc[3]),
                                                  pack_bf16x2(acc[4], acc[5]), pack_bf16x2(acc[6], acc[7]));
  }
}

// dbias[v] += sum over slab rows of dZ[row, v].  HBM bound: a block covers 256 columns x rows_per_block rows, a lane
// 8 consecutive columns (16-byte loads, a warp reads 512 contiguous bytes of a row, four rows in flight per warp), its
// eight warps take every eighth row; one shared-memory reduction and one atomicAdd per column and block.
__global__ void __launch_bounds__(256) dbias_kernel(const __nv_bfloat16* __restrict__ dz, int rows, int V, int ld,
                                                    int rows_per_block, float* __restrict__ dbias) {
  __shared__ float part[8][257];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int col = blockIdx.x * 256 + lane * 8;
  const int r0 = blockIdx.y * rows_per_block;
  const int r1 = min(rows, r0 + rows_per_block);
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
  if (col < ld) {
#pragma unroll 4
    for (int r = r0 + warp; r < r1; r += 8) {
      const uint4 w = __ldg(reinterpret_cast<const uint4*>(dz + static_cast<size_t>(r) * ld + col));
      const uint32_t u[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&u[e]);
        acc[2 * e] += __low2float(h);
        acc[2 * e + 1] += __high2float(h);
      }
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) part[warp][lane * 8 + e] = acc[e];
  __syncthreads();
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c < V) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += part[w][threadIdx.x];
    atomicAdd(dbias + c, t);
  }
}


// Kept numerators -> dZ in place: dZ[r, c] = e[r, c] * exp(cref[tile(c)][r] - lse[r]) * scale_r - (c == target_r) * scale_r,
// scale_r = 0 on ignored rows.  HBM-bound.  One block per row, 16-byte accesses, four independent loads per thread in
// flight; the per-tile factor is recomputed per vector (32 consecutive lanes share a tile: one broadcast load of cref
// and one MUFU per 8 elements).  No shared memory and 40 registers, so its blocks fit on SMs whose shared memory is
// almost entirely held by a persistent GEMM CTA: the fused path runs it on a second stream beside the next slab's sweep.
__global__ void __launch_bounds__(256) probs_to_dz_kernel(__nv_bfloat16* __restrict__ probs,
                                                          const float* __restrict__ cref,
                                                          const float* __restrict__ lse,
                                                          const int64_t* __restrict__ targets, int rows, int V, int ld,
                                                          int Mpad, int n_tiles, float scale,
                                                          const float* __restrict__ scale_dev,
                                                          const int32_t* __restrict__ nvalid_dev) {
  const int r = blockIdx.x;
  if (r >= rows) return;
  const int tgt = static_cast<int>(targets[r]);
  float sc = scale;
  if (scale_dev != nullptr) sc *= __ldg(scale_dev);
  if (nvalid_dev != nullptr) sc /= static_cast<float>(__ldg(nvalid_dev));
  if (tgt < 0) sc = 0.f;
  const float l2 = -__ldg(lse + r) * kLog2e;
  const float* cr = cref + r;
  uint4* row = reinterpret_cast<uint4*>(probs + static_cast<size_t>(r) * ld);
  const int nvec = ld / 8;  // ld is a multiple of 64
  for (int i0 = threadIdx.x; i0 < nvec; i0 += 256 * 4) {
    uint4 in[4];
    float f[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * 256;
      if (i < nvec) {
        in[u] = __ldcs(row + i);
        const int t = min((i * 8) >> 8, n_tiles - 1);  // 256-column tiles; pad columns past V reuse the last tile
        f[u] = fast_exp2(fmaf(__ldg(cr + static_cast<size_t>(t) * Mpad), kLog2e, l2)) * sc;
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * 256;
      if (i >= nvec) break;
      const int c8 = i * 8;
      const uint32_t iw[4] = {in[u].x, in[u].y, in[u].z, in[u].w};
      uint32_t ow[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&iw[e]);
        float a = __low2float(h) * f[u], b = __high2float(h) * f[u];
        if (c8 + 2 * e == tgt) a -= sc;
        if (c8 + 2 * e + 1 == tgt) b -= sc;
        ow[e] = pack_bf16x2(a, b);
      }
      row[i] = make_uint4(ow[0], ow[1], ow[2], ow[3]);
    }
  }
}

// ------------------------------------------------------------------------------------------------ host side
// SMs the persistent kernels leave idle (FTK_OPT_SM_RESERVE): room for a concurrent NCCL kernel, whose CTAs cannot
// share an SM with ours (registers) and would otherwise wait for a whole GEMM to finish.
inline int sm_reserve() { return ftk::option(FTK_OPT_SM_RESERVE); }

template <int kCG, int BN, int kStages, bool kAMN, bool kBMN, class Epi>
int launch_umma(const CUtensorMap& tmA, const CUtensorMap& tmB, const Sched& sc, const typename Epi::Params& ep,
                int sm_count, cudaStream_t st, int reserve = -1) {
  using L = SmemLayout<kCG, BN, kStages, Epi::kSmemBytes>;
  auto kern = umma_kernel<kCG, BN, kStages, kAMN, kBMN, Epi>;
  static bool attr_done[64] = {false};  // per instantiation and device
  int dev = 0;
  FTK_CUDA_CHECK(cudaGetDevice(&dev));
  if (!attr_done[dev & 63]) {
    FTK_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(L::BYTES)));
    attr_done[dev & 63] = true;
  }
  int usable = sm_count - (reserve >= 0 ? reserve : sm_reserve());
  if (usable < kCG) usable = kCG;
  int n_cl = usable / kCG;
  if (n_cl > sc.n_units) n_cl = sc.n_units;
  if (n_cl <= 0) return FTK_OK;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>(n_cl * kCG));
  cfg.blockDim = dim3(threads_of<Epi>());
  cfg.dynamicSmemBytes = L::BYTES;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kCG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  FTK_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, sc, ep));
  return FTK_OK;
}

// Which CTA grouping to use: 2 (CTA pairs, 256-row tiles) unless ftk_set_option(FTK_OPT_CTA_GROUP, 1) was called.
int cta_group() { return ftk::option(FTK_OPT_CTA_GROUP); }

inline int ceil_div(int64_t a, int64_t b) { return static_cast<int>((a + b - 1) / b); }
inline int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

// Split the vocabulary sweep of each m block into chunks so the (m block x chunk) units fill the machine evenly.
void plan_chunks(int n_mblk, int n_ntile, int n_groups, int* n_chunks, int* tiles_per_chunk) {
  int best_s = 1;
  int64_t best_cost = INT64_MAX;
  const int s_max = n_ntile < kMaxChunks ? n_ntile : kMaxChunks;
  for (int s = 1; s <= s_max; ++s) {
    const int tpc = ceil_div(n_ntile, s);
    const int s_eff = ceil_div(n_ntile, tpc);
    const int64_t units = static_cast<int64_t>(n_mblk) * s_eff;
    const int64_t cost = ((units + n_groups - 1) / n_groups) * tpc;
    if (cost < best_cost) {
      best_cost = cost;
      best_s = s_eff;
    }
  }
  *n_chunks = best_s;
  *tiles_per_chunk = ceil_div(n_ntile, best_s);
  *n_chunks = ceil_div(n_ntile, *tiles_per_chunk);
}

// L2 eviction policy per operand role (a, b, stores): letters n/f/l.  evict_first / evict_last on any role measured
// no different from normal (profiles/r01_l2hint_experiment.txt), so every kernel passes "nnn".
uint64_t hint_of(char c) { return c == 'f' ? kEvictFirst : c == 'l' ? kEvictLast : kEvictNormal; }
void pick_hints(const char* src, int, uint64_t* a, uint64_t* b, uint64_t* s) {
  *a = hint_of(src[0]);
  *b = hint_of(src[1]);
  if (s != nullptr) *s = hint_of(src[2]);
}

// m blocks per scheduling group: as many as keep the group's A blocks (rows x K bf16) within ~24 MB of L2.
int pick_m_group(int n_mblk, int rows_per_mblk, int k_elems) {
  const int64_t blk_bytes = static_cast<int64_t>(rows_per_mblk) * k_elems * 2;
  int64_t g = (int64_t(24) << 20) / (blk_bytes > 0 ? blk_bytes : 1);
  if (g < 1) g = 1;
  if (g > n_mblk) g = n_mblk;
  return static_cast<int>(g);
}

struct FwdWorkspace {
  float* pmax;
  float* psum;
  float* ztgt;
  float* blk_loss;
  int* blk_cnt;
  unsigned int* ticket;
  int64_t Mpad;
  int n_blk;
  size_t bytes;
};

FwdWorkspace carve_fwd(void* base, int64_t M) {
  FwdWorkspace w;
  w.Mpad = round_up(M, 256);
  w.n_blk = static_cast<int>(w.Mpad / 256);
  const uintptr_t p = reinterpret_cast<uintptr_t>(base);  // integer arithmetic: base may be null for a size query
  size_t off = 0;
  w.pmax = reinterpret_cast<float*>(p + off);
  off += sizeof(float) * kMaxChunks * w.Mpad;
  w.psum = reinterpret_cast<float*>(p + off);
  off += sizeof(float) * kMaxChunks * w.Mpad;
  w.ztgt = reinterpret_cast<float*>(p + off);
  off += sizeof(float) * w.Mpad;
  w.blk_loss = reinterpret_cast<float*>(p + off);
  off += round_up(sizeof(float) * w.n_blk, 256);
  w.blk_cnt = reinterpret_cast<int*>(p + off);
  off += round_up(sizeof(int) * w.n_blk, 256);
  w.ticket = reinterpret_cast<unsigned int*>(p + off);
  off += 256;
  w.bytes = off;
  return w;
}

template <int kCG>
int run_fwd(const void* H, const void* W, const void* bias, const int64_t* targets, int64_t M, int d, int V,
            float* lse, float* loss_sum, int32_t* n_valid, const FwdWorkspace& ws, int sm_count, cudaStream_t st) {
  constexpr int BN = 256;
  constexpr int kStages = (kCG == 1) ? 4 : 6;
  CUtensorMap tmA, tmB;
  if (int rc = make_tmap(&tmA, H, d, M, static_cast<uint64_t>(d) * 2, BK, BM)) return rc;
  if (int rc = make_tmap(&tmB, W, d, V, static_cast<uint64_t>(d) * 2, BK, BN / kCG)) return rc;
  Sched sc;
  sc.k_split = 1;
  const int n_mblk = ceil_div(M, BM * kCG);
  sc.n_ntile = ceil_div(V, BN);
  int n_chunks;
  plan_chunks(n_mblk, sc.n_ntile, (sm_count - sm_reserve() > kCG ? sm_count - sm_reserve() : kCG) / kCG, &n_chunks,
              &sc.tiles_per_chunk);
  sc.n_units = n_mblk * n_chunks;
  sc.n_mblk = n_mblk;
  sc.n_chunks = n_chunks;
  sc.m_group = pick_m_group(n_mblk, BM * kCG, d);
  sc.n_kblk = ceil_div(d, BK);
  pick_hints("nnn", 0, &sc.hint_a, &sc.hint_b, nullptr);
  FwdEpi::Params ep;
  ep.targets = targets;
  ep.bias = static_cast<const __nv_bfloat16*>(bias);
  ep.pmax = ws.pmax;
  ep.psum = ws.psum;
  ep.ztgt = ws.ztgt;
  ep.M = static_cast<int>(M);
  ep.V = V;
  ep.Mpad = static_cast<int>(ws.Mpad);
  FTK_CUDA_CHECK(cudaMemsetAsync(ws.ticket, 0, sizeof(unsigned int), st));
{
    ftk::TimelineScope tl_(FTK_TL_FWD, st);
    if (int rc = launch_umma<kCG, BN, kStages, false, false, FwdEpi>(tmA, tmB, sc, ep, sm_count, st)) return rc;
  }
  fwd_finalize_kernel<<<ceil_div(M, 256), 256, 0, st>>>(ws.pmax, ws.psum, ws.ztgt, targets, static_cast<int>(M),
                                                        static_cast<int>(ws.Mpad), V, n_chunks, lse, ws.blk_loss,
                                                        ws.blk_cnt, ws.ticket, loss_sum, n_valid, 0);
  FTK_CUDA_CHECK(cudaGetLastError());
  return FTK_OK;
}

// Forward that keeps the softmax numerators (FwdKeepEpi) for a recompute-free backward.
template <int kCG>
int run_fwd_keep(const void* H, const void* W, const void* bias, const int64_t* targets, int64_t M, int d, int V,
                 float* lse, float* loss_sum, int32_t* n_valid, __nv_bfloat16* probs, float* cref,
                 const FwdWorkspace& ws, int sm_count, cudaStream_t st, int add_to_loss = 0, int reserve = -1) {
  constexpr int BN = 256;
  constexpr int kStages = (kCG == 1) ? 4 : 6;
  const int64_t ld = round_up(V, 64);
  CUtensorMap tmA, tmB;
  if (int rc = make_tmap(&tmA, H, d, M, static_cast<uint64_t>(d) * 2, BK, BM)) return rc;
  if (int rc = make_tmap(&tmB, W, d, V, static_cast<uint64_t>(d) * 2, BK, BN / kCG)) return rc;
  Sched sc;
  sc.k_split = 1;
  const int n_mblk = ceil_div(M, BM * kCG);
  sc.n_ntile = ceil_div(V, BN);
  int n_chunks;
  const int usable = sm_count - (reserve >= 0 ? reserve : sm_reserve());
  plan_chunks(n_mblk, sc.n_ntile, (usable > kCG ? usable : kCG) / kCG, &n_chunks, &sc.tiles_per_chunk);
  sc.n_units = n_mblk * n_chunks;
  sc.n_mblk = n_mblk;
  sc.n_chunks = n_chunks;
  sc.m_group = pick_m_group(n_mblk, BM * kCG, d);
  sc.n_kblk = ceil_div(d, BK);
  FwdKeepEpi::Params ep;
  pick_hints("nnn", 0, &sc.hint_a, &sc.hint_b, &ep.store_hint);
  if (int rc = make_tmap(&ep.tm_out, probs, ld, M, static_cast<uint64_t>(ld) * 2, 64, 32)) return rc;
  ep.targets = targets;
  ep.bias = static_cast<const __nv_bfloat16*>(bias);
  ep.pmax = ws.pmax;
  ep.psum = ws.psum;
  ep.ztgt = ws.ztgt;
  ep.cref = cref;
  ep.M = static_cast<int>(M);
  ep.V = V;
  ep.Mpad = static_cast<int>(ws.Mpad);
  ep.ld = static_cast<int>(ld);
  FTK_CUDA_CHECK(cudaMemsetAsync(ws.ticket, 0, sizeof(unsigned int), st));
{
    ftk::TimelineScope tl_(FTK_TL_FWD_KEEP, st);
    if (int rc = launch_umma<kCG, BN, kStages, false, false, FwdKeepEpi>(tmA, tmB, sc, ep, sm_count, st, reserve)) return rc;
  }
  fwd_finalize_kernel<<<ceil_div(M, 256), 256, 0, st>>>(ws.pmax, ws.psum, ws.ztgt, targets, static_cast<int>(M),
                                                        static_cast<int>(ws.Mpad), V, n_chunks, lse, ws.blk_loss,
                                                        ws.blk_cnt, ws.ticket, loss_sum, n_valid, add_to_loss);
  FTK_CUDA_CHECK(cudaGetLastError());
  return FTK_OK;
}

constexpr int64_t kDefaultSlabRows = 16384;
inline int64_t dz_pitch(int V) { return round_up(V, 64); }  // bf16 elements; rows start 128-byte aligned

// How the two gradient GEMMs of one slab are issued.  Default: dH then dW, both on `sm_count - FTK_OPT_SM_RESERVE`
// SMs.  overlap_tail (the last slab of a data-parallel caller's fused call): dW first on every SM, `dw_done`
// recorded behind it, then dH on `sm_count - FTK_OPT_SM_RESERVE` SMs, so the all-reduce of dW can start on the SMs
// dH leaves idle while dH is still running.
struct GradGemmPlan {
  bool overlap_tail = false;
  bool full_sms = false;  // ignore the SM reserve for both GEMMs (slabs whose gradients nobody is waiting for yet)
  cudaEvent_t dw_done = nullptr;
  float* dh_partials = nullptr;  // scratch for a split-K dH GEMM (plan_dh_split), or null: never split
  size_t dh_partial_bytes = 0;
};

template <int kCG>
int run_grad_gemms(const void* H, const void* W, int64_t rows, int d, int V, void* dH, void* dW, int dw_dtype,
                   int accumulate, float* dbias, __nv_bfloat16* dz, int sm_count, cudaStream_t st,
                   const GradGemmPlan& plan);

template <int kCG>
int run_bwd_slab(const void* H, const void* W, const void* bias, const int64_t* targets, const float* lse,
                 int64_t rows, int d, int V, float grad_scale, const float* grad_scale_dev, const int32_t* n_valid_dev,
                 void* dH, void* dW, int dw_dtype, int accumulate, float* dbias, __nv_bfloat16* dz, int sm_count,
                 cudaStream_t st) {
  constexpr int BN = 256;
  constexpr int kStages = (kCG == 1) ? 4 : 6;
  const int64_t ld = dz_pitch(V);
  const int n_groups = sm_count / kCG;

  // ---- stage 1: dZ slab = (softmax - onehot) * scale, logits recomputed on the tensor cores
  {
    CUtensorMap tmA, tmB;
    if (int rc = make_tmap(&tmA, H, d, rows, static_cast<uint64_t>(d) * 2, BK, BM)) return rc;
    if (int rc = make_tmap(&tmB, W, d, V, static_cast<uint64_t>(d) * 2, BK, BN / kCG)) return rc;
    Sched sc;
    sc.k_split = 1;
    const int n_mblk = ceil_div(rows, BM * kCG);
    sc.n_ntile = ceil_div(V, BN);
    int n_chunks;
    plan_chunks(n_mblk, sc.n_ntile, n_groups, &n_chunks, &sc.tiles_per_chunk);
    sc.n_units = n_mblk * n_chunks;
    sc.n_mblk = n_mblk;
    sc.n_chunks = n_chunks;
    sc.m_group = pick_m_group(n_mblk, BM * kCG, d);
    sc.n_kblk = ceil_div(d, BK);
    DzEpi::Params ep;
    pick_hints("nnn", 1, &sc.hint_a, &sc.hint_b, &ep.store_hint);
    ep.targets = targets;
    ep.bias = static_cast<const __nv_bfloat16*>(bias);
    ep.lse = lse;
    if (int rc = make_tmap(&ep.tm_out, dz, ld, rows, static_cast<uint64_t>(ld) * 2, 64, 32)) return rc;
    ep.scale_dev = grad_scale_dev;
    ep.nvalid_dev = n_valid_dev;
    ep.scale = grad_scale;
    ep.rows = static_cast<int>(rows);
    ep.V = V;
    ep.ld = static_cast<int>(ld);
    ftk::TimelineScope tl_(FTK_TL_DZ, st);
    if (int rc = launch_umma<kCG, BN, kStages, false, false, DzEpi>(tmA, tmB, sc, ep, sm_count, st)) return rc;
  }
  return run_grad_gemms<kCG>(H, W, rows, d, V, dH, dW, dw_dtype, accumulate, dbias, dz, sm_count, st, GradGemmPlan());
}

// Split-K plan of the dH GEMM.  Its units are (row block x N tile) with the whole vocabulary as K: at d = 768 a slab
// of 10,496 rows is 41 x 3 = 123 units, 1.66 waves over 74 SM pairs charged as 2.  Cutting every unit into k_split
// pieces over disjoint vocabulary ranges fills the last wave (123 x 3 = 369 = 4.99 waves of a third the length); each
// piece writes fp32 partial sums to its own copy of the output and dh_reduce_kernel adds the copies in a fixed order
// (deterministic, no atomics).  Used when it cuts the charged time by more than 8 % and the copies stay small.
constexpr size_t kDhPartialBudget = size_t(160) << 20;
inline int plan_dh_split(int64_t rows, int d, int V, int n_groups, int kCG) {
  const int64_t units = static_cast<int64_t>(ceil_div(rows, BM * kCG)) * ceil_div(d, 256);
  const int n_kblk = ceil_div(V, BK);
  int best = 1;
  double best_t = static_cast<double>(ceil_div(units, n_groups)) * n_kblk;
  for (int ks = 2; ks <= 8; ++ks) {
    if (n_kblk < 16 * ks) break;  // every split must own K blocks
    if (static_cast<size_t>(ks) * round_up(rows, BM * kCG) * round_up(d, 256) * 4 > kDhPartialBudget) break;
    // each piece pays its own pipeline fill and an fp32 epilogue: ~24 K blocks' worth
    const double t = static_cast<double>(ceil_div(units * ks, n_groups)) * (static_cast<double>(ceil_div(n_kblk, ks)) + 24);
    if (t < best_t * 0.92 && t < static_cast<double>(ceil_div(units, n_groups)) * n_kblk * 0.92) {
      best_t = t;
      best = ks;
    }
  }
  return best;
}
inline size_t dh_partial_bytes(int64_t rows, int d, int V, int n_groups, int kCG) {
  const int ks = plan_dh_split(rows, d, V, n_groups, kCG);
  return ks > 1 ? static_cast<size_t>(ks) * round_up(rows, BM * kCG) * d * 4 : 0;
}

// dH[rows, d] = dZ[rows, V] . W[V, d]: A K-major (K = V, the tail zero-filled by TMA), B MN-major.
// partials != null and k_split > 1: split-K into fp32 copies (see plan_dh_split), reduced into dH afterwards.
template <int kCG, int BN>
int run_dh(const void* W, int64_t rows, int d, int V, void* dH, __nv_bfloat16* dz, int sm_count, cudaStream_t st,
           int reserve, float* partials = nullptr, int k_split = 1) {
  constexpr int kStages = (BN == 256) ? ((kCG == 1) ? 4 : 6) : ((kCG == 1) ? 6 : 8);
  const int64_t ld = dz_pitch(V);
  CUtensorMap tmA, tmB;
  if (int rc = make_tmap(&tmA, dz, V, rows, static_cast<uint64_t>(ld) * 2, BK, BM)) return rc;
  if (int rc = make_tmap(&tmB, W, d, V, static_cast<uint64_t>(d) * 2, 64, BK)) return rc;
  if (partials == nullptr) k_split = 1;
  Sched sc;
  const int n_mblk = ceil_div(rows, BM * kCG);
  sc.n_ntile = ceil_div(d, BN);
  sc.tiles_per_chunk = 1;
  sc.k_split = k_split;
  sc.n_chunks = sc.n_ntile * k_split;  // (N tile, split) pairs
  sc.n_units = n_mblk * sc.n_chunks;
  sc.n_mblk = n_mblk;
  sc.m_group = 1;  // all pieces of one m block are adjacent: its dZ rows stream from HBM once
  sc.n_kblk = ceil_div(V, BK);
  pick_hints("nnn", 2, &sc.hint_a, &sc.hint_b, nullptr);
  ftk::TimelineScope tl_(FTK_TL_DH, st);
  if (k_split > 1) {
    StoreEpi<true>::Params ep;
    ep.out = partials;
    ep.rows = static_cast<int>(rows);
    ep.cols = d;
    ep.ld = d;
    ep.accumulate = 0;
    ep.split_stride = static_cast<size_t>(round_up(rows, BM * kCG)) * d;
    if (int rc = launch_umma<kCG, BN, kStages, false, true, StoreEpi<true>>(tmA, tmB, sc, ep, sm_count, st, reserve))
      return rc;
    const size_t n_vec = static_cast<size_t>(rows) * d / 8;
    size_t blocks = (n_vec + 255) / 256;
    if (blocks > static_cast<size_t>(sm_count) * 8) blocks = static_cast<size_t>(sm_count) * 8;
    dh_reduce_kernel<<<static_cast<unsigned>(blocks), 256, 0, st>>>(partials, k_split, ep.split_stride, n_vec,
                                                                  static_cast<__nv_bfloat16*>(dH));
    FTK_CUDA_CHECK(cudaGetLastError());
    return FTK_OK;
  }
  StoreEpi<false>::Params ep;
  ep.out = dH;
  ep.rows = static_cast<int>(rows);
  ep.cols = d;
  ep.ld = d;
  ep.accumulate = 0;
  ep.split_stride = 0;
  return launch_umma<kCG, BN, kStages, false, true, StoreEpi<false>>(tmA, tmB, sc, ep, sm_count, st, reserve);
}

// dW[V, d] (+)= dZ^T[V, rows] . H[rows, d]: A and B MN-major.
template <int kCG>
int run_dw(const void* H, int64_t rows, int d, int V, void* dW, int dw_dtype, int accumulate, __nv_bfloat16* dz,
           int sm_count, cudaStream_t st, int reserve) {
  constexpr int BN = 256;
  constexpr int kStages = (kCG == 1) ? 4 : 6;
  const int64_t ld = dz_pitch(V);
  CUtensorMap tmA, tmB;
  if (int rc = make_tmap(&tmA, dz, V, rows, static_cast<uint64_t>(ld) * 2, 64, BK)) return rc;
  if (int rc = make_tmap(&tmB, H, d, rows, static_cast<uint64_t>(d) * 2, 64, BK)) return rc;
  Sched sc;
  sc.k_split = 1;
  const int n_mblk = ceil_div(V, BM * kCG);
  sc.n_ntile = ceil_div(d, BN);
  sc.tiles_per_chunk = 1;
  sc.n_units = n_mblk * sc.n_ntile;
  sc.n_mblk = n_mblk;
  sc.n_chunks = sc.n_ntile;
  sc.m_group = 1;
  sc.n_kblk = ceil_div(rows, BK);
  pick_hints("nnn", 3, &sc.hint_a, &sc.hint_b, nullptr);
  ftk::TimelineScope tl_(FTK_TL_DW, st);
  if (dw_dtype == FTK_F32) {
    StoreEpi<true>::Params ep;
    ep.out = dW;
    ep.rows = V;
    ep.cols = d;
    ep.ld = d;
    ep.accumulate = accumulate;
  ep.split_stride = 0;
    return launch_umma<kCG, BN, kStages, true, true, StoreEpi<true>>(tmA, tmB, sc, ep, sm_count, st, reserve);
  }
  StoreEpi<false>::Params ep;
  ep.out = dW;
  ep.rows = V;
  ep.cols = d;
  ep.ld = d;
  ep.accumulate = accumulate;
  ep.split_stride = 0;
  return launch_umma<kCG, BN, kStages, true, true, StoreEpi<false>>(tmA, tmB, sc, ep, sm_count, st, reserve);
}

// Tile width of the dH GEMM.  Its units are (row block x N tile) with the whole vocabulary as K, so few rows give
// few, long units.  128-wide tiles double the unit count but move 1.5x the operand bytes per flop from L2 to shared
// memory, which is what bounds these kernels at full clock: measured 1.6x slower per flop (8192 rows, d = 2048:
// 7 half-waves of 128 took 3.9 ms against 2.8 ms for 4 waves of 256).  They pay only when the 256-wide launch would
// leave most SM pairs idle (a few hundred rows).
constexpr int kNarrowTilePenaltyPct = 160;
inline int pick_dh_bn(int64_t rows, int d, int n_groups, int kCG) {
  const int64_t n_mblk = ceil_div(rows, BM * kCG);
  const int64_t w256 = ceil_div(n_mblk * ceil_div(d, 256), n_groups) * 2;
  const int64_t w128 = ceil_div(n_mblk * ceil_div(d, 128), n_groups);
  return (w128 * kNarrowTilePenaltyPct < w256 * 100) ? 128 : 256;
}

template <int kCG>
int run_grad_gemms(const void* H, const void* W, int64_t rows, int d, int V, void* dH, void* dW, int dw_dtype,
                   int accumulate, float* dbias, __nv_bfloat16* dz, int sm_count, cudaStream_t st,
                   const GradGemmPlan& plan) {
  const int res_dh = plan.full_sms ? 0 : -1;
  const int res_dw = (plan.full_sms || plan.overlap_tail) ? 0 : -1;
  const int usable = sm_count - (res_dh == 0 ? 0 : sm_reserve());
  const int n_groups = (usable > kCG ? usable : kCG) / kCG;
  const int dh_bn = pick_dh_bn(rows, d, n_groups, kCG);
  if (plan.overlap_tail) {
    if (int rc = run_dw<kCG>(H, rows, d, V, dW, dw_dtype, accumulate, dz, sm_count, st, res_dw)) return rc;
    if (plan.dw_done != nullptr) FTK_CUDA_CHECK(cudaEventRecord(plan.dw_done, st));
  }
  if (dh_bn == 128) {
    if (int rc = run_dh<kCG, 128>(W, rows, d, V, dH, dz, sm_count, st, res_dh)) return rc;
  } else {
    int ks = 1;
    if (plan.dh_partials != nullptr) {
      ks = plan_dh_split(rows, d, V, n_groups, kCG);
      if (static_cast<size_t>(ks) * round_up(rows, BM * kCG) * d * 4 > plan.dh_partial_bytes) ks = 1;
    }
    if (int rc = run_dh<kCG, 256>(W, rows, d, V, dH, dz, sm_count, st, res_dh, plan.dh_partials, ks)) return rc;
  }
  if (!plan.overlap_tail)
    if (int rc = run_dw<kCG>(H, rows, d, V, dW, dw_dtype, accumulate, dz, sm_count, st, res_dw)) return rc;
  // ---- optional: dbias += column sums of the slab
  if (dbias != nullptr) {
    ftk::TimelineScope tl_(FTK_TL_DBIAS, st);
    const int rows_per_block = 256;
    dim3 grid(ceil_div(V, 256), ceil_div(rows, rows_per_block));
    dbias_kernel<<<grid, 256, 0, st>>>(dz, static_cast<int>(rows), V, static_cast<int>(dz_pitch(V)), rows_per_block,
                                       dbias);
    FTK_CUDA_CHECK(cudaGetLastError());
  }
  return FTK_OK;
}

int check_common(const void* H, const void* W, int64_t M, int d, int V, ftk::DeviceInfo* di) {
  FTK_REQUIRE(H && W, "lmhead_ce: null H or W");
  FTK_REQUIRE(M > 0 && M < (int64_t(1) << 30), "lmhead_ce: M=%lld out of range", static_cast<long long>(M));
  FTK_REQUIRE(d > 0 && d % 64 == 0, "lmhead_ce: d=%d must be a positive multiple of 64", d);
  FTK_REQUIRE(V > 0 && V <= (1 << 24), "lmhead_ce: V=%d out of range", V);
  if (int rc = ftk::get_device_info(di)) return rc;
  if (di->cc_major != 10)
    return ftk::fail(FTK_EUNSUPPORTED, "lmhead_ce needs an sm_100 device (tcgen05/TMEM); found sm_%d%d", di->cc_major,
                     di->cc_minor);
  return FTK_OK;
}

}  // namespace

extern "C" size_t ftk_lmhead_ce_fwd_workspace(int64_t M, int32_t d, int32_t V) {
  (void)d;
  (void)V;
  if (M <= 0) return 0;
  return carve_fwd(nullptr, M).bytes;
}

extern "C" int ftk_lmhead_ce_fwd(const void* H, const void* W, const void* bias, const int64_t* targets, int64_t M,
                                 int32_t d, int32_t V, float* lse, float* loss_sum, int32_t* n_valid, void* workspace,
                                 size_t workspace_bytes, void* stream) {
  ftk::DeviceInfo di;
  if (int rc = check_common(H, W, M, d, V, &di)) return rc;
  FTK_REQUIRE(targets && lse && loss_sum && n_valid && workspace, "ftk_lmhead_ce_fwd: null pointer argument");
  FTK_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "ftk_lmhead_ce_fwd: workspace must be 256-byte aligned");
  const FwdWorkspace ws = carve_fwd(workspace, M);
  if (workspace_bytes < ws.bytes)
    return ftk::fail(FTK_EWORKSPACE, "ftk_lmhead_ce_fwd: workspace %zu bytes < required %zu", workspace_bytes,
                     ws.bytes);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (cta_group() == 2 && di.sm_count >= 2)
    return run_fwd<2>(H, W, bias, targets, M, d, V, lse, loss_sum, n_valid, ws, di.sm_count, st);
  return run_fwd<1>(H, W, bias, targets, M, d, V, lse, loss_sum, n_valid, ws, di.sm_count, st);
}

extern "C" size_t ftk_lmhead_ce_bwd_workspace(int64_t M, int32_t d, int32_t V, int64_t slab_rows) {
  (void)d;
  if (M <= 0 || V <= 0) return 0;
  int64_t rows = slab_rows > 0 ? slab_rows : kDefaultSlabRows;
  if (rows > M) rows = M;
  rows = round_up(rows, 256);
  return static_cast<size_t>(rows) * static_cast<size_t>(dz_pitch(V)) * 2;
}

extern "C" int ftk_lmhead_ce_bwd(const void* H, const void* W, const void* bias, const int64_t* targets,
                                 const float* lse, int64_t M, int32_t d, int32_t V, float grad_scale,
                                 const float* grad_scale_dev, const int32_t* n_valid_dev, void* dH, void* dW,
                                 int32_t dw_dtype, int32_t accumulate, float* dbias, void* workspace,
                                 size_t workspace_bytes, void* stream) {
  ftk::DeviceInfo di;
  if (int rc = check_common(H, W, M, d, V, &di)) return rc;
  FTK_REQUIRE(targets && lse && dH && dW && workspace, "ftk_lmhead_ce_bwd: null pointer argument");
  FTK_REQUIRE(dw_dtype == FTK_BF16 || dw_dtype == FTK_F32, "ftk_lmhead_ce_bwd: bad dw_dtype %d", dw_dtype);
  FTK_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "ftk_lmhead_ce_bwd: workspace must be 256-byte aligned");
  FTK_REQUIRE((reinterpret_cast<uintptr_t>(dH) & 15) == 0 && (reinterpret_cast<uintptr_t>(dW) & 15) == 0,
              "ftk_lmhead_ce_bwd: dH/dW must be 16-byte aligned");
  const int64_t ld = dz_pitch(V);
  int64_t slab = static_cast<int64_t>(workspace_bytes / (static_cast<size_t>(ld) * 2));
  if (slab >= M)
    slab = M;
  else
    slab = slab / 256 * 256;
  if (slab <= 0)
    return ftk::fail(FTK_EWORKSPACE, "ftk_lmhead_ce_bwd: workspace %zu bytes holds no 256-row dZ slab (pitch %lld)",
                     workspace_bytes, static_cast<long long>(ld));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dbias != nullptr && !accumulate) FTK_CUDA_CHECK(cudaMemsetAsync(dbias, 0, sizeof(float) * V, st));
  const bool pair = (cta_group() == 2 && di.sm_count >= 2);
  const size_t dh_esize = 2;
  for (int64_t r0 = 0; r0 < M; r0 += slab) {
    const int64_t rows = (M - r0 < slab) ? (M - r0) : slab;
    const void* Hs = static_cast<const uint8_t*>(H) + static_cast<size_t>(r0) * d * 2;
    void* dHs = static_cast<uint8_t*>(dH) + static_cast<size_t>(r0) * d * dh_esize;
    const int acc = (accumulate || r0 > 0) ? 1 : 0;
    int rc;
    if (pair)
      rc = run_bwd_slab<2>(Hs, W, bias, targets + r0, lse + r0, rows, d, V, grad_scale, grad_scale_dev, n_valid_dev,
                           dHs, dW, dw_dtype, acc, dbias, static_cast<__nv_bfloat16*>(workspace), di.sm_count, st);
    else
      rc = run_bwd_slab<1>(Hs, W, bias, targets + r0, lse + r0, rows, d, V, grad_scale, grad_scale_dev, n_valid_dev,
                           dHs, dW, dw_dtype, acc, dbias, static_cast<__nv_bfloat16*>(workspace), di.sm_count, st);
    if (rc) return rc;
  }
  return FTK_OK;
}

// ------------------------------------------------------------------------------------------------
// Opt-in pair: the forward keeps the softmax numerators, the backward skips the logit recompute.
extern "C" size_t ftk_lmhead_ce_probs_bytes(int64_t M, int32_t V) {
  if (M <= 0 || V <= 0) return 0;
  return static_cast<size_t>(M) * static_cast<size_t>(round_up(V, 64)) * 2;
}

extern "C" size_t ftk_lmhead_ce_cref_bytes(int64_t M, int32_t V) {
  if (M <= 0 || V <= 0) return 0;
  return static_cast<size_t>(ceil_div(V, 256)) * static_cast<size_t>(round_up(M, 256)) * sizeof(float);
}

extern "C" int ftk_lmhead_ce_fwd_keep(const void* H, const void* W, const void* bias, const int64_t* targets,
                                      int64_t M, int32_t d, int32_t V, float* lse, float* loss_sum, int32_t* n_valid,
                                      void* probs, float* cref, void* workspace, size_t workspace_bytes,
                                      void* stream) {
  ftk::DeviceInfo di;
  if (int rc = check_common(H, W, M, d, V, &di)) return rc;
  FTK_REQUIRE(targets && lse && loss_sum && n_valid && probs && cref && workspace,
              "ftk_lmhead_ce_fwd_keep: null pointer argument");
  FTK_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0 && (reinterpret_cast<uintptr_t>(probs) & 127) == 0,
              "ftk_lmhead_ce_fwd_keep: workspace must be 256-byte and probs 128-byte aligned");
  const FwdWorkspace ws = carve_fwd(workspace, M);
  if (workspace_bytes < ws.bytes)
    return ftk::fail(FTK_EWORKSPACE, "ftk_lmhead_ce_fwd_keep: workspace %zu bytes < required %zu", workspace_bytes,
                     ws.bytes);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (cta_group() == 2 && di.sm_count >= 2)
    return run_fwd_keep<2>(H, W, bias, targets, M, d, V, lse, loss_sum, n_valid, static_cast<__nv_bfloat16*>(probs),
                           cref, ws, di.sm_count, st);
  return run_fwd_keep<1>(H, W, bias, targets, M, d, V, lse, loss_sum, n_valid, static_cast<__nv_bfloat16*>(probs), cref,
                         ws, di.sm_count, st);
}

extern "C" int ftk_lmhead_ce_bwd_kept(const void* H, const void* W, const int64_t* targets, const float* lse,
                                      void* probs, const float* cref, int64_t M, int32_t d, int32_t V,
                                      float grad_scale, const float* grad_scale_dev, const int32_t* n_valid_dev,
                                      void* dH, void* dW, int32_t dw_dtype, int32_t accumulate, float* dbias,
                                      void* stream) {
  ftk::DeviceInfo di;
  if (int rc = check_common(H, W, M, d, V, &di)) return rc;
  FTK_REQUIRE(targets && lse && probs && cref && dH && dW, "ftk_lmhead_ce_bwd_kept: null pointer argument");
  FTK_REQUIRE(dw_dtype == FTK_BF16 || dw_dtype == FTK_F32, "ftk_lmhead_ce_bwd_kept: bad dw_dtype %d", dw_dtype);
  FTK_REQUIRE((reinterpret_cast<uintptr_t>(dH) & 15) == 0 && (reinterpret_cast<uintptr_t>(dW) & 15) == 0,
              "ftk_lmhead_ce_bwd_kept: dH/dW must be 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t ld = dz_pitch(V);
  if (dbias != nullptr && !accumulate) FTK_CUDA_CHECK(cudaMemsetAsync(dbias, 0, sizeof(float) * V, st));
  {
    const int n_tiles = ceil_div(V, 256);
    ftk::TimelineScope tl_(FTK_TL_CONVERT, st);
    probs_to_dz_kernel<<<static_cast<unsigned>(M), 256, 0, st>>>(
        static_cast<__nv_bfloat16*>(probs), cref, lse, targets, static_cast<int>(M), V, static_cast<int>(ld),
        static_cast<int>(round_up(M, 256)), n_tiles, grad_scale, grad_scale_dev, n_valid_dev);
    FTK_CUDA_CHECK(cudaGetLastError());
  }
  if (cta_group() == 2 && di.sm_count >= 2)
    return run_grad_gemms<2>(H, W, M, d, V, dH, dW, dw_dtype, accumulate, dbias, static_cast<__nv_bfloat16*>(probs),
                             di.sm_count, st, GradGemmPlan());
  return run_grad_gemms<1>(H, W, M, d, V, dH, dW, dw_dtype, accumulate, dbias, static_cast<__nv_bfloat16*>(probs),
                           di.sm_count, st, GradGemmPlan());
}

// ------------------------------------------------------------------------------------------------
// Forward and backward in one call, with a bounded numerators slab (ftk.h: ftk_lmhead_ce_fused).
namespace {

constexpr size_t kFusedSlabBudget = size_t(1) << 30;  // bytes of one bf16 numerators slab; two are alive at a time

// Charged time of the forward sweep and the dH GEMM over `r` rows, in 256 x 256 x 64 MMA blocks per SM pair: whole
// waves over the SM pairs, so a launch that needs 3.46 waves is charged 4.
int64_t slab_cost(int64_t r, int d, int V, int n_groups, int kCG) {
  const int n_mblk = ceil_div(r, BM * kCG);
  int n_chunks, tpc;
  plan_chunks(n_mblk, ceil_div(V, 256), n_groups, &n_chunks, &tpc);
  int64_t cost = static_cast<int64_t>(ceil_div(static_cast<int64_t>(n_mblk) * n_chunks, n_groups)) * tpc * (d / 64);
  {  // dH: whole waves of (row block x N tile) units, split over K when that fills the last wave (plan_dh_split)
    const int64_t units = static_cast<int64_t>(n_mblk) * ceil_div(d, 256);
    const int ks = plan_dh_split(r, d, V, n_groups, kCG);
    const int64_t n_kblk = ceil_div(V, 64);
    cost += ceil_div(units * ks, n_groups) * (ceil_div(n_kblk, ks) + (ks > 1 ? 24 : 0));
  }
  // dW: K = the slab's rows.  The work is the same whatever the split, but a 256 x 256 output tile's epilogue
  // (TMEM read, read-modify-write of dW) takes about as long as 48 K blocks of MMAs, so slabs under ~3,000 rows make
  // the dW GEMM epilogue-paced (measured: 4,096 rows as 2 x 2,048 is 30 % slower than as one slab).
  const int64_t k_dw = ceil_div(r, 64);
  cost += static_cast<int64_t>(ceil_div(static_cast<int64_t>(ceil_div(V, BM * kCG)) * ceil_div(d, 256), n_groups)) *
          (k_dw > 48 ? k_dw : 48);
  return cost;
}

// Rows per slab: the rows are cut into slabs of R rows plus a remainder, R a multiple of the 256-row tile within the
// byte budget; R is the value whose forward and dH launches waste the least of their last waves (the dW launches do
// the same total work whatever the split).  At M = 16,384, d = 2048, 148 SMs: R = 9,472 = 37 row blocks, whose
// 37 x 8 dH tiles are exactly 4 waves over 74 SM pairs and whose forward is 74 units of half the vocabulary each.
int64_t plan_fused_slab(int64_t M, int d, int V, int sm_count, int kCG) {
  const int n_groups = (sm_count > kCG ? sm_count : kCG) / kCG;
  const int64_t ld = dz_pitch(V);
  const int64_t step = BM * kCG;
  int64_t cap = static_cast<int64_t>(kFusedSlabBudget / (static_cast<size_t>(ld) * 2)) / step * step;
  if (cap < step) cap = step;
  const int64_t m_pad = round_up(M, step);
  if (cap > m_pad) cap = m_pad;
  int64_t best_rows = cap, best_cost = INT64_MAX;
  for (int64_t rows = cap; rows >= step; rows -= step) {
    const int64_t n_full = M / rows, rem = M - n_full * rows;
    if (n_full + (rem > 0) > 256) break;  // ever smaller slabs only add dW read-modify-write passes
    int64_t cost = n_full * slab_cost(rows, d, V, n_groups, kCG);
    if (rem > 0) cost += slab_cost(rem, d, V, n_groups, kCG);
    // each extra slab re-reads and re-writes dW once: about V * d * 8 bytes of HBM traffic, charged as tensor time
    // at ~100 flop per byte (it overlaps the dW GEMM)
    cost += (n_full + (rem > 0) - 1) * (static_cast<int64_t>(V) * d * 8 * 100 / (2ll * 256 * 256 * 64)) / n_groups;
    // a single slab has no other slab's sweep to hide its conversion pass behind: 4 bytes of HBM traffic per
    // (row, column) at ~6.2 TB/s, in units of the 0.45 us one SM pair takes for a 256 x 256 x 64 block
    if (n_full + (rem > 0) == 1) cost += static_cast<int64_t>(static_cast<double>(M) * ld * 1.45e-6);
    if (cost < best_cost) {
      best_cost = cost;
      best_rows = rows;
    }
  }
  return best_rows;
}

struct FusedWorkspace {
  void* fwd;
  float* cref[2];
  __nv_bfloat16* probs[2];
  float* dh_partials;
  size_t dh_partial_bytes;
  int64_t slab;
  int n_buf;
  size_t fwd_bytes, bytes;
};

// One set of forward statistics (a slab's finalize has run before the next sweep starts), and a numerators slab plus
// its reference-maxima table per buffer: two buffers when the rows make two or more slabs, so that the conversion of
// one slab can run on the side stream while the next slab's sweep fills the other.
FusedWorkspace carve_fused(void* base, int64_t M, int d, int V, int64_t slab, int sm_count, int kCG) {
  FusedWorkspace w;
  if (slab > M) slab = M;
  w.slab = slab;
  w.n_buf = (M > slab) ? 2 : 1;
  const uintptr_t p = reinterpret_cast<uintptr_t>(base);
  size_t off = 0;
  w.fwd = reinterpret_cast<void*>(p + off);
  w.fwd_bytes = carve_fwd(nullptr, slab).bytes;
  off += round_up(w.fwd_bytes, 256);
  for (int b = 0; b < 2; ++b) {
    if (b >= w.n_buf) {
      w.cref[b] = w.cref[0];
      w.probs[b] = w.probs[0];
      continue;
    }
    w.cref[b] = reinterpret_cast<float*>(p + off);
    off += round_up(static_cast<size_t>(ceil_div(V, 256)) * round_up(slab, 256) * sizeof(float), 256);
    w.probs[b] = reinterpret_cast<__nv_bfloat16*>(p + off);
    off += static_cast<size_t>(round_up(slab, 256)) * dz_pitch(V) * 2;
  }
  // split-K scratch of the dH GEMM: the larger need of the two slab sizes in use (full slab, remainder)
  const int n_groups = (sm_count > kCG ? sm_count : kCG) / kCG;
  w.dh_partial_bytes = dh_partial_bytes(slab, d, V, n_groups, kCG);
  if (M % slab != 0) {
    const size_t r = dh_partial_bytes(M % slab, d, V, n_groups, kCG);
    if (r > w.dh_partial_bytes) w.dh_partial_bytes = r;
  }
  w.dh_partials = w.dh_partial_bytes != 0 ? reinterpret_cast<float*>(p + off) : nullptr;
  off += round_up(w.dh_partial_bytes, 256);
  w.bytes = off;
  return w;
}

// The side stream the conversion passes run on: one per device, created on first use and kept.
int fused_side_stream(cudaStream_t* out) {
  static cudaStream_t streams[64] = {nullptr};
  static std::mutex mu;
  int dev = 0;
  FTK_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return ftk::fail(FTK_EINVAL, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lk(mu);
  if (streams[dev] == nullptr) FTK_CUDA_CHECK(cudaStreamCreateWithFlags(&streams[dev], cudaStreamNonBlocking));
  *out = streams[dev];
  return FTK_OK;
}

int fused_slab_rows(int64_t M, int d, int V, int64_t slab_rows, int64_t* out) {
  ftk::DeviceInfo di;
  if (int rc = ftk::get_device_info(&di)) return rc;
  int64_t slab = slab_rows > 0 ? round_up(slab_rows, 256) : plan_fused_slab(M, d, V, di.sm_count, cta_group());
  if (slab > round_up(M, 256)) slab = round_up(M, 256);
  *out = slab;
  return FTK_OK;
}

}  // namespace

extern "C" int64_t ftk_lmhead_ce_fused_slab_rows(int64_t M, int32_t d, int32_t V, int64_t slab_rows) {
  int64_t slab = 0;
  if (M <= 0 || d <= 0 || V <= 0 || fused_slab_rows(M, d, V, slab_rows, &slab)) return 0;
  return slab;
}

extern "C" int64_t ftk_lmhead_ce_fused_plan(int64_t M, int32_t d, int32_t V, int32_t sm_count, int32_t cta_group) {
  if (M <= 0 || d <= 0 || V <= 0 || sm_count <= 0 || (cta_group != 1 && cta_group != 2)) return 0;
  return plan_fused_slab(M, d, V, sm_count, cta_group);
}

extern "C" size_t ftk_lmhead_ce_fused_workspace(int64_t M, int32_t d, int32_t V, int64_t slab_rows) {
  const int64_t slab = ftk_lmhead_ce_fused_slab_rows(M, d, V, slab_rows);
  if (slab <= 0) return 0;
  ftk::DeviceInfo di;
  if (ftk::get_device_info(&di)) return 0;
  // the base only matters for the pointers, not for the size
  return carve_fused(reinterpret_cast<void*>(uintptr_t(256)), M, d, V, slab, di.sm_count, cta_group()).bytes;
}

extern "C" int ftk_lmhead_ce_fused(const void* H, const void* W, const void* bias, const int64_t* targets, int64_t M,
                                   int32_t d, int32_t V, float grad_scale, const float* grad_scale_dev, float* lse,
                                   float* loss_sum, int32_t* n_valid, void* dH, void* dW, int32_t dw_dtype,
                                   int32_t accumulate, float* dbias, void* dw_done_event, void* workspace,
                                   size_t workspace_bytes, int64_t slab_rows, void* stream) {
  ftk::DeviceInfo di;
  if (int rc = check_common(H, W, M, d, V, &di)) return rc;
  FTK_REQUIRE(targets && lse && loss_sum && n_valid && dH && dW && workspace,
              "ftk_lmhead_ce_fused: null pointer argument");
  FTK_REQUIRE(dw_dtype == FTK_BF16 || dw_dtype == FTK_F32, "ftk_lmhead_ce_fused: bad dw_dtype %d", dw_dtype);
  FTK_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "ftk_lmhead_ce_fused: workspace must be 256-byte aligned");
  FTK_REQUIRE((reinterpret_cast<uintptr_t>(dH) & 15) == 0 && (reinterpret_cast<uintptr_t>(dW) & 15) == 0,
              "ftk_lmhead_ce_fused: dH/dW must be 16-byte aligned");
  int64_t slab = 0;
  if (int rc = fused_slab_rows(M, d, V, slab_rows, &slab)) return rc;
  const FusedWorkspace fw = carve_fused(workspace, M, d, V, slab, di.sm_count, cta_group());
  if (workspace_bytes < fw.bytes)
    return ftk::fail(FTK_EWORKSPACE, "ftk_lmhead_ce_fused: workspace %zu bytes < required %zu", workspace_bytes,
                     fw.bytes);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool pair = (cta_group() == 2 && di.sm_count >= 2);
  // the gradient scale 1 / n_valid is needed by the first slab, before the forward has seen every row: count first
  FTK_CUDA_CHECK(cudaMemsetAsync(n_valid, 0, sizeof(int32_t), st));
  FTK_CUDA_CHECK(cudaMemsetAsync(loss_sum, 0, sizeof(float), st));
  {
    const int64_t blocks = ceil_div(M, 2048);
    count_valid_kernel<<<static_cast<unsigned>(blocks < 128 ? blocks : 128), 256, 0, st>>>(targets, M, n_valid);
    FTK_CUDA_CHECK(cudaGetLastError());
  }
  if (dbias != nullptr && !accumulate) FTK_CUDA_CHECK(cudaMemsetAsync(dbias, 0, sizeof(float) * V, st));
  const int64_t ld = dz_pitch(V);
  const int n_tiles = ceil_div(V, 256);
  // Slabs: the remainder (M mod slab rows) first, full slabs after it, so the last slab -- the one whose dH GEMM a
  // data-parallel caller overlaps with the all-reduce of dW -- is a full one.
  // Schedule, with two slab buffers: sweep(0), sweep(1), dH(0), dW(0), sweep(2), dH(1), dW(1), ... on `stream`, and
  // convert(i) on the side stream between sweep(i) and dH(i), i.e. beside sweep(i+1) (or beside the gradient GEMMs of
  // slab i-1 for the last one).  The conversion is HBM-bound and shared-memory free, the sweeps are tensor-bound and
  // hold the shared memory: they share the SMs.  With one slab everything is in stream order.
  const int64_t rem = M % fw.slab;
  const int n_slabs = static_cast<int>(M / fw.slab + (rem != 0));
  cudaStream_t side = st;
  cudaEvent_t ev_swept[2] = {nullptr, nullptr}, ev_conv[2] = {nullptr, nullptr};
  const bool overlap = fw.n_buf == 2 && n_slabs >= 2;
  auto cleanup = [&]() {
    for (int b = 0; b < 2; ++b) {
      if (ev_swept[b] != nullptr) cudaEventDestroy(ev_swept[b]);  // pending work keeps what it needs
      if (ev_conv[b] != nullptr) cudaEventDestroy(ev_conv[b]);
    }
  };
  if (overlap) {
    if (int rc = fused_side_stream(&side)) return rc;
    for (int b = 0; b < 2; ++b) {
      cudaError_t e = cudaEventCreateWithFlags(&ev_swept[b], cudaEventDisableTiming);
      if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ev_conv[b], cudaEventDisableTiming);
      if (e != cudaSuccess) {
        cleanup();
        return ftk::fail(FTK_ECUDA, "ftk_lmhead_ce_fused: cudaEventCreate failed: %s", cudaGetErrorString(e));
      }
    }
  }
  struct Slab {
    int64_t r0, rows;
  };
  auto slab_at = [&](int i) {
    Slab sl;
    if (rem != 0) {
      sl.r0 = (i == 0) ? 0 : rem + static_cast<int64_t>(i - 1) * fw.slab;
      sl.rows = (i == 0) ? rem : fw.slab;
    } else {
      sl.r0 = static_cast<int64_t>(i) * fw.slab;
      sl.rows = fw.slab;
    }
    return sl;
  };
  // (1) logits tile by tile: softmax statistics + bf16 numerators of slab i; (2) lse and the slab's loss;
  // (3) numerators -> dZ in place, on the side stream
  auto sweep_and_convert = [&](int i) -> int {
    const Slab sl = slab_at(i);
    const int b = i & (fw.n_buf - 1);
    const void* Hs = static_cast<const uint8_t*>(H) + static_cast<size_t>(sl.r0) * d * 2;
    const FwdWorkspace ws = carve_fwd(fw.fwd, sl.rows);
    int rc = pair ? run_fwd_keep<2>(Hs, W, bias, targets + sl.r0, sl.rows, d, V, lse + sl.r0, loss_sum, nullptr,
                                    fw.probs[b], fw.cref[b], ws, di.sm_count, st, 1, 0)
                  : run_fwd_keep<1>(Hs, W, bias, targets + sl.r0, sl.rows, d, V, lse + sl.r0, loss_sum, nullptr,
                                    fw.probs[b], fw.cref[b], ws, di.sm_count, st, 1, 0);
    if (rc) return rc;
    if (overlap) {
      FTK_CUDA_CHECK(cudaEventRecord(ev_swept[b], st));
      FTK_CUDA_CHECK(cudaStreamWaitEvent(side, ev_swept[b], 0));
    }
    {
      ftk::TimelineScope tl_(FTK_TL_CONVERT, side);
      probs_to_dz_kernel<<<static_cast<unsigned>(sl.rows), 256, 0, side>>>(
          fw.probs[b], fw.cref[b], lse + sl.r0, targets + sl.r0, static_cast<int>(sl.rows), V, static_cast<int>(ld),
          static_cast<int>(round_up(sl.rows, 256)), n_tiles, grad_scale, grad_scale_dev, n_valid);
      FTK_CUDA_CHECK(cudaGetLastError());
    }
    if (overlap) FTK_CUDA_CHECK(cudaEventRecord(ev_conv[b], side));
    return FTK_OK;
  };
  // (4) dH of slab i's rows, dW accumulated over the slabs
  auto grad_gemms = [&](int i) -> int {
    const Slab sl = slab_at(i);
    const int b = i & (fw.n_buf - 1);
    if (overlap) FTK_CUDA_CHECK(cudaStreamWaitEvent(st, ev_conv[b], 0));
    const void* Hs = static_cast<const uint8_t*>(H) + static_cast<size_t>(sl.r0) * d * 2;
    void* dHs = static_cast<uint8_t*>(dH) + static_cast<size_t>(sl.r0) * d * 2;
    const int acc = (accumulate || i > 0) ? 1 : 0;
    GradGemmPlan plan;
    plan.overlap_tail = (i == n_slabs - 1) && dw_done_event != nullptr;
    plan.full_sms = !plan.overlap_tail;  // the SM reserve only matters where a collective can run beside us
    plan.dw_done = static_cast<cudaEvent_t>(dw_done_event);
    plan.dh_partials = fw.dh_partials;
    plan.dh_partial_bytes = fw.dh_partial_bytes;
    return pair ? run_grad_gemms<2>(Hs, W, sl.rows, d, V, dHs, dW, dw_dtype, acc, dbias, fw.probs[b], di.sm_count, st,
                                    plan)
                : run_grad_gemms<1>(Hs, W, sl.rows, d, V, dHs, dW, dw_dtype, acc, dbias, fw.probs[b], di.sm_count, st,
                                    plan);
  };
  int rc = sweep_and_convert(0);
  for (int i = 0; i < n_slabs && rc == FTK_OK; ++i) {
    if (overlap && i + 1 < n_slabs) rc = sweep_and_convert(i + 1);  // fills the other buffer while convert(i) runs
    if (rc == FTK_OK) rc = grad_gemms(i);
    if (!overlap && rc == FTK_OK && i + 1 < n_slabs) rc = sweep_and_convert(i + 1);
  }
  cleanup();
  return rc;
}

extern "C" int ftk_scale_unless_one(void* buf, size_t n, int32_t dtype, const float* factor_dev, int32_t poison,
                                    void* stream) {
  FTK_REQUIRE(buf && factor_dev, "ftk_scale_unless_one: null pointer argument");
  FTK_REQUIRE(dtype == FTK_BF16 || dtype == FTK_F32, "ftk_scale_unless_one: bad dtype %d", dtype);
  const size_t per_vec = dtype == FTK_F32 ? 4 : 8;
  FTK_REQUIRE(n % per_vec == 0 && (reinterpret_cast<uintptr_t>(buf) & 15) == 0,
              "ftk_scale_unless_one: buffer must be 16-byte aligned and hold a multiple of %zu elements", per_vec);
  if (n == 0) return FTK_OK;
  ftk::DeviceInfo di;
  if (int rc = ftk::get_device_info(&di)) return rc;
  const size_t n_vec = n / per_vec;
  size_t blocks = (n_vec + 255) / 256;
  if (blocks > static_cast<size_t>(di.sm_count) * 8) blocks = static_cast<size_t>(di.sm_count) * 8;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == FTK_F32)
    scale_unless_one_kernel<float><<<static_cast<unsigned>(blocks), 256, 0, st>>>(static_cast<float*>(buf), n_vec, factor_dev,
                                                                                  poison);
  else
    scale_unless_one_kernel<__nv_bfloat16><<<static_cast<unsigned>(blocks), 256, 0, st>>>(
        static_cast<__nv_bfloat16*>(buf), n_vec, factor_dev, poison);
  FTK_CUDA_CHECK(cudaGetLastError());
  return FTK_OK;
}

// ------------------------------------------------------------------------------------------------
// Linear layers of the transformer trunk on the same kernel template (include/ftk.h: ftk_linear_fwd / ftk_linear_bwd).
namespace {

int check_linear(const void* X, const void* W, int64_t M, int K, int N, ftk::DeviceInfo* di) {
  FTK_REQUIRE(X && W, "ftk_linear: null X or W");
  FTK_REQUIRE(M > 0 && M < (int64_t(1) << 30), "ftk_linear: M=%lld out of range", static_cast<long long>(M));
  FTK_REQUIRE(K > 0 && K % 64 == 0 && N > 0 && N % 64 == 0, "ftk_linear: K=%d and N=%d must be positive multiples of 64",
              K, N);
  if (int rc = ftk::get_device_info(di)) return rc;
  if (di->cc_major != 10)
    return ftk::fail(FTK_EUNSUPPORTED, "ftk_linear needs an sm_100 device (tcgen05/TMEM); found sm_%d%d", di->cc_major,
                     di->cc_minor);
  return FTK_OK;
}

// out[M, N] = f(A[M, Kd] . B^T), A K-major; B K-major ([N, Kd] rows) or MN-major ([Kd, N] rows).
template <int kCG, bool kBMN>
int run_linear_gemm(const void* A, const void* B, int64_t M, int Kd, int N, const LinearEpi::Params& ep_in, int sm_count,
                    cudaStream_t st, int tl_tag) {
  constexpr int BN = 256;
  constexpr int kStages = (kCG == 1) ? 4 : 6;
  CUtensorMap tmA, tmB;
  if (int rc = make_tmap(&tmA, A, Kd, M, static_cast<uint64_t>(Kd) * 2, BK, BM)) return rc;
  if constexpr (kBMN) {
    if (int rc = make_tmap(&tmB, B, N, Kd, static_cast<uint64_t>(N) * 2, 64, BK)) return rc;
  } else {
    if (int rc = make_tmap(&tmB, B, Kd, N, static_cast<uint64_t>(Kd) * 2, BK, BN / kCG)) return rc;
  }
  Sched sc;
  sc.k_split = 1;
  const int n_mblk = ceil_div(M, BM * kCG);
  sc.n_ntile = ceil_div(N, BN);
  const int usable = sm_count - sm_reserve();
  int n_chunks;
  plan_chunks(n_mblk, sc.n_ntile, (usable > kCG ? usable : kCG) / kCG, &n_chunks, &sc.tiles_per_chunk);
  sc.n_units = n_mblk * n_chunks;
  sc.n_mblk = n_mblk;
  sc.n_chunks = n_chunks;
  sc.m_group = pick_m_group(n_mblk, BM * kCG, Kd);
  sc.n_kblk = ceil_div(Kd, BK);
  pick_hints("nnn", 0, &sc.hint_a, &sc.hint_b, nullptr);
  ftk::TimelineScope tl_(tl_tag, st);
  return launch_umma<kCG, BN, kStages, false, kBMN, LinearEpi>(tmA, tmB, sc, ep_in, sm_count, st);
}

}  // namespace

extern "C" int ftk_linear_fwd(const void* X, const void* W, const void* bias, const void* residual, int64_t M, int32_t K,
                              int32_t N, int32_t act, void* Y, void* pre, void* stream) {
  ftk::DeviceInfo di;
  if (int rc = check_linear(X, W, M, K, N, &di)) return rc;
  FTK_REQUIRE(Y != nullptr, "ftk_linear_fwd: null output");
  FTK_REQUIRE(act == FTK_ACT_NONE || act == FTK_ACT_GELU_NEW, "ftk_linear_fwd: bad activation code %d", act);
  FTK_REQUIRE((reinterpret_cast<uintptr_t>(bias) & 15) == 0, "ftk_linear_fwd: bias must be 16-byte aligned");
  FTK_REQUIRE(((reinterpret_cast<uintptr_t>(Y) | reinterpret_cast<uintptr_t>(pre) | reinterpret_cast<uintptr_t>(residual)) & 31) == 0,
              "ftk_linear_fwd: Y / pre / residual must be 32-byte aligned");
  LinearEpi::Params ep;
  ep.out = static_cast<__nv_bfloat16*>(Y);
  ep.pre = static_cast<__nv_bfloat16*>(pre);
  ep.bias = static_cast<const __nv_bfloat16*>(bias);
  ep.mul_pre = nullptr;
  ep.residual = static_cast<const __nv_bfloat16*>(residual);
  ep.rows = static_cast<int>(M);
  ep.cols = N;
  ep.ld = N;
  ep.act = act;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (cta_group() == 2 && di.sm_count >= 2)
    return run_linear_gemm<2, false>(X, W, M, K, N, ep, di.sm_count, st, FTK_TL_LINEAR_FWD);
  return run_linear_gemm<1, false>(X, W, M, K, N, ep, di.sm_count, st, FTK_TL_LINEAR_FWD);
}

extern "C" int ftk_linear_bwd(const void* X, const void* W, const void* dY, const void* dx_mul_pre, const void* dx_add,
                              int64_t M, int32_t K, int32_t N, void* dX, void* dW, int32_t dw_dtype, int32_t accumulate,
                              float* dbias, void* stream) {
  ftk::DeviceInfo di;
  if (int rc = check_linear(X, W, M, K, N, &di)) return rc;
  FTK_REQUIRE(dY != nullptr, "ftk_linear_bwd: null dY");
  FTK_REQUIRE(dw_dtype == FTK_BF16 || dw_dtype == FTK_F32, "ftk_linear_bwd: bad dw_dtype %d", dw_dtype);
  FTK_REQUIRE(((reinterpret_cast<uintptr_t>(dX) | reinterpret_cast<uintptr_t>(dW) | reinterpret_cast<uintptr_t>(dx_mul_pre) |
                reinterpret_cast<uintptr_t>(dx_add)) & 31) == 0,
              "ftk_linear_bwd: dX / dW / dx_mul_pre / dx_add must be 32-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool pair = cta_group() == 2 && di.sm_count >= 2;
  if (dX != nullptr) {  // dX[M, K] = dY[M, N] . W[N, K]  (times gelu_new'(dx_mul_pre) when given)
    LinearEpi::Params ep;
    ep.out = static_cast<__nv_bfloat16*>(dX);
    ep.pre = nullptr;
    ep.bias = nullptr;
    ep.mul_pre = static_cast<const __nv_bfloat16*>(dx_mul_pre);
    ep.residual = static_cast<const __nv_bfloat16*>(dx_add);
    ep.rows = static_cast<int>(M);
    ep.cols = K;
    ep.ld = K;
    ep.act = FTK_ACT_NONE;
    const int rc = pair ? run_linear_gemm<2, true>(dY, W, M, N, K, ep, di.sm_count, st, FTK_TL_LINEAR_DX)
                        : run_linear_gemm<1, true>(dY, W, M, N, K, ep, di.sm_count, st, FTK_TL_LINEAR_DX);
    if (rc) return rc;
  }
  if (dW != nullptr) {  // dW[N, K] (+)= dY^T[N, M] . X[M, K]: the lm_head's dW GEMM with V = N, d = K
    __nv_bfloat16* dy = const_cast<__nv_bfloat16*>(static_cast<const __nv_bfloat16*>(dY));
    // run_dw reads its A operand through a [rows][round_up(V, 64)] pitch: N is a multiple of 64, so the pitch is N
    const int rc = pair ? run_dw<2>(X, M, K, N, dW, dw_dtype, accumulate, dy, di.sm_count, st, -1)
                        : run_dw<1>(X, M, K, N, dW, dw_dtype, accumulate, dy, di.sm_count, st, -1);
    if (rc) return rc;
  }
  if (dbias != nullptr) {  // dbias[N] (+)= column sums of dY
    if (!accumulate) FTK_CUDA_CHECK(cudaMemsetAsync(dbias, 0, sizeof(float) * N, st));
    ftk::TimelineScope tl_(FTK_TL_DBIAS, st);
    const int rows_per_block = 256;
    dim3 grid(ceil_div(N, 256), ceil_div(M, rows_per_block));
    dbias_kernel<<<grid, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(dY), static_cast<int>(M), N, N, rows_per_block,
                                       dbias);
    FTK_CUDA_CHECK(cudaGetLastError());
  }
  return FTK_OK;
}

// ------------------------------------------------------------------------------------------------
// Forward pass for callers that cannot allocate device memory.
//
// ftk_lmhead_ce_fwd needs a caller-provided workspace, which in practice means a cudaMalloc (or a framework
// allocator call) somewhere up the stack.  Some hosts have no way to make one: a plugin embedded in a serving or
// training runtime that claimed all of HBM for its own pool at start-up and hands its plugins no allocator, or a
// latency-critical loop that must not hit the device-wide synchronisation cudaMalloc/cudaFree imply.  For them the
// scratch below is part of the library image: a fixed array in device global memory, created when the module is
// loaded on a GPU and held until the process exits (every GPU that loads the library pays for it once).
//
// Limit: M <= FTK_STATIC_MAX_ROWS (16,384 rows, one 8 x 2048 micro-batch, 8.5 MB of scratch); larger calls are
// rejected.
// The array is shared by every call and keeps whatever the last call left in it.  The only part that must start
// at zero -- the finalize kernel's arrival counter -- is cleared on the stream at the start of each call; the
// per-chunk statistics are fully rewritten before they are read.  Two calls whose kernels overlap on the GPU
// (different streams, or different host threads) would write the same scratch and race: serialise them.
namespace {

constexpr int64_t kStaticMaxRows = FTK_STATIC_MAX_ROWS;
// same carve-up as carve_fwd(): 2 * kMaxChunks + 1 floats per (padded) row, block partials, counter
constexpr size_t kStaticScratchBytes =
    sizeof(float) * (2 * kMaxChunks + 1) * static_cast<size_t>(kStaticMaxRows) + 2 * 256 + 256;

static __device__ __align__(256) unsigned char g_fwd_static_scratch[kStaticScratchBytes];

// In host code `g_fwd_static_scratch` is only a handle for the runtime, not a device address; the address a kernel
// can use comes from cudaGetSymbolAddress, looked up once per device with that device current.
int static_scratch_ptr(void** out) {
  static void* ptr_of_device[64] = {nullptr};
  int dev = 0;
  FTK_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return ftk::fail(FTK_EINVAL, "device ordinal %d out of range", dev);
  if (ptr_of_device[dev] == nullptr) {
    void* p = nullptr;
    FTK_CUDA_CHECK(cudaGetSymbolAddress(&p, g_fwd_static_scratch));
    if (p == nullptr) return ftk::fail(FTK_ECUDA, "cudaGetSymbolAddress returned a null scratch address");
    ptr_of_device[dev] = p;
  }
  *out = ptr_of_device[dev];
  return FTK_OK;
}

}  // namespace

extern "C" int ftk_lmhead_ce_fwd_static(const void* H, const void* W, const void* bias, const int64_t* targets,
                                        int64_t M, int32_t d, int32_t V, float* lse, float* loss_sum,
                                        int32_t* n_valid, void* stream) {
  ftk::DeviceInfo di;
  if (int rc = check_common(H, W, M, d, V, &di)) return rc;
  FTK_REQUIRE(targets && lse && loss_sum && n_valid, "ftk_lmhead_ce_fwd_static: null pointer argument");
  if (M > kStaticMaxRows)
    return ftk::fail(FTK_EWORKSPACE,
                     "ftk_lmhead_ce_fwd_static: M=%lld exceeds the built-in scratch limit of %lld rows; use "
                     "ftk_lmhead_ce_fwd with a caller-provided workspace",
                     static_cast<long long>(M), static_cast<long long>(kStaticMaxRows));
  void* scratch = nullptr;
  if (int rc = static_scratch_ptr(&scratch)) return rc;
  const FwdWorkspace ws = carve_fwd(scratch, M);
  if (ws.bytes > kStaticScratchBytes)
    return ftk::fail(FTK_EWORKSPACE, "ftk_lmhead_ce_fwd_static: internal scratch too small (%zu > %zu)", ws.bytes,
                     kStaticScratchBytes);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (cta_group() == 2 && di.sm_count >= 2)
    return run_fwd<2>(H, W, bias, targets, M, d, V, lse, loss_sum, n_valid, ws, di.sm_count, st);
  return run_fwd<1>(H, W, bias, targets, M, d, V, lse, loss_sum, n_valid, ws, di.sm_count, st);
}
