// G1: context-window gather from the HBM-resident uint16 token stream.
//
// One launch builds input_ids (int64), labels (int64, -100 where masked; optionally pre-shifted for the
// fused loss) and attention_mask (bool) for B windows chosen by an index list. Replaces
// TokenizedDataset.load (FT.py:662-692), collector (FT.py:1030-1035) and the masked_fill of
// ModifiedTrainer.compute_loss (FT.py:476-477).
//
// HBM-bound byte work: 2 B read + 17 B written per token. Each block stages a 2048-token piece of one
// window in shared memory with 16-byte loads, then writes the three outputs with 16-byte stores in which
// consecutive threads cover consecutive addresses (a full 512 B per warp instruction).
#include <stdlib.h>

#include "ftk_host.h"

namespace {

constexpr int kThreads = 256;
constexpr int kTile = 2048;  // tokens staged per block iteration

struct GatherArgs {
  const uint16_t* tokens;
  const int64_t* ctx_idx;
  int64_t n_ctx;
  int T;
  int B;
  int pad_id;
  int pad_is_eos;
  int shift;
  int tiles_per_window;
  int64_t* ids;
  int64_t* labels;
  uint8_t* mask;
  int* n_valid;
};

template <bool kCS, typename T>
__device__ __forceinline__ void store_out(T* p, const T& v) {
  if constexpr (kCS)
    __stcs(p, v);  // streaming (evict-first) store: the outputs are not re-read by this kernel
  else
    *p = v;
}

template <bool kVec, bool kCS = false>
__global__ void __launch_bounds__(kThreads) gather_windows_kernel(GatherArgs a) {
  __shared__ __align__(16) uint16_t tok[kTile + 8];
  __shared__ int warp_cnt[kThreads / 32];
  const int tid = threadIdx.x;
  const int64_t n_tiles = static_cast<int64_t>(a.B) * a.tiles_per_window;

  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int b = static_cast<int>(tile / a.tiles_per_window);
    const int t0 = static_cast<int>(tile % a.tiles_per_window) * kTile;
    const int len = min(kTile, a.T - t0);  // tokens of this piece
    const int64_t idx = a.ctx_idx[b];
    const uint16_t* src = a.tokens + idx * a.T + t0;
    const bool all_ones = a.pad_is_eos && (idx != a.n_ctx - 1);  // FT.py:674
    const int pad = a.pad_id;

    // ---- stage: coalesced loads into shared memory (+1 halo token for the shifted labels)
    if constexpr (kVec) {
      const int nvec = len >> 3;  // T % 16 == 0 -> len % 16 == 0
      if (tid < nvec) reinterpret_cast<uint4*>(tok)[tid] = __ldg(reinterpret_cast<const uint4*>(src) + tid);
    } else {
      for (int i = tid; i < len; i += kThreads) tok[i] = __ldg(src + i);
    }
    if (tid == 0) tok[len] = (t0 + len < a.T) ? __ldg(src + len) : static_cast<uint16_t>(0);
    __syncthreads();

    const bool halo_valid = (t0 + len < a.T);  // is there a token after this piece inside the window?
    int64_t* ids_row = a.ids + static_cast<int64_t>(b) * a.T + t0;
    int64_t* lab_row = a.labels + static_cast<int64_t>(b) * a.T + t0;
    uint8_t* msk_row = a.mask + static_cast<int64_t>(b) * a.T + t0;
    int cnt = 0;

    if constexpr (kVec) {
      // ids / labels: 2 tokens (16 B) per thread per step
      for (int i = tid * 2; i < len; i += kThreads * 2) {
        const uint32_t pr = *reinterpret_cast<const uint32_t*>(tok + i);
        const int v0 = pr & 0xFFFF, v1 = pr >> 16;
        longlong2 o;
        o.x = v0;
        o.y = v1;
        store_out<kCS>(reinterpret_cast<longlong2*>(ids_row + i), o);
        longlong2 l;
        if (a.shift) {
          const int v2 = tok[i + 2];
          const bool ok1 = (i + 1 < len) || halo_valid;  // i+1 <= len always here
          const bool ok2 = (i + 2 < len) || (i + 2 == len && halo_valid);
          l.x = (ok1 && (all_ones || v1 != pad)) ? v1 : -100;
          l.y = (ok2 && (all_ones || v2 != pad)) ? v2 : -100;
        } else {
          l.x = (all_ones || v0 != pad) ? v0 : -100;
          l.y = (all_ones || v1 != pad) ? v1 : -100;
        }
        cnt += (l.x != -100) + (l.y != -100);
        store_out<kCS>(reinterpret_cast<longlong2*>(lab_row + i), l);
      }
      // mask: 16 tokens (16 B) per thread per step
      for (int i = tid * 16; i < len; i += kThreads * 16) {
        const uint4 p0 = *reinterpret_cast<const uint4*>(tok + i);
        const uint4 p1 = *reinterpret_cast<const uint4*>(tok + i + 8);
        const uint32_t w[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
        uint32_t m[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint32_t a0 = w[2 * q], a1 = w[2 * q + 1];
          const uint32_t b0 = (all_ones || static_cast<int>(a0 & 0xFFFF) != pad) ? 1u : 0u;
          const uint32_t b1 = (all_ones || static_cast<int>(a0 >> 16) != pad) ? 1u : 0u;
          const uint32_t b2 = (all_ones || static_cast<int>(a1 & 0xFFFF) != pad) ? 1u : 0u;
          const uint32_t b3 = (all_ones || static_cast<int>(a1 >> 16) != pad) ? 1u : 0u;
          m[q] = b0 | (b1 << 8) | (b2 << 16) | (b3 << 24);
        }
        store_out<kCS>(reinterpret_cast<uint4*>(msk_row + i), make_uint4(m[0], m[1], m[2], m[3]));
      }
    } else {
      for (int i = tid; i < len; i += kThreads) {
        const int v0 = tok[i];
        ids_row[i] = v0;
        int64_t l;
        if (a.shift) {
          const int v1 = tok[i + 1];
          const bool ok = (i + 1 < len) || halo_valid;
          l = (ok && (all_ones || v1 != pad)) ? v1 : -100;
        } else {
          l = (all_ones || v0 != pad) ? v0 : -100;
        }
        cnt += (l != -100);
        lab_row[i] = l;
        msk_row[i] = (all_ones || v0 != pad) ? 1 : 0;
      }
    }

    if (a.n_valid != nullptr) {  // block-reduce the count, one atomic per piece
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
      if ((tid & 31) == 0) warp_cnt[tid >> 5] = cnt;
      __syncthreads();
      if (tid == 0) {
        int s = 0;
#pragma unroll
        for (int w = 0; w < kThreads / 32; ++w) s += warp_cnt[w];
        atomicAdd(a.n_valid + b, s);
      }
    }
    __syncthreads();  // tok[] and warp_cnt[] are reused by the next piece
  }
}


// Variant without shared memory: one warp per 512-token piece, each lane owning 4 consecutive tokens per 128-token
// row of the piece. Loads are 8 B per lane (256 B per warp instruction); ids and labels leave as 32-byte stores
// (STG.256: 1 KB contiguous per warp instruction), the mask as 4-byte stores. All loads of a piece are issued before
// its stores. Needs T % 4 == 0.
constexpr int kPiece = 512;

template <bool kCS>
__device__ __forceinline__ void st_global_256(int64_t* p, int64_t a, int64_t b, int64_t c, int64_t d) {
  if constexpr (kCS)
    asm volatile("st.global.cs.v4.b64 [%0], {%1, %2, %3, %4};" ::"l"(p), "l"(a), "l"(b), "l"(c), "l"(d) : "memory");
  else
    asm volatile("st.global.v4.b64 [%0], {%1, %2, %3, %4};" ::"l"(p), "l"(a), "l"(b), "l"(c), "l"(d) : "memory");
}

template <bool kCS>
__global__ void __launch_bounds__(kThreads) gather_windows_direct_kernel(GatherArgs a) {
  const int lane = threadIdx.x & 31;
  const int warps_per_block = kThreads / 32;
  const int pieces_per_window = (a.T + kPiece - 1) / kPiece;
  const int64_t n_pieces = static_cast<int64_t>(a.B) * pieces_per_window;
  const int pad = a.pad_id;
  for (int64_t piece = static_cast<int64_t>(blockIdx.x) * warps_per_block + (threadIdx.x >> 5); piece < n_pieces;
       piece += static_cast<int64_t>(gridDim.x) * warps_per_block) {
    const int b = static_cast<int>(piece / pieces_per_window);
    const int t0 = static_cast<int>(piece % pieces_per_window) * kPiece;
    const int64_t idx = a.ctx_idx[b];
    const uint16_t* src = a.tokens + idx * a.T;
    const bool all_ones = a.pad_is_eos && (idx != a.n_ctx - 1);  // FT.py:674
    uint2 tk[4];
    uint32_t nxt[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {  // loads first
      const int t = t0 + k * 128 + lane * 4;
      tk[k] = make_uint2(0u, 0u);
      nxt[k] = 0u;
      if (t < a.T) {
        tk[k] = __ldg(reinterpret_cast<const uint2*>(src + t));
        if (a.shift && t + 4 < a.T) nxt[k] = __ldg(src + t + 4);
      }
    }
    int cnt = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int t = t0 + k * 128 + lane * 4;
      if (t >= a.T) continue;
      const int v0 = tk[k].x & 0xFFFF, v1 = tk[k].x >> 16, v2 = tk[k].y & 0xFFFF, v3 = tk[k].y >> 16;
      const int v4 = static_cast<int>(nxt[k]);
      const bool m0 = all_ones || v0 != pad, m1 = all_ones || v1 != pad, m2 = all_ones || v2 != pad,
                 m3 = all_ones || v3 != pad;
      const int64_t row = static_cast<int64_t>(b) * a.T + t;
      st_global_256<kCS>(a.ids + row, v0, v1, v2, v3);
      int64_t l0, l1, l2, l3;
      if (a.shift) {
        const bool m4 = (t + 4 < a.T) && (all_ones || v4 != pad);
        l0 = m1 ? v1 : -100;
        l1 = m2 ? v2 : -100;
        l2 = m3 ? v3 : -100;
        l3 = m4 ? v4 : -100;
      } else {
        l0 = m0 ? v0 : -100;
        l1 = m1 ? v1 : -100;
        l2 = m2 ? v2 : -100;
        l3 = m3 ? v3 : -100;
      }
      st_global_256<kCS>(a.labels + row, l0, l1, l2, l3);
      store_out<kCS>(reinterpret_cast<uint32_t*>(a.mask + row),
                     (m0 ? 1u : 0u) | (m1 ? 0x100u : 0u) | (m2 ? 0x10000u : 0u) | (m3 ? 0x1000000u : 0u));
      cnt += (l0 != -100) + (l1 != -100) + (l2 != -100) + (l3 != -100);
    }
    if (a.n_valid != nullptr) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
      if (lane == 0 && cnt != 0) atomicAdd(a.n_valid + b, cnt);
    }
  }
}

}  // namespace

extern "C" int ftk_gather_windows(const uint16_t* tokens, int64_t n_ctx, int32_t T, const int64_t* ctx_idx, int32_t B,
                                  int32_t pad_id, int32_t pad_is_eos, int32_t shift_labels, int64_t* input_ids,
                                  int64_t* labels, uint8_t* mask, int32_t* n_valid, void* stream) {
  FTK_REQUIRE(tokens && ctx_idx && input_ids && labels && mask, "ftk_gather_windows: null pointer argument");
  FTK_REQUIRE(n_ctx > 0 && T > 0 && B >= 0, "ftk_gather_windows: bad sizes n_ctx=%lld T=%d B=%d",
              static_cast<long long>(n_ctx), T, B);
  if (B == 0) return FTK_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ftk::DeviceInfo di;
  if (int rc = ftk::get_device_info(&di)) return rc;

  GatherArgs a;
  a.tokens = tokens;
  a.ctx_idx = ctx_idx;
  a.n_ctx = n_ctx;
  a.T = T;
  a.B = B;
  a.pad_id = pad_id;
  a.pad_is_eos = pad_is_eos ? 1 : 0;
  a.shift = shift_labels ? 1 : 0;
  a.tiles_per_window = (T + kTile - 1) / kTile;
  a.ids = input_ids;
  a.labels = labels;
  a.mask = mask;
  a.n_valid = n_valid;

  if (n_valid != nullptr) FTK_CUDA_CHECK(cudaMemsetAsync(n_valid, 0, sizeof(int32_t) * B, st));

  const int64_t max_blocks = static_cast<int64_t>(di.sm_count) * 8;  // 8 resident 256-thread blocks per SM
  const bool aligned32 = ((reinterpret_cast<uintptr_t>(tokens) & 31) == 0) &&
                         ((reinterpret_cast<uintptr_t>(input_ids) & 31) == 0) &&
                         ((reinterpret_cast<uintptr_t>(labels) & 31) == 0) &&
                         ((reinterpret_cast<uintptr_t>(mask) & 31) == 0);
  // T % 16 == 0: the shared-memory staged kernel (measured 4-5 % faster on B200); T % 4 == 0: the direct kernel;
  // anything else: the scalar path. ftk_set_option(FTK_OPT_GATHER_VARIANT, 1) forces the direct kernel where it applies.
  // Outputs too large to stay in L2 (>= 4 M tokens, 80 MB) leave through streaming stores (st.global.cs), which
  // measured +11 % on a 2^26-token launch; small batches use plain stores so the consumers find them in L2.
  // FTK_OPT_GATHER_STORES overrides.
  // Large launches (outputs beyond L2) take the direct kernel by default: with the token source read from HBM
  // it measured 5.66-5.69 TB/s against 5.43-5.48 for the staged one (round 2); small batches keep the staged kernel.
  // FTK_OPT_GATHER_VARIANT: 0 = this rule, 1 = always direct, 2 = always staged.
  const int variant = ftk::option(FTK_OPT_GATHER_VARIANT);
  const bool big = static_cast<int64_t>(B) * T >= (int64_t(1) << 22);
  const bool want_direct = variant == 1 || (variant == 0 && big);
  const int stores = ftk::option(FTK_OPT_GATHER_STORES);
  bool cs = static_cast<int64_t>(B) * T >= (int64_t(1) << 22);
  if (stores != 0) cs = (stores == 1);
  ftk::TimelineScope tl_(FTK_TL_GATHER, st);
  if (T % 4 == 0 && aligned32 && (want_direct || T % 16 != 0)) {
    const int64_t n_pieces = static_cast<int64_t>(B) * ((T + kPiece - 1) / kPiece);
    const int64_t blocks = (n_pieces + kThreads / 32 - 1) / (kThreads / 32);
    const int grid = static_cast<int>(blocks < max_blocks ? blocks : max_blocks);
    if (cs)
      gather_windows_direct_kernel<true><<<grid, kThreads, 0, st>>>(a);
    else
      gather_windows_direct_kernel<false><<<grid, kThreads, 0, st>>>(a);
  } else {
    const int64_t n_tiles = static_cast<int64_t>(B) * a.tiles_per_window;
    const int grid = static_cast<int>(n_tiles < max_blocks ? n_tiles : max_blocks);
    const bool vec = (T % 16 == 0) && aligned32;
    if (vec && cs)
      gather_windows_kernel<true, true><<<grid, kThreads, 0, st>>>(a);
    else if (vec)
      gather_windows_kernel<true><<<grid, kThreads, 0, st>>>(a);
    else
      gather_windows_kernel<false><<<grid, kThreads, 0, st>>>(a);
  }
  FTK_CUDA_CHECK(cudaGetLastError());
  return FTK_OK;
}
