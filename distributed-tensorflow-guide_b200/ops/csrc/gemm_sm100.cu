// bf16 GEMM for sm_100a: tcgen05.mma (single-thread issue) with the accumulator in
// TMEM, operands staged by TMA into 128B-swizzled shared memory, persistent CTAs,
// warp-specialised (1 TMA warp, 1 MMA warp, 4 epilogue warps), double-buffered TMEM
// so the epilogue of tile i overlaps the main loop of tile i+1.
//
// Work list rows W1-W4 of SURVEY section 8:
//   W1 forward  Y = act(X W^T + b)                A,B K-major
//   W2 dgrad    dX = (dY W) * relu'/gelu'          A K-major, B MN-major
//   W3 wgrad + Hogwild push: dW = dY^T X, epilogue issues red.global.add.v4.f32 of
//      -lr*dW straight into the ps shard (a peer GPU's HBM over NVLink)   A,B MN-major
//   W4 wgrad -> window accumulator (+=) or scaled store into the symmetric grad
//      buffer the reduce-scatter kernel reads
//
// Replaces: K1/K2 matmuls, K3 ApplyGradientDescent-on-ps (fused), K8
// window accumulate (fused), K13 per-worker lr scale (alpha).
#include "common.cuh"

#include <algorithm>
#include <mutex>
#include <vector>
#include <unordered_map>
#include <string>
#include <cstring>
#include <cstdlib>

namespace dtg {

constexpr int BM = 128;
constexpr int BK = 64;       // 64 bf16 = 128 B = one swizzle atom
constexpr int UMMA_K = 16;
constexpr int GEMM_THREADS = 320;   // warps 0-3 and 6-9: epilogue; 4: TMA; 5: MMA
#ifndef DTG_EPI_ONE_LANE_POLL
#define DTG_EPI_ONE_LANE_POLL 0
#endif
// Epilogue staging: 8 warps x (32 rows x 32 floats), no padding.  Row r keeps its j-th 16-byte
// column at slot j ^ (r & 7): conflict-free both for the row-per-thread writes and for the
// 4-rows-x-128-B reads, and 4 KB smaller than a padded pitch -- which is what lets the 2-CTA
// kernel hold a sixth operand stage (profiles/gemm_2cta_8k_ncu_summary.md).
constexpr int EPI_PITCH = 32;
constexpr int EPI_FLOATS = 8 * 32 * EPI_PITCH;
__device__ __forceinline__ int epi_slot(int row, int col4) {      // float index of (row, 4*col4)
  return row * EPI_PITCH + ((col4 ^ (row & 7)) << 2);
}

// EPI_F32_RED_COAL: same reduction-add as EPI_F32_RED but staged through shared memory so
// that each instruction covers 4 rows x 128 contiguous bytes -- 128 B NVLink payloads for a
// target in peer HBM (measured 187 -> see profiles/ GB/s for the scattered 16 B form).
enum EpiMode : int { EPI_BF16 = 0, EPI_F32 = 1, EPI_F32_ACC = 2, EPI_F32_RED = 3, EPI_F32_RED_COAL = 4,
                     // attention epilogues of the batched kernel (scores never reach HBM in fp32):
                     EPI_ATTN_STATS = 5,   // per (row, n-tile) running max / sum-exp of scale*QK^T
                     EPI_ATTN_P = 6,       // P = exp2(scale2*acc - m[row]) * invl[row]  -> bf16
                     EPI_ATTN_DS = 7,      // dS = alpha * P * (acc - D[row])             -> bf16
                     // push as a TMA reduction: the warp stages its 32x32 fp32 chunk in shared
                     // memory (128B-swizzle layout) and ONE lane issues
                     // cp.reduce.async.bulk.tensor.2d ... .add: the add into the (peer) ps shard
                     // leaves the SM as full 128 B rows from the TMA unit, not from the LSU
                     EPI_F32_RED_TMA = 8 };
enum TmaOut : int { TMAOUT_NONE = 0, TMAOUT_RED = 1, TMAOUT_F32 = 2, TMAOUT_BF16 = 3,
                    // same staging, but the warp reads its chunk back and stores it itself with
                    // full-row 16-byte vectors: no wait for a bulk store to drain the staging
                    // buffer (what bounded the exposed epilogue of the tall kernel: 8 chunks x
                    // one TMA-store round trip per warp and tile)
                    COALOUT_F32 = 4, COALOUT_BF16 = 5 };
enum Act : int { ACT_NONE = 0, ACT_RELU = 1, ACT_GELU = 2 };
enum MaskMode : int { MASK_NONE = 0, MASK_RELU = 1, MASK_GELU = 2 };

struct GemmEpilogue {
  void* out;                    // bf16* (EPI_BF16) or float*
  void* out2;                   // optional bf16*: pre-activation copy (needed by gelu')
  const float* bias;            // [N] fp32, added before the activation
  const __nv_bfloat16* mask;    // [M, ldmask]: MASK_RELU: saved activation, MASK_GELU: saved pre-activation
  const __nv_bfloat16* residual;  // [M, ldo] added after activation (bf16 mode)
  long long ldo;
  long long ldmask;
  float alpha;
  int mode;
  int act;
  int mask_mode;
  int vec;                      // 16B-vector stores allowed (alignment checked on host)
  int vec_mask;                 // 8B-vector mask loads allowed
  int fancy;                    // any of bias / act / mask / residual / out2 in use
  int direct;                   // bf16/fp32 store with 16B-aligned rows: register-direct epilogue
  int nb1, nb2;                 // batched form: nb1 * nb2 independent problems
  const float* rowv1;           // attention: m[row] (ATTN_P) or D[row] (ATTN_DS); row = bz*M + r
  const float* rowv2;           // attention: 1/l[row] (ATTN_P)
  float* stats;                 // ATTN_STATS output [nb*M][tiles_n][2]
  float scale2;                 // softmax scale * log2(e)
  int causal;                   // keys <= query only; tiles above the diagonal are skipped
  long long o_bs1, o_bs2;       // output batch strides (elements of the output type)
  int ksplit;                   // split-K: work unit = (tile, K range); partials meet by red.add
  unsigned long long l2hint;    // 2-CTA kernel: L2 eviction hint for operand loads (0 = none)
  int group;                    // 2-CTA kernel: m-blocks per tile-order group
  // tall (512 x 256 per pair) kernel schedule, filled by the host (see TallSched)
  int tall_big_mblocks;         // 512-row m-blocks done as big tiles
  int tall_nbig, tall_nsmall;   // tile counts
  int tall_extra;               // rounds of small tiles given first to the clusters with fewer big tiles
  // implicit-GEMM convolution (conv_which: 0 none; 1: the A operand is im2col(x), forward;
  // 2: the B operand is im2col(x), weight gradient).  x is NHWC bf16; the operand's tensor map
  // is an im2col-mode map over x and its k index / row index is a pixel or a (tap, channel) pair.
  int conv_which, cv_C, cv_k, cv_stride, cv_pad, cv_Ho, cv_Wo;
  unsigned int epi_sleep_ns;    // suspend-time hint of the epilogue warps' accumulator wait (0 = plain polling)
  int tma_single;               // one staging buffer per warp (64-wide tiles: one chunk per tile)
  int tma_out;                  // TmaOut: the chunk leaves through the TMA unit (0 = LSU stores)
  CUtensorMap tmO;              // [M, N] view of the output, box 32 x 32 (fp32: SW128, bf16: SW64)
  // shared-memory descriptor strides (bytes), filled on the host from the operand majors
  uint32_t a_lbo, a_sbo, a_kstep, b_lbo, b_sbo, b_kstep;
};

template <int BN>
struct GemmCfg {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_BYTES = 256;
  static constexpr int BUDGET = 227 * 1024 - 1024 /*alignment slack*/ - EPI_FLOATS * 4 - BAR_BYTES;
  static constexpr int STAGES_RAW = BUDGET / STAGE_BYTES;
  static constexpr int STAGES = STAGES_RAW > 8 ? 8 : STAGES_RAW;
  static constexpr int TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;   // 128 / 256 / 512
  static constexpr int SMEM_BYTES = 1024 + STAGES * STAGE_BYTES + EPI_FLOATS * 4 + BAR_BYTES;
};

__device__ __forceinline__ float gelu_tanh(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * (x + k1 * x * x * x);
  return 0.5f * x * (1.0f + tanhf(u));
}
__device__ __forceinline__ float gelu_tanh_grad(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * (x + k1 * x * x * x);
  float t = tanhf(u);
  return 0.5f * (1.0f + t) + 0.5f * x * (1.0f - t * t) * k0 * (1.0f + 3.0f * k1 * x * x);
}

// Grouped tile order: `group` consecutive m-blocks are swept across all n-blocks before
// moving on, so the tiles in flight at any time share a compact set of A and B panels
// (L2 reuse) instead of one B panel and every A panel.
__host__ __device__ __forceinline__ void tile_coords(int tile, int tiles_m, int tiles_n, int group,
                                                     int& mb, int& nb) {
  const int per_group = group * tiles_n;
  const int g = tile / per_group;
  const int first_m = g * group;
  const int gsz = group < tiles_m - first_m ? group : tiles_m - first_m;
  const int r = tile - g * per_group;
  mb = first_m + r % gsz;
  nb = r / gsz;
}

__device__ __forceinline__ void unpack_bf16x4(const __nv_bfloat16* p, float (&f)[4]) {
  const uint2 raw = *reinterpret_cast<const uint2*>(p);
  const __nv_bfloat162 lo = *reinterpret_cast<const __nv_bfloat162*>(&raw.x);
  const __nv_bfloat162 hi = *reinterpret_cast<const __nv_bfloat162*>(&raw.y);
  f[0] = __low2float(lo); f[1] = __high2float(lo);
  f[2] = __low2float(hi); f[3] = __high2float(hi);
}

// One 32x32 fp32 chunk of the accumulator: registers (thread = row) -> per-warp smem
// staging -> (4 rows x 128 B per instruction) coalesced global accesses with the fused
// epilogue.  Shared by the 1-CTA and the 2-CTA kernels.
// Global reads the epilogue of one chunk needs (ReLU'/GELU' mask, residual).  They do not
// depend on the accumulator, so the kernels issue them one chunk ahead -- for the first
// chunk of a tile even before the accumulator is complete -- and their latency hides
// behind the previous chunk's work instead of serialising the row iterations.
__device__ __forceinline__ bool chunk_is_direct(const GemmEpilogue& ep, int row_base, int cbase,
                                                int M, int N) {
  return ep.direct && (row_base + 31 < M) && (cbase + 31 < N);
}

__device__ __forceinline__ void epilogue_prefetch(const GemmEpilogue& ep, int lane, int row_base,
                                                  int cbase, int M, int N, uint2 (&mraw)[8],
                                                  uint2 (&rraw)[8]) {
  if (chunk_is_direct(ep, row_base, cbase, M, N)) {
    // direct layout: thread = row, its 32 columns = 64 contiguous bytes of mask / residual
    const long long row = row_base + lane;
    if (ep.mask_mode != MASK_NONE) {
      const uint4* mp = reinterpret_cast<const uint4*>(ep.mask + row * ep.ldmask + cbase);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint4 t = mp[i];
        mraw[2 * i] = make_uint2(t.x, t.y);
        mraw[2 * i + 1] = make_uint2(t.z, t.w);
      }
    }
    if (ep.mode == EPI_BF16 && ep.residual != nullptr) {
      const uint4* rp = reinterpret_cast<const uint4*>(ep.residual + row * ep.ldo + cbase);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint4 t = rp[i];
        rraw[2 * i] = make_uint2(t.x, t.y);
        rraw[2 * i + 1] = make_uint2(t.z, t.w);
      }
    }
    return;
  }
}

// Edge chunks (tile crosses M or N, or unaligned output): every access is guarded.
__device__ __noinline__ void epilogue_chunk_generic(const GemmEpilogue& ep, const float* st,
                                                    int lane, int row_base, int cbase, int M,
                                                    int N, long long ooff) {
  const int cg = lane & 7;       // 4-column group inside the 32-column chunk
  const int rsub = lane >> 3;    // 0..3
  const int col = cbase + cg * 4;
  float bz[4] = {0.f, 0.f, 0.f, 0.f};
  if (ep.bias != nullptr) {
#pragma unroll
    for (int e = 0; e < 4; ++e)
      if (col + e < N) bz[e] = __ldg(ep.bias + col + e);
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int rr = 4 * i + rsub;
    const int row = row_base + rr;
    float4 v4 = *reinterpret_cast<const float4*>(st + epi_slot(rr, cg));
    if (row >= M || col >= N) continue;
    float v[4] = {v4.x, v4.y, v4.z, v4.w};
    const bool full4 = (col + 3 < N) && ep.vec;
    float pre[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float x = ep.alpha * v[e] + bz[e];
      pre[e] = x;
      if (ep.act == ACT_RELU) x = fmaxf(x, 0.f);
      else if (ep.act == ACT_GELU) x = gelu_tanh(x);
      v[e] = x;
    }
    if (ep.mask_mode != MASK_NONE) {
      const __nv_bfloat16* mp = ep.mask + (long long)row * ep.ldmask + col;
      float mv[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if (col + e < N) mv[e] = __bfloat162float(mp[e]);
#pragma unroll
      for (int e = 0; e < 4; ++e)
        v[e] *= (ep.mask_mode == MASK_RELU) ? (mv[e] > 0.f ? 1.f : 0.f) : gelu_tanh_grad(mv[e]);
    }
    const long long off = ooff + (long long)row * ep.ldo + col;
    if (ep.mode == EPI_BF16) {
      __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(ep.out) + off;
      if (ep.residual != nullptr) {
        const __nv_bfloat16* rp = ep.residual + off;
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (col + e < N) v[e] += __bfloat162float(rp[e]);
      }
      if (full4) {
        *reinterpret_cast<uint2*>(o) =
            make_uint2(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]));
        if (ep.out2 != nullptr)
          *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(ep.out2) + off) =
              make_uint2(pack_bf16x2(pre[0], pre[1]), pack_bf16x2(pre[2], pre[3]));
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (col + e < N) {
            o[e] = __float2bfloat16(v[e]);
            if (ep.out2 != nullptr)
              reinterpret_cast<__nv_bfloat16*>(ep.out2)[off + e] = __float2bfloat16(pre[e]);
          }
      }
    } else {
      float* o = reinterpret_cast<float*>(ep.out) + off;
      if (ep.mode == EPI_F32) {
        if (full4) {
          *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (col + e < N) o[e] = v[e];
        }
      } else if (ep.mode == EPI_F32_ACC) {
        if (full4) {
          const float4 c = *reinterpret_cast<const float4*>(o);
          *reinterpret_cast<float4*>(o) =
              make_float4(c.x + v[0], c.y + v[1], c.z + v[2], c.w + v[3]);
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (col + e < N) o[e] += v[e];
        }
      } else {  // EPI_F32_RED: lock-free push, the target may be a peer GPU's HBM
        if (full4) {
          red_add_v4_f32(o, v[0], v[1], v[2], v[3]);
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (col + e < N) red_add_f32(o + e, v[e]);
        }
      }
    }
  }
}

// Direct path (bf16 / fp32 stores on interior, aligned chunks): no shared-memory staging at
// all.  A thread owns one accumulator row: 32 consecutive columns = 64 B (bf16) or 128 B
// (fp32) of output written with 16 B vector stores, in groups of 8 columns to keep the
// register footprint small.  Cheapest epilogue; used by every forward and dgrad GEMM.
//
// STAGED = true: the same values go to this warp's shared-memory staging instead (bf16: 32 rows
// x 64 B in the 64B-swizzle layout, fp32: 32 rows x 128 B in the 128B-swizzle layout) and the
// caller hands the box to the TMA unit -- full-line writes issued by one lane instead of 32
// partial-line STG wavefronts per instruction.
template <int MODE, bool FANCY, bool STAGED = false>
__device__ __forceinline__ void epilogue_chunk_direct(const GemmEpilogue& ep, int lane,
                                                      const uint32_t (&r)[32], int row_base,
                                                      int cbase, long long ooff,
                                                      const uint2 (&mraw)[8],
                                                      const uint2 (&rraw)[8],
                                                      float* stg = nullptr) {
  const long long off = ooff + (long long)(row_base + lane) * ep.ldo + cbase;
  const float alpha = ep.alpha;
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    float v[8], pre[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = alpha * __uint_as_float(r[8 * g + e]);
    if (FANCY) {
      if (ep.bias != nullptr) {
        const float4 b0 = __ldg(reinterpret_cast<const float4*>(ep.bias + cbase + 8 * g));
        const float4 b1 = __ldg(reinterpret_cast<const float4*>(ep.bias + cbase + 8 * g + 4));
        v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w;
        v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) pre[e] = v[e];
      if (ep.act == ACT_RELU) {
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = fmaxf(v[e], 0.f);
      } else if (ep.act == ACT_GELU) {
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = gelu_tanh(v[e]);
      }
      if (ep.mask_mode != MASK_NONE) {
        float mv[8];
        unpack_bf16x4(reinterpret_cast<const __nv_bfloat16*>(&mraw[2 * g]), *reinterpret_cast<float(*)[4]>(&mv[0]));
        unpack_bf16x4(reinterpret_cast<const __nv_bfloat16*>(&mraw[2 * g + 1]), *reinterpret_cast<float(*)[4]>(&mv[4]));
        if (ep.mask_mode == MASK_RELU) {
#pragma unroll
          for (int e = 0; e < 8; ++e) v[e] = mv[e] > 0.f ? v[e] : 0.f;
        } else {
#pragma unroll
          for (int e = 0; e < 8; ++e) v[e] *= gelu_tanh_grad(mv[e]);
        }
      }
    }
    if (MODE == EPI_BF16) {
      if (FANCY && ep.residual != nullptr) {
        float rv[8];
        unpack_bf16x4(reinterpret_cast<const __nv_bfloat16*>(&rraw[2 * g]), *reinterpret_cast<float(*)[4]>(&rv[0]));
        unpack_bf16x4(reinterpret_cast<const __nv_bfloat16*>(&rraw[2 * g + 1]), *reinterpret_cast<float(*)[4]>(&rv[4]));
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] += rv[e];
      }
      const uint4 pk = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]),
                                  pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
      if (STAGED)      // row = lane, 16-byte column g of 4, 64B swizzle
        *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(stg) + lane * 64 +
                                  ((g ^ ((lane >> 1) & 3)) << 4)) = pk;
      else
        *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(ep.out) + off + 8 * g) = pk;
      if (FANCY && ep.out2 != nullptr)
        *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(ep.out2) + off + 8 * g) =
            make_uint4(pack_bf16x2(pre[0], pre[1]), pack_bf16x2(pre[2], pre[3]),
                       pack_bf16x2(pre[4], pre[5]), pack_bf16x2(pre[6], pre[7]));
    } else {
      float* o = reinterpret_cast<float*>(ep.out) + off + 8 * g;
      if (STAGED) {    // EPI_F32 store or reduce-add: 128B swizzle, 16-byte columns 2g, 2g+1
        *reinterpret_cast<float4*>(stg + epi_slot(lane, 2 * g)) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(stg + epi_slot(lane, 2 * g + 1)) = make_float4(v[4], v[5], v[6], v[7]);
      } else if (MODE == EPI_F32) {
        *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
      } else if (MODE == EPI_F32_ACC) {
        const float4 c0 = *reinterpret_cast<const float4*>(o);
        const float4 c1 = *reinterpret_cast<const float4*>(o + 4);
        *reinterpret_cast<float4*>(o) = make_float4(c0.x + v[0], c0.y + v[1], c0.z + v[2], c0.w + v[3]);
        *reinterpret_cast<float4*>(o + 4) = make_float4(c1.x + v[4], c1.y + v[5], c1.z + v[6], c1.w + v[7]);
      } else {   // EPI_F32_RED: lock-free push; the target may be a peer GPU's HBM
        red_add_v4_f32(o, v[0], v[1], v[2], v[3]);
        red_add_v4_f32(o + 4, v[4], v[5], v[6], v[7]);
      }
    }
  }
}

// TMA reduce-add of one staged 32 x 32 fp32 box at (row_base, cbase); rows / columns past the
// tensor's bounds are clipped by the TMA unit, so edge chunks need no special case.
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* m, const void* smem_src,
                                                  int32_t c0, int32_t c1) {
  asm volatile(
      "cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.bulk_group [%0, {%2, %3}], [%1];\n" ::
          "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, const void* smem_src, int32_t c0,
                                             int32_t c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];\n" ::
          "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void bulk_wait_group_read1() {   // all but the newest group's sources
  asm volatile("cp.async.bulk.wait_group.read 1;\n" ::: "memory");
}
__device__ __forceinline__ void bulk_commit_group() {
  asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
}
__device__ __forceinline__ void bulk_wait_group_read0() {   // sources may be overwritten again
  asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory");
}
__device__ __forceinline__ void bulk_wait_group0() {        // everything this thread issued is done
  asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory");
}
// End of an epilogue warp's life: its outstanding bulk reductions read its shared memory.
__device__ __forceinline__ void epilogue_drain(const GemmEpilogue& ep, int lane) {
  if (ep.tma_out != 0 && ep.tma_out < COALOUT_F32 && lane == 0) bulk_wait_group0();
  __syncwarp();
}

__device__ __forceinline__ void epilogue_chunk(const GemmEpilogue& ep, float* st, int lane,
                                               const uint32_t (&r)[32], int row_base, int cbase,
                                               int M, int N, long long ooff,
                                               const uint2 (&mraw)[8], const uint2 (&rraw)[8]) {
  if (ep.tma_out >= COALOUT_F32 && chunk_is_direct(ep, row_base, cbase, M, N)) {
    // registers -> swizzled staging (thread = row) -> coalesced 16-byte stores (thread = column
    // group): every store instruction writes whole 64 B / 128 B row segments
    const bool bf = ep.tma_out == COALOUT_BF16;
    if (bf) {
      if (ep.fancy) epilogue_chunk_direct<EPI_BF16, true, true>(ep, lane, r, row_base, cbase, ooff, mraw, rraw, st);
      else          epilogue_chunk_direct<EPI_BF16, false, true>(ep, lane, r, row_base, cbase, ooff, mraw, rraw, st);
      __syncwarp();
      __nv_bfloat16* ob = reinterpret_cast<__nv_bfloat16*>(ep.out) + ooff + (long long)row_base * ep.ldo + cbase;
      const int j = lane & 3, rsub = lane >> 2;            // 4 x 16 B per 64 B row, 8 rows per instruction
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int rr = 8 * i + rsub;
        const uint4 v = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(st) + rr * 64 +
                                                        ((j ^ ((rr >> 1) & 3)) << 4));
        *reinterpret_cast<uint4*>(ob + (long long)rr * ep.ldo + 8 * j) = v;
      }
    } else {
      if (ep.fancy) epilogue_chunk_direct<EPI_F32, true, true>(ep, lane, r, row_base, cbase, ooff, mraw, rraw, st);
      else          epilogue_chunk_direct<EPI_F32, false, true>(ep, lane, r, row_base, cbase, ooff, mraw, rraw, st);
      __syncwarp();
      float* of = reinterpret_cast<float*>(ep.out) + ooff + (long long)row_base * ep.ldo + cbase;
      const int cg = lane & 7, rsub = lane >> 3;           // 8 x 16 B per 128 B row, 4 rows per instruction
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int rr = 4 * i + rsub;
        *reinterpret_cast<float4*>(of + (long long)rr * ep.ldo + 4 * cg) =
            *reinterpret_cast<const float4*>(st + epi_slot(rr, cg));
      }
    }
    __syncwarp();                        // the staging buffer is free again
    return;
  }
  if (ep.tma_out != 0 && ep.tma_out < COALOUT_F32 &&
      (ep.tma_out == TMAOUT_RED || chunk_is_direct(ep, row_base, cbase, M, N))) {
    // Output through the TMA unit.  Reductions take every chunk (the unit clips boxes at the
    // tensor's edge); stores take interior chunks only, because their bias / mask / residual
    // READS are not clipped.
    const bool bf = ep.tma_out == TMAOUT_BF16;
    // bf16 boxes are 2 KB: two per warp, alternating, so one store may still be reading
    const bool two_bufs = bf && !ep.tma_single;      // consecutive chunks of a warp alternate
    float* stg = two_bufs ? st + ((cbase >> 5) & 1) * 512 : st;
    if (lane == 0) { if (two_bufs) bulk_wait_group_read1(); else bulk_wait_group_read0(); }
    __syncwarp();
    if (ep.tma_out == TMAOUT_RED) {
      const float alpha = ep.alpha;
#pragma unroll
      for (int j = 0; j < 8; ++j)
        *reinterpret_cast<float4*>(stg + epi_slot(lane, j)) =
            make_float4(alpha * __uint_as_float(r[4 * j]), alpha * __uint_as_float(r[4 * j + 1]),
                        alpha * __uint_as_float(r[4 * j + 2]), alpha * __uint_as_float(r[4 * j + 3]));
    } else if (bf) {
      if (ep.fancy) epilogue_chunk_direct<EPI_BF16, true, true>(ep, lane, r, row_base, cbase, ooff, mraw, rraw, stg);
      else          epilogue_chunk_direct<EPI_BF16, false, true>(ep, lane, r, row_base, cbase, ooff, mraw, rraw, stg);
    } else {
      if (ep.fancy) epilogue_chunk_direct<EPI_F32, true, true>(ep, lane, r, row_base, cbase, ooff, mraw, rraw, stg);
      else          epilogue_chunk_direct<EPI_F32, false, true>(ep, lane, r, row_base, cbase, ooff, mraw, rraw, stg);
    }
    fence_proxy_async_smem();            // generic-proxy writes -> visible to the async proxy
    __syncwarp();
    if (lane == 0) {
      if (ep.tma_out == TMAOUT_RED) tma_reduce_add_2d(&ep.tmO, stg, cbase, row_base);
      else                          tma_store_2d(&ep.tmO, stg, cbase, row_base);
      bulk_commit_group();
    }
    return;
  }
  if (ep.mode == EPI_F32_RED_COAL && chunk_is_direct(ep, row_base, cbase, M, N)) {
    // coalesced remote push: registers -> smem (thread = row) -> 4 rows x 128 B per red
#pragma unroll
    for (int j = 0; j < 8; ++j)
      *reinterpret_cast<float4*>(st + epi_slot(lane, j)) =
          make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                      __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
    __syncwarp();
    const int cg = lane & 7, rsub = lane >> 3;
    const float alpha = ep.alpha;
    float* base = reinterpret_cast<float*>(ep.out) + ooff + (long long)row_base * ep.ldo + cbase + cg * 4;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int rr = 4 * i + rsub;
      const float4 a = *reinterpret_cast<const float4*>(st + epi_slot(rr, cg));
      red_add_v4_f32(base + (long long)rr * ep.ldo, alpha * a.x, alpha * a.y, alpha * a.z, alpha * a.w);
    }
    __syncwarp();
    return;
  }
  if (chunk_is_direct(ep, row_base, cbase, M, N)) {
    if (ep.fancy) {
      switch (ep.mode) {
        case EPI_BF16: epilogue_chunk_direct<EPI_BF16, true>(ep, lane, r, row_base, cbase, ooff, mraw, rraw); break;
        case EPI_F32: epilogue_chunk_direct<EPI_F32, true>(ep, lane, r, row_base, cbase, ooff, mraw, rraw); break;
        case EPI_F32_ACC: epilogue_chunk_direct<EPI_F32_ACC, true>(ep, lane, r, row_base, cbase, ooff, mraw, rraw); break;
        default: epilogue_chunk_direct<EPI_F32_RED, true>(ep, lane, r, row_base, cbase, ooff, mraw, rraw); break;
      }
    } else {
      switch (ep.mode) {
        case EPI_BF16: epilogue_chunk_direct<EPI_BF16, false>(ep, lane, r, row_base, cbase, ooff, mraw, rraw); break;
        case EPI_F32: epilogue_chunk_direct<EPI_F32, false>(ep, lane, r, row_base, cbase, ooff, mraw, rraw); break;
        case EPI_F32_ACC: epilogue_chunk_direct<EPI_F32_ACC, false>(ep, lane, r, row_base, cbase, ooff, mraw, rraw); break;
        default: epilogue_chunk_direct<EPI_F32_RED, false>(ep, lane, r, row_base, cbase, ooff, mraw, rraw); break;
      }
    }
    return;
  }
  // edge / unaligned chunk: stage through shared memory, every access guarded
  if (ep.tma_out != 0 && ep.tma_out < COALOUT_F32) {   // a TMA store may still be reading this staging buffer
    if (lane == 0) bulk_wait_group_read0();
    __syncwarp();
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    float4 v = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                           __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
    *reinterpret_cast<float4*>(st + epi_slot(lane, j)) = v;
  }
  __syncwarp();
  epilogue_chunk_generic(ep, st, lane, row_base, cbase, M, N, ooff);
  __syncwarp();
}

// ---- attention epilogues (batched kernel only) ---------------------------------------
template <int BN>
__device__ __forceinline__ bool causal_skip(const GemmEpilogue& ep, int tile, int tiles_per_batch,
                                            int tiles_m, int tiles_n) {
  if (!ep.causal) return false;
  const int bz = tile / tiles_per_batch;
  int mb, nb;
  tile_coords(tile - bz * tiles_per_batch, tiles_m, tiles_n, 16, mb, nb);
  return nb * BN >= mb * BM + BM;          // every key of the tile is after every query
}

// One thread owns one query row of the 128 x BN score tile in TMEM.
template <int BN>
__device__ __forceinline__ void attn_epilogue_tile(const GemmEpilogue& ep, uint32_t tmem_tile,
                                                   int row, int warp_row0, int M, int N, int n0,
                                                   int nb, int tiles_n, long long grow,
                                                   long long ooff) {
  const bool row_ok = row < M;
  const int climit = ep.causal ? row : N - 1;          // last visible key of this row
  if (ep.mode == EPI_ATTN_STATS) {
    float m = -INFINITY, l = 0.f;
#pragma unroll 1
    for (int chunk = 0; chunk < BN / 32; ++chunk) {
      const int cbase = n0 + chunk * 32;
      if (cbase >= N || (ep.causal && cbase > warp_row0 + 31)) break;   // warp-uniform
      uint32_t r[32];
      tmem_ld_32x32(tmem_tile + chunk * 32, r);
      tmem_ld_wait();
      float sv[32], cm = -INFINITY;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const int col = cbase + j;
        sv[j] = (col < N && col <= climit) ? __uint_as_float(r[j]) * ep.scale2 : -INFINITY;
        cm = fmaxf(cm, sv[j]);
      }
      const float mn = fmaxf(m, cm);
      if (mn > -INFINITY) {
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) acc += exp2f(sv[j] - mn);
        l = l * exp2f(m - mn) + acc;
        m = mn;
      }
    }
    if (row_ok) {
      float* sp = ep.stats + (grow * tiles_n + nb) * 2;
      sp[0] = m;
      sp[1] = l;
    }
    return;
  }
  const float rv1 = row_ok ? ep.rowv1[grow] : 0.f;
  const float rv2 = (row_ok && ep.mode == EPI_ATTN_P) ? ep.rowv2[grow] : 0.f;
  __nv_bfloat16* orow = reinterpret_cast<__nv_bfloat16*>(ep.out) + ooff + (long long)row * ep.ldo;
  const __nv_bfloat16* prow = ep.mask + ooff + (long long)row * ep.ldo;   // P, same layout as dS
#pragma unroll 1
  for (int chunk = 0; chunk < BN / 32; ++chunk) {
    const int cbase = n0 + chunk * 32;
    if (cbase >= N || (ep.causal && cbase > warp_row0 + 31)) break;
    uint32_t r[32];
    tmem_ld_32x32(tmem_tile + chunk * 32, r);
    tmem_ld_wait();
    if (!row_ok) continue;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const int col = cbase + 8 * g;
      if (col >= N) break;
      float v[8];
      if (ep.mode == EPI_ATTN_P) {
#pragma unroll
        for (int e = 0; e < 8; ++e)
          v[e] = (col + e <= climit && col + e < N)
                     ? exp2f(__uint_as_float(r[8 * g + e]) * ep.scale2 - rv1) * rv2 : 0.f;
      } else {
        float pv[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (col + 7 < N) {
          const uint4 raw = *reinterpret_cast<const uint4*>(prow + col);
          unpack_bf16x4(reinterpret_cast<const __nv_bfloat16*>(&raw.x), *reinterpret_cast<float(*)[4]>(&pv[0]));
          unpack_bf16x4(reinterpret_cast<const __nv_bfloat16*>(&raw.z), *reinterpret_cast<float(*)[4]>(&pv[4]));
        } else {
          for (int e = 0; e < 8 && col + e < N; ++e) pv[e] = __bfloat162float(prow[col + e]);
        }
#pragma unroll
        for (int e = 0; e < 8; ++e)
          v[e] = ep.alpha * pv[e] * (__uint_as_float(r[8 * g + e]) - rv1);
      }
      if (col + 7 < N) {
        *reinterpret_cast<uint4*>(orow + col) =
            make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                       pack_bf16x2(v[6], v[7]));
      } else {
        for (int e = 0; e < 8 && col + e < N; ++e) orow[col + e] = __float2bfloat16(v[e]);
      }
    }
  }
}

__device__ __forceinline__ void tma_load_4d(void* smem_dst, const CUtensorMap* m, uint64_t* bar,
                                            int32_t c0, int32_t c1, int32_t c2, int32_t c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];\n" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

// im2col-mode TMA load (SASS: UTMALDG.4D.IM2COL): `pixelsPerColumn` consecutive output pixels
// starting at the one whose filter window has its corner at input position (w, h) of image n,
// `channelsPerPixel` channels from c, filter tap (w_off, h_off); out-of-image taps read zero.
// The tile arrives exactly as the tiled load of the materialised patch matrix would leave it.
__device__ __forceinline__ void tma_load_im2col(void* smem_dst, const CUtensorMap* m, uint64_t* bar,
                                                int32_t c, int32_t w, int32_t h, int32_t n,
                                                uint16_t w_off, uint16_t h_off) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};\n" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c), "r"(w), "r"(h), "r"(n),
      "h"(w_off), "h"(h_off)
      : "memory");
}
__device__ __forceinline__ void tma_load_im2col_2sm(void* smem_dst, const CUtensorMap* m,
                                                    uint64_t* bar, int32_t c, int32_t w, int32_t h,
                                                    int32_t n, uint16_t w_off, uint16_t h_off) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.im2col.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};\n" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c), "r"(w), "r"(h),
      "r"(n), "h"(w_off), "h"(h_off)
      : "memory");
}
// Output pixel index -> (image, corner of its filter window in input coordinates).
struct ConvPixel { int n, h, w; };
__device__ __forceinline__ ConvPixel conv_pixel(const GemmEpilogue& ep, int m) {
  const int per = ep.cv_Ho * ep.cv_Wo;
  ConvPixel px;
  px.n = m / per;
  const int rem = m - px.n * per;
  const int p = rem / ep.cv_Wo;
  px.h = p * ep.cv_stride - ep.cv_pad;
  px.w = (rem - p * ep.cv_Wo) * ep.cv_stride - ep.cv_pad;
  return px;
}

// BATCHED: nb1 * nb2 independent GEMMs of the same shape in one launch (attention:
// one problem per (head, batch)); the tensor maps are then 4-D {inner, rows, b1, b2}.
template <int BN, bool A_MN, bool B_MN, bool BATCHED>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 int M, int N, int K, const __grid_constant__ GemmEpilogue ep) {
  using C = GemmCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = sA + C::STAGES * C::A_BYTES;
  float* sEpi = reinterpret_cast<float*>(sB + C::STAGES * C::B_BYTES);
  uint64_t* full = reinterpret_cast<uint64_t*>(sEpi + EPI_FLOATS);
  uint64_t* empty = full + C::STAGES;
  uint64_t* tfull = empty + C::STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int tiles_m = (M + BM - 1) / BM;
  const int tiles_n = (N + BN - 1) / BN;
  const int tiles_per_batch = tiles_m * tiles_n;
  const int num_tiles = BATCHED ? tiles_per_batch * ep.nb1 * ep.nb2 : tiles_per_batch;
  const int kblocks = (K + BK - 1) / BK;
  // split-K (wgrad with a small output and a huge reduction, e.g. conv filters): a unit is
  // (tile, K range); every unit adds its partial tile with red.global.add (host sets the mode)
  const int ksplit = (!BATCHED && ep.ksplit > 1) ? ep.ksplit : 1;
  const int kper = (kblocks + ksplit - 1) / ksplit;
  const int num_units = num_tiles * ksplit;

  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 5 && lane == 0) {
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], 8);
    }
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(tmem_slot, C::TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

  if (warp == 4) {
    // ===================== TMA producer (one lane) =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x) {
        const int tile = unit % num_tiles;
        const int kb0 = (unit / num_tiles) * kper, kb1 = min(kblocks, kb0 + kper);
        if (BATCHED && causal_skip<BN>(ep, tile, tiles_per_batch, tiles_m, tiles_n)) continue;
        int mb, nb;
        const int bz = BATCHED ? tile / tiles_per_batch : 0;
        tile_coords(BATCHED ? tile - bz * tiles_per_batch : tile, tiles_m, tiles_n, 16, mb, nb);
        const int b1 = BATCHED ? bz % ep.nb1 : 0, b2 = BATCHED ? bz / ep.nb1 : 0;
        const int m0 = mb * BM;
        const int n0 = nb * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full[stage], C::STAGE_BYTES);
          uint8_t* a_dst = sA + stage * C::A_BYTES;
          uint8_t* b_dst = sB + stage * C::B_BYTES;
          auto load = [&](void* dst, const CUtensorMap* map, int c0, int c1) {
            if constexpr (BATCHED) tma_load_4d(dst, map, &full[stage], c0, c1, b1, b2);
            else tma_load_2d(dst, map, &full[stage], c0, c1);
          };
          if (!BATCHED && !A_MN && ep.conv_which == 1) {
            // forward conv: rows = output pixels from m0, k-block = (filter tap, 64 channels)
            const int kcol = kb * BK, tap = kcol / ep.cv_C;
            const ConvPixel px = conv_pixel(ep, m0);
            tma_load_im2col(a_dst, &tmA, &full[stage], kcol - tap * ep.cv_C, px.w, px.h, px.n,
                            (uint16_t)(tap % ep.cv_k), (uint16_t)(tap / ep.cv_k));
          } else if constexpr (!A_MN) {
            load(a_dst, &tmA, kb * BK, m0);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j)
              load(a_dst + j * (64 * BK * 2), &tmA, m0 + j * 64, kb * BK);
          }
          if (!BATCHED && B_MN && ep.conv_which == 2) {
            // weight gradient: k index = output pixel (64 per block), columns = (tap, channel)
            const ConvPixel px = conv_pixel(ep, kb * BK);
#pragma unroll
            for (int j = 0; j < BN / 64; ++j) {
              const int col = n0 + j * 64, tap = col / ep.cv_C;
              tma_load_im2col(b_dst + j * (64 * BK * 2), &tmB, &full[stage], col - tap * ep.cv_C,
                              px.w, px.h, px.n, (uint16_t)(tap % ep.cv_k), (uint16_t)(tap / ep.cv_k));
            }
          } else if constexpr (!B_MN) {
            load(b_dst, &tmB, kb * BK, n0);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              load(b_dst + j * (64 * BK * 2), &tmB, n0 + j * 64, kb * BK);
          }
          if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 5) {
    // ===================== MMA issuer (one lane) =====================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN, B_MN);
      // K-major: rows of 128 B, 8-row groups 1024 B apart (SBO); LBO unused.
      // MN-major: 64-element (128 B) MN atoms LBO apart, 8-k-row groups SBO apart.
      const uint32_t A_LBO = ep.a_lbo, A_SBO = ep.a_sbo, A_KSTEP = ep.a_kstep;
      const uint32_t B_LBO = ep.b_lbo, B_SBO = ep.b_sbo, B_KSTEP = ep.b_kstep;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x) {
        const int tile = unit % num_tiles;
        const int kb0 = (unit / num_tiles) * kper, kb1 = min(kblocks, kb0 + kper);
        if (BATCHED && causal_skip<BN>(ep, tile, tiles_per_batch, tiles_m, tiles_n)) continue;
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_base = smem_u32(sA + stage * C::A_BYTES);
          const uint32_t b_base = smem_u32(sB + stage * C::B_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t adesc = make_smem_desc_sw128(a_base + k * A_KSTEP, A_LBO, A_SBO);
            const uint64_t bdesc = make_smem_desc_sw128(b_base + k * B_KSTEP, B_LBO, B_SBO);
            umma_bf16_ss(d_tmem, adesc, bdesc, idesc, (kb != kb0 || k != 0) ? 1u : 0u);
          }
          umma_commit(&empty[stage]);   // frees the smem slot when these MMAs retire
          if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull[acc]);       // accumulator complete -> epilogue
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
    __syncwarp();
  } else {
    // ===================== epilogue warps 0..3 and 6..9 =====================
    // Warp w may touch TMEM lane quadrant w % 4; warps 0-3 take the left half of the tile's
    // columns, warps 6-9 the right half (two warps per scheduler hide each other's latency).
    const int q = warp & 3;
    const int hf = warp >= 6 ? 1 : 0;
    constexpr int HCH = BN / 64;            // 32-column chunks per half
    float* st = sEpi + (hf * 4 + q) * 32 * EPI_PITCH;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x) {
      const int tile = unit % num_tiles;
      if (BATCHED && causal_skip<BN>(ep, tile, tiles_per_batch, tiles_m, tiles_n)) continue;
      int mb, nb;
      const int bz = BATCHED ? tile / tiles_per_batch : 0;
      tile_coords(BATCHED ? tile - bz * tiles_per_batch : tile, tiles_m, tiles_n, 16, mb, nb);
      const long long ooff =
          BATCHED ? (long long)(bz % ep.nb1) * ep.o_bs1 + (long long)(bz / ep.nb1) * ep.o_bs2 : 0ll;
      const int m0 = mb * BM;
      const int n0 = nb * BN;
      uint2 mN[8], rN[8];
      const int rbase = m0 + q * 32;
      epilogue_prefetch(ep, lane, rbase, n0 + hf * (BN / 2), M, N, mN, rN);   // before the accumulator is ready
      // one polling lane per warp: 8 idle epilogue warps x 32 spinning lanes cost issue slots
      // and power while the main loop runs under the power cap
      if (DTG_EPI_ONE_LANE_POLL) {
        if (lane == 0) mbar_wait(&tfull[acc], acc_phase);
        __syncwarp();
      } else {
        mbar_wait_parked(&tfull[acc], acc_phase, ep.epi_sleep_ns);
      }
      tc_fence_after();
      const bool attn = BATCHED && ep.mode >= EPI_ATTN_STATS && ep.mode <= EPI_ATTN_DS;
      if (attn && hf == 0)                  // legacy per-tile softmax epilogues: one thread per row
        attn_epilogue_tile<BN>(ep, tmem_base + (uint32_t(q * 32) << 16) + acc * BN, rbase + lane,
                               rbase, M, N, n0, nb, tiles_n, (long long)bz * M + rbase + lane, ooff);
#pragma unroll 1
      for (int hc = 0; hc < (attn ? 0 : HCH); ++hc) {
        const int chunk = hf * HCH + hc;
        const int cbase = n0 + chunk * 32;
        if (cbase >= N) break;
        if (ep.mode >= 100) {                 // timing experiments: 100 = skip, 101 = TMEM read only
          if (ep.mode == 101) {
            uint32_t rr[32];
            tmem_ld_32x32(tmem_base + (uint32_t(q * 32) << 16) + acc * BN + chunk * 32, rr);
            tmem_ld_wait();
            if (rr[0] == 0x7fc12345u) st[lane] = __uint_as_float(rr[1]);
          }
          continue;
        }
        uint2 mC[8], rC[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { mC[i] = mN[i]; rC[i] = rN[i]; }
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + (uint32_t(q * 32) << 16) + acc * BN + chunk * 32, r);
        if (hc + 1 < HCH) epilogue_prefetch(ep, lane, rbase, cbase + 32, M, N, mN, rN);
        tmem_ld_wait();
        epilogue_chunk(ep, st, lane, r, rbase, cbase, M, N, ooff, mC, rC);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    epilogue_drain(ep, lane);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
}

// ---------------------------------------------------------------------------
// 2-CTA variant: a cluster of two CTAs (one SM pair) computes a 256 x BN tile with
// tcgen05.mma.cta_group::2.  Each CTA stages its own 128 rows of A and HALF of the B
// tile; the pair's tensor cores read both halves, so per-SM shared-memory traffic is
// ~2/3 of the 1-CTA kernel's -- the difference between ~75% and ~90% of cuBLAS.
//   * TMA loads use .cta_group::2 and complete on the LEADER's full barrier
//   * tcgen05.commit multicasts "slot free" / "accumulator ready" to both CTAs
//   * epilogue warps of both CTAs arrive remotely on the leader's tmem-empty barrier
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d_2sm(void* smem_dst, const CUtensorMap* m,
                                                uint64_t* bar, int32_t c0, int32_t c1) {
  // the barrier operand addresses the leader CTA's copy (peer bit cleared)
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];\n" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1)
      : "memory");
}
// Same load with an L2 eviction-priority hint (CUTLASS's encodings: evict_first
// 0x12F0000000000000, evict_last 0x14F0000000000000); hint 0 = the plain instruction.
__device__ __forceinline__ void tma_load_2d_2sm_hint(void* smem_dst, const CUtensorMap* m,
                                                     uint64_t* bar, int32_t c0, int32_t c1,
                                                     uint64_t hint) {
  if (hint == 0) { tma_load_2d_2sm(smem_dst, m, bar, c0, c1); return; }
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      ".L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;\n" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1),
      "l"(hint)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_ss_2sm(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                                 uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64"
      " [%0], %1;\n" ::"r"(smem_u32(bar)),
      "h"((uint16_t)3)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_cta(uint64_t* bar, uint32_t cta) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}\n" ::"r"(smem_u32(bar)),
      "r"(cta)
      : "memory");
}

// NST = 0: as many operand stages as fit (6 x 32 KB for the 256-wide tile); NST > 0: exactly
// that many (A/B experiments: DTG_GEMM2_STAGES=5 is the round-1 depth).
template <int BN, int NST = 0>
struct Gemm2Cfg {
  static constexpr int A_BYTES = BM * BK * 2;            // this CTA's 128 rows of A
  static constexpr int B_BYTES = (BN / 2) * BK * 2;      // this CTA's half of B
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_BYTES = 256;
  static constexpr int BUDGET = 227 * 1024 - 1024 - EPI_FLOATS * 4 - BAR_BYTES;
  static constexpr int STAGES_RAW = BUDGET / STAGE_BYTES;
  static constexpr int STAGES_MAX = STAGES_RAW > 8 ? 8 : STAGES_RAW;
  static constexpr int STAGES = (NST > 0 && NST < STAGES_MAX) ? NST : STAGES_MAX;
  static constexpr int TMEM_COLS = 2 * BN;
  static constexpr int SMEM_BYTES = 1024 + STAGES * STAGE_BYTES + EPI_FLOATS * 4 + BAR_BYTES;
};

template <int BN, bool A_MN, bool B_MN, int NST = 0>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_2cta_kernel(const __grid_constant__ CUtensorMap tmA,
                      const __grid_constant__ CUtensorMap tmB, int M, int N, int K,
                      const __grid_constant__ GemmEpilogue ep) {
  using C = Gemm2Cfg<BN, NST>;
  constexpr int BM2 = 2 * BM;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = sA + C::STAGES * C::A_BYTES;
  float* sEpi = reinterpret_cast<float*>(sB + C::STAGES * C::B_BYTES);
  uint64_t* full = reinterpret_cast<uint64_t*>(sEpi + EPI_FLOATS);
  uint64_t* empty = full + C::STAGES;
  uint64_t* tfull = empty + C::STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();
  const bool leader = cta == 0;
  const int tiles_m = (M + BM2 - 1) / BM2;
  const int tiles_n = (N + BN - 1) / BN;
  const int num_tiles = tiles_m * tiles_n;
  const int kblocks = (K + BK - 1) / BK;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;

  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 5 && lane == 0) {
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(&full[s], 1);     // leader's producer arms it with both CTAs' bytes
      mbar_init(&empty[s], 1);    // one multicast commit per use
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], 16);  // 8 epilogue warps x 2 CTAs (leader's copy is used)
    }
    fence_barrier_init();
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(
                     smem_u32(tmem_slot)),
                 "r"((uint32_t)C::TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;\n" ::: "memory");
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

  if (warp == 4) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        int mb, nb;
        tile_coords(tile, tiles_m, tiles_n, ep.group, mb, nb);
        const int m0 = mb * BM2 + (int)cta * BM;
        const int n0 = nb * BN + (int)cta * (BN / 2);
        const uint64_t hint = ep.l2hint;
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          if (leader) mbar_arrive_expect_tx(&full[stage], 2 * C::STAGE_BYTES);
          uint8_t* a_dst = sA + stage * C::A_BYTES;
          uint8_t* b_dst = sB + stage * C::B_BYTES;
          if (!A_MN && ep.conv_which == 1) {
            const int kcol = kb * BK, tap = kcol / ep.cv_C;
            const ConvPixel px = conv_pixel(ep, m0);
            tma_load_im2col_2sm(a_dst, &tmA, &full[stage], kcol - tap * ep.cv_C, px.w, px.h, px.n,
                                (uint16_t)(tap % ep.cv_k), (uint16_t)(tap / ep.cv_k));
          } else if constexpr (!A_MN) {
            tma_load_2d_2sm_hint(a_dst, &tmA, &full[stage], kb * BK, m0, hint);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j)
              tma_load_2d_2sm_hint(a_dst + j * (64 * BK * 2), &tmA, &full[stage], m0 + j * 64, kb * BK, hint);
          }
          if constexpr (!B_MN) {
            tma_load_2d_2sm_hint(b_dst, &tmB, &full[stage], kb * BK, n0, hint);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 128; ++j)
              tma_load_2d_2sm_hint(b_dst + j * (64 * BK * 2), &tmB, &full[stage], n0 + j * 64, kb * BK, hint);
          }
          if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 5) {
    // ===================== MMA issuer (leader CTA, one lane) =====================
    if (leader && lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM2, BN, A_MN, B_MN);
      const uint32_t A_LBO = ep.a_lbo, A_SBO = ep.a_sbo, A_KSTEP = ep.a_kstep;
      const uint32_t B_LBO = ep.b_lbo, B_SBO = ep.b_sbo, B_KSTEP = ep.b_kstep;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_base = smem_u32(sA + stage * C::A_BYTES);
          const uint32_t b_base = smem_u32(sB + stage * C::B_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t adesc = make_smem_desc_sw128(a_base + k * A_KSTEP, A_LBO, A_SBO);
            const uint64_t bdesc = make_smem_desc_sw128(b_base + k * B_KSTEP, B_LBO, B_SBO);
            umma_bf16_ss_2sm(d_tmem, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit_2sm(&empty[stage]);
          if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit_2sm(&tfull[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
    __syncwarp();
  } else {
    // ===================== epilogue warps 0..3, 6..9 (both CTAs) =====================
    const int q = warp & 3;
    const int hf = warp >= 6 ? 1 : 0;
    constexpr int HCH = BN / 64;
    float* st = sEpi + (hf * 4 + q) * 32 * EPI_PITCH;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      int mb, nb;
      tile_coords(tile, tiles_m, tiles_n, ep.group, mb, nb);
      const int m0 = mb * BM2 + (int)cta * BM;
      const int n0 = nb * BN;
      uint2 mN[8], rN[8];
      const int rbase = m0 + q * 32;
      epilogue_prefetch(ep, lane, rbase, n0 + hf * (BN / 2), M, N, mN, rN);   // before the accumulator is ready
      // one polling lane per warp: 8 idle epilogue warps x 32 spinning lanes cost issue slots
      // and power while the main loop runs under the power cap
      if (DTG_EPI_ONE_LANE_POLL) {
        if (lane == 0) mbar_wait(&tfull[acc], acc_phase);
        __syncwarp();
      } else {
        mbar_wait_parked(&tfull[acc], acc_phase, ep.epi_sleep_ns);
      }
      tc_fence_after();
#pragma unroll 1
      for (int hc = 0; hc < HCH; ++hc) {
        const int chunk = hf * HCH + hc;
        const int cbase = n0 + chunk * 32;
        if (cbase >= N) break;
        if (ep.mode >= 100) {                 // timing experiments: 100 = skip, 101 = TMEM read only
          if (ep.mode == 101) {
            uint32_t rr[32];
            tmem_ld_32x32(tmem_base + (uint32_t(q * 32) << 16) + acc * BN + chunk * 32, rr);
            tmem_ld_wait();
            if (rr[0] == 0x7fc12345u) st[lane] = __uint_as_float(rr[1]);
          }
          continue;
        }
        uint2 mC[8], rC[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { mC[i] = mN[i]; rC[i] = rN[i]; }
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + (uint32_t(q * 32) << 16) + acc * BN + chunk * 32, r);
        if (hc + 1 < HCH) epilogue_prefetch(ep, lane, rbase, cbase + 32, M, N, mN, rN);
        tmem_ld_wait();
        epilogue_chunk(ep, st, lane, r, rbase, cbase, M, N, 0ll, mC, rC);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cta(&tempty[acc], 0);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    epilogue_drain(ep, lane);
  }
  tc_fence_before();
  cluster_sync_all();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base),
                 "r"((uint32_t)C::TMEM_COLS)
                 : "memory");
  }
}

// ---------------------------------------------------------------------------
// "Tall" pair kernel: 512 x 256 output per CTA pair (256 rows per CTA), the shape cuBLAS'
// own sm_100 kernel uses for large bf16 GEMMs (nvjet ..256x256_64x4..2cta; ncu side by side in
// profiles/gemm_tall_vs_cublas.md).  Why: with 128 rows per CTA every SM has to take in
// 32 KB of operands per 512 tensor-core cycles = 64 B/clk, and the SM's operand intake tops out
// near 48 B/clk (the 256 x 256 pair tile sits at 75 % tensor-pipe activity whatever the
// pipeline depth, L2 hints or multicast).  With 256 rows per CTA the B half is reused by two
// MMAs: 48 KB per 1024 cycles = 48 B/clk.
//
// TMEM: both 256-column halves hold one accumulator each (rows 0-127 and 128-255 of the CTA),
// so a big tile's epilogue is not hidden behind the next big tile's main loop (cuBLAS does the
// same).  Two things soften that: the TMA producer keeps filling the ring during the epilogue,
// and the schedule ends with 256-row "small" tiles that use ONE half and alternate halves, so
// the tail of the problem (where 512-row tiles would leave SMs idle for up to a whole tile)
// is both finer-grained and double-buffered.
//
// Static schedule, computed identically by every role of both CTAs (no communication):
//   big tiles    b = c, c+C, ...            < n_big        (c = cluster id, C = clusters)
//   small tiles, phase A: clusters that got one big tile fewer (c >= first_short) take `extra`
//                rounds among themselves first, then phase B: round robin over all clusters.
// The host picks (rows done as big tiles, extra) by minimising the slowest cluster's load.
// ---------------------------------------------------------------------------
struct TallCfg {
  static constexpr int BN = 256;
  static constexpr int A_BYTES = 2 * BM * BK * 2;        // this CTA's 256 rows of A
  static constexpr int B_BYTES = (BN / 2) * BK * 2;      // this CTA's half of B
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;  // 48 KB
  static constexpr int STAGES = 4;
  static constexpr int BAR_BYTES = 256;
  static constexpr int TMEM_COLS = 512;
  static constexpr int SMEM_BYTES = 1024 + STAGES * STAGE_BYTES + EPI_FLOATS * 4 + BAR_BYTES;
};
static_assert(TallCfg::SMEM_BYTES <= 227 * 1024, "tall GEMM: shared memory budget");

struct TallSched {
  int n_big, n_small, clusters, c, first_short, n_short, extra;
  int ib = 0, ia = 0, ip = 0, small_seen = 0;
  __host__ __device__ TallSched(const GemmEpilogue& ep, int cluster_id, int num_clusters)
      : n_big(ep.tall_nbig), n_small(ep.tall_nsmall), clusters(num_clusters), c(cluster_id),
        first_short(ep.tall_nbig % num_clusters), extra(ep.tall_extra) {
    n_short = clusters - first_short;
  }
  // kind: 1 = big tile `idx`; 0 = small tile `idx`, accumulator half `half`
  __host__ __device__ bool next(int& kind, int& idx, int& half) {
    const int b = c + ib * clusters;
    if (b < n_big) { ++ib; kind = 1; idx = b; half = 0; return true; }
    if (c >= first_short) {
      while (ia < extra) {
        const int j = (c - first_short) + ia * n_short;
        ++ia;
        if (j < n_small) { kind = 0; idx = j; half = (small_seen++) & 1; return true; }
      }
    }
    const int j = extra * n_short + c + ip * clusters;
    if (j < n_small) { ++ip; kind = 0; idx = j; half = (small_seen++) & 1; return true; }
    return false;
  }
};

template <bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_tall_kernel(const __grid_constant__ CUtensorMap tmA,
                      const __grid_constant__ CUtensorMap tmB, int M, int N, int K,
                      const __grid_constant__ GemmEpilogue ep) {
  using C = TallCfg;
  constexpr int BN = C::BN;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = sA + C::STAGES * C::A_BYTES;
  float* sEpi = reinterpret_cast<float*>(sB + C::STAGES * C::B_BYTES);
  uint64_t* full = reinterpret_cast<uint64_t*>(sEpi + EPI_FLOATS);
  uint64_t* empty = full + C::STAGES;
  uint64_t* tfull = empty + C::STAGES;      // [2]: one per accumulator half
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();
  const bool leader = cta == 0;
  const int tiles_n = (N + BN - 1) / BN;
  const int big_rows = ep.tall_big_mblocks * 4 * BM;                 // rows covered by 512-row tiles
  const int small_m = (M - big_rows + 2 * BM - 1) / (2 * BM);        // 256-row blocks after them
  const int kblocks = (K + BK - 1) / BK;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;

  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 5 && lane == 0) {
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], 16);  // 8 epilogue warps x 2 CTAs (leader's copy is used)
    }
    fence_barrier_init();
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(
                     smem_u32(tmem_slot)),
                 "r"((uint32_t)C::TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;\n" ::: "memory");
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

  // tile index -> first row of this CTA's block, first column of the pair's tile
  auto coords = [&](int kind, int idx, int& m0, int& n0) {
    int mb, nb;
    if (kind) {
      tile_coords(idx, ep.tall_big_mblocks, tiles_n, ep.group > 1 ? ep.group / 2 : 1, mb, nb);
      m0 = mb * 4 * BM + (int)cta * 2 * BM;
    } else {
      tile_coords(idx, small_m, tiles_n, ep.group, mb, nb);
      m0 = big_rows + mb * 2 * BM + (int)cta * BM;
    }
    n0 = nb * BN;
  };

  if (warp == 4) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      TallSched sch(ep, cluster_id, num_clusters);
      int stage = 0;
      uint32_t phase = 0;
      int kind, idx, half;
      while (sch.next(kind, idx, half)) {
        int m0, n0;
        coords(kind, idx, m0, n0);
        const int nb0 = n0 + (int)cta * (BN / 2);
        const uint32_t tx = kind ? 2 * C::STAGE_BYTES : 2 * (BM * BK * 2 + C::B_BYTES);
        const int a_boxes = kind ? 2 : 1;            // 128-row boxes of A for this CTA
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          if (leader) mbar_arrive_expect_tx(&full[stage], tx);
          uint8_t* a_dst = sA + stage * C::A_BYTES;
          uint8_t* b_dst = sB + stage * C::B_BYTES;
          for (int h = 0; h < a_boxes; ++h) {
            uint8_t* ah = a_dst + h * (BM * BK * 2);
            if constexpr (!A_MN) {
              tma_load_2d_2sm(ah, &tmA, &full[stage], kb * BK, m0 + h * BM);
            } else {
#pragma unroll
              for (int j = 0; j < BM / 64; ++j)
                tma_load_2d_2sm(ah + j * (64 * BK * 2), &tmA, &full[stage], m0 + h * BM + j * 64,
                                kb * BK);
            }
          }
          if constexpr (!B_MN) {
            tma_load_2d_2sm(b_dst, &tmB, &full[stage], kb * BK, nb0);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 128; ++j)
              tma_load_2d_2sm(b_dst + j * (64 * BK * 2), &tmB, &full[stage], nb0 + j * 64, kb * BK);
          }
          if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 5) {
    // ===================== MMA issuer (leader CTA, one lane) =====================
    if (leader && lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(2 * BM, BN, A_MN, B_MN);
      const uint32_t A_LBO = ep.a_lbo, A_SBO = ep.a_sbo, A_KSTEP = ep.a_kstep;
      const uint32_t B_LBO = ep.b_lbo, B_SBO = ep.b_sbo, B_KSTEP = ep.b_kstep;
      TallSched sch(ep, cluster_id, num_clusters);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t eph[2] = {0, 0};               // phase of tempty[h] this role waits for next
      int kind, idx, half;
      while (sch.next(kind, idx, half)) {
        const int h0 = kind ? 0 : half, h1 = kind ? 1 : half;
        for (int h = h0; h <= h1; ++h) {
          mbar_wait(&tempty[h], eph[h] ^ 1);
          eph[h] ^= 1;
        }
        tc_fence_after();
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_base = smem_u32(sA + stage * C::A_BYTES);
          const uint32_t b_base = smem_u32(sB + stage * C::B_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t bdesc = make_smem_desc_sw128(b_base + k * B_KSTEP, B_LBO, B_SBO);
            // big: rows 0-127 -> columns [0,256), rows 128-255 -> [256,512); small: one half
            const uint64_t adesc0 = make_smem_desc_sw128(a_base + k * A_KSTEP, A_LBO, A_SBO);
            umma_bf16_ss_2sm(tmem_base + h0 * BN, adesc0, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
            if (kind) {
              const uint64_t adesc1 =
                  make_smem_desc_sw128(a_base + BM * BK * 2 + k * A_KSTEP, A_LBO, A_SBO);
              umma_bf16_ss_2sm(tmem_base + BN, adesc1, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
            }
          }
          umma_commit_2sm(&empty[stage]);
          if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit_2sm(&tfull[h0]);
        if (kind) umma_commit_2sm(&tfull[1]);
      }
    }
    __syncwarp();
  } else {
    // ===================== epilogue warps 0..3, 6..9 (both CTAs) =====================
    const int q = warp & 3;
    const int hf = warp >= 6 ? 1 : 0;
    constexpr int HCH = BN / 64;
    float* st = sEpi + (hf * 4 + q) * 32 * EPI_PITCH;
    TallSched sch(ep, cluster_id, num_clusters);
    uint32_t fph[2] = {0, 0};                 // phase of tfull[h] to wait for next
    int kind, idx, half;
    while (sch.next(kind, idx, half)) {
      int m0, n0;
      coords(kind, idx, m0, n0);
      const int h0 = kind ? 0 : half, h1 = kind ? 1 : half;
      for (int h = h0; h <= h1; ++h) {
        // rows of this warp: big tile -> CTA block row 128*h + 32*q; small tile -> 32*q
        const int rbase = m0 + (kind ? h * BM : 0) + q * 32;
        uint2 mN[8], rN[8];
        epilogue_prefetch(ep, lane, rbase, n0 + hf * (BN / 2), M, N, mN, rN);
        mbar_wait_parked(&tfull[h], fph[h], ep.epi_sleep_ns);
        fph[h] ^= 1;
        tc_fence_after();
#pragma unroll 1
        for (int hc = 0; hc < HCH; ++hc) {
          const int chunk = hf * HCH + hc;
          const int cbase = n0 + chunk * 32;
          if (cbase >= N) break;
          if (ep.mode >= 100) {               // timing experiments: 100 = skip, 101 = TMEM read only
            if (ep.mode == 101) {
              uint32_t rr[32];
              tmem_ld_32x32(tmem_base + (uint32_t(q * 32) << 16) + h * BN + chunk * 32, rr);
              tmem_ld_wait();
              if (rr[0] == 0x7fc12345u) st[lane] = __uint_as_float(rr[1]);
            }
            continue;
          }
          uint2 mC[8], rC[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) { mC[i] = mN[i]; rC[i] = rN[i]; }
          uint32_t r[32];
          tmem_ld_32x32(tmem_base + (uint32_t(q * 32) << 16) + h * BN + chunk * 32, r);
          if (hc + 1 < HCH) epilogue_prefetch(ep, lane, rbase, cbase + 32, M, N, mN, rN);
          tmem_ld_wait();
          epilogue_chunk(ep, st, lane, r, rbase, cbase, M, N, 0ll, mC, rC);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cta(&tempty[h], 0);
      }
    }
    epilogue_drain(ep, lane);
  }
  tc_fence_before();
  cluster_sync_all();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base),
                 "r"((uint32_t)C::TMEM_COLS)
                 : "memory");
  }
}


// ---------------------------------------------------------------------------
// Host side: tensor-map cache + dispatch
// ---------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  });
  return fn;
}

struct MapKey {
  const void* ptr;
  uint64_t inner, outer, ld;
  uint32_t box_inner, box_outer;
  uint64_t nb1, nb2, bs1, bs2;   // batch dims/strides (elements); nb1 = 0 -> plain 2-D map
  uint64_t f32;                  // 1: fp32 elements (outputs written through TMA), 0: bf16
  uint64_t sw64;                 // 1: 64B swizzle (bf16 output boxes), 0: 128B
  bool operator==(const MapKey& o) const { return std::memcmp(this, &o, sizeof(MapKey)) == 0; }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    const uint64_t* p = reinterpret_cast<const uint64_t*>(&k);
    size_t h = 1469598103934665603ull;
    for (size_t i = 0; i < sizeof(MapKey) / 8; ++i) h = (h ^ p[i]) * 1099511628211ull;
    return h;
  }
};
static std::mutex g_map_mu;
static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> g_maps;

// 2D bf16 tensor [outer, inner] with row pitch `ld` elements; box = [box_outer, box_inner].
static int get_tensor_map(const void* ptr, uint64_t inner, uint64_t outer, uint64_t ld,
                          uint32_t box_inner, uint32_t box_outer, CUtensorMap* out,
                          uint64_t nb1 = 0, uint64_t nb2 = 0, uint64_t bs1 = 0, uint64_t bs2 = 0,
                          bool f32 = false, bool sw64 = false) {
  MapKey key;
  std::memset(&key, 0, sizeof(key));
  key.ptr = ptr; key.inner = inner; key.outer = outer; key.ld = ld;
  key.box_inner = box_inner; key.box_outer = box_outer;
  key.nb1 = nb1; key.nb2 = nb2; key.bs1 = bs1; key.bs2 = bs2; key.f32 = f32 ? 1 : 0; key.sw64 = sw64 ? 1 : 0;
  {
    std::lock_guard<std::mutex> g(g_map_mu);
    auto it = g_maps.find(key);
    if (it != g_maps.end()) { *out = it->second; return OK; }
  }
  PFN_encodeTiled enc = get_encode();
  if (enc == nullptr) {
    dtg_set_last_error("cuTensorMapEncodeTiled", "driver entry point not available");
    return ERR_TMA_ENCODE;
  }
  const int rank = nb1 ? 4 : 2;
  cuuint64_t dims[4] = {inner, outer, nb1 ? nb1 : 1, nb2 ? nb2 : 1};
  const uint64_t esz = f32 ? 4 : 2;
  cuuint64_t strides[3] = {ld * esz, bs1 * esz, bs2 * esz};
  cuuint32_t box[4] = {box_inner, box_outer, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(out, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16,
                   rank, const_cast<void*>(ptr), dims,
                   strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   sw64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[160];
    snprintf(buf, sizeof(buf), "CUresult %d (ptr %p inner %llu outer %llu ld %llu box %ux%u)",
             (int)r, ptr, (unsigned long long)inner, (unsigned long long)outer,
             (unsigned long long)ld, box_inner, box_outer);
    dtg_set_last_error("cuTensorMapEncodeTiled", buf);
    return ERR_TMA_ENCODE;
  }
  std::lock_guard<std::mutex> g(g_map_mu);
  if (g_maps.size() > 4096) g_maps.clear();
  g_maps.emplace(key, *out);
  return OK;
}

// im2col-mode tensor map over an NHWC bf16 activation tensor (implicit-GEMM convolution): the
// bounding box of filter-window corners is [-pad, dim - 1 + pad - (k - 1)], traversed with the
// convolution stride; a load delivers `pixels` consecutive output pixels x 64 channels.
struct ConvDesc {
  const void* x;
  int N, H, W, C, k, stride, pad;
  int which;          // 1: A operand = im2col(x) (forward), 2: B operand = im2col(x) (wgrad)
};
typedef CUresult (*PFN_encodeIm2col)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                     const cuuint64_t*, const cuuint64_t*, const int*, const int*,
                                     cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                     CUtensorMapSwizzle, CUtensorMapL2promotion,
                                     CUtensorMapFloatOOBfill);
static int get_im2col_map(const ConvDesc& cv, uint32_t pixels, CUtensorMap* out) {
  static PFN_encodeIm2col fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &p, cudaEnableDefault, &qres) ==
            cudaSuccess && qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeIm2col>(p);
  });
  if (fn == nullptr) {
    dtg_set_last_error("cuTensorMapEncodeIm2col", "driver entry point not available");
    return ERR_TMA_ENCODE;
  }
  struct Key { const void* x; int g[8]; uint32_t pixels; };
  static std::mutex mu;
  static std::vector<std::pair<Key, CUtensorMap>> cache;     // a few dozen layers: linear scan
  Key key;
  std::memset(&key, 0, sizeof(key));
  key.x = cv.x; key.pixels = pixels;
  const int g[8] = {cv.N, cv.H, cv.W, cv.C, cv.k, cv.stride, cv.pad, 0};
  std::memcpy(key.g, g, sizeof(g));
  {
    std::lock_guard<std::mutex> lk(mu);
    for (auto& kv : cache)
      if (std::memcmp(&kv.first, &key, sizeof(Key)) == 0) { *out = kv.second; return OK; }
  }
  cuuint64_t dims[4] = {(cuuint64_t)cv.C, (cuuint64_t)cv.W, (cuuint64_t)cv.H, (cuuint64_t)cv.N};
  cuuint64_t strides[3] = {(cuuint64_t)cv.C * 2, (cuuint64_t)cv.W * cv.C * 2,
                           (cuuint64_t)cv.H * cv.W * cv.C * 2};
  int lower[2] = {-cv.pad, -cv.pad};
  int upper[2] = {cv.pad - (cv.k - 1), cv.pad - (cv.k - 1)};
  cuuint32_t estr[4] = {1, (cuuint32_t)cv.stride, (cuuint32_t)cv.stride, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(cv.x), dims, strides,
                  lower, upper, 64, pixels, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[200];
    snprintf(buf, sizeof(buf), "CUresult %d (x %p N %d H %d W %d C %d k %d stride %d pad %d pixels %u)",
             (int)r, cv.x, cv.N, cv.H, cv.W, cv.C, cv.k, cv.stride, cv.pad, pixels);
    dtg_set_last_error("cuTensorMapEncodeIm2col", buf);
    return ERR_TMA_ENCODE;
  }
  std::lock_guard<std::mutex> lk(mu);
  if (cache.size() > 512) cache.clear();
  cache.emplace_back(key, *out);
  return OK;
}

template <int BN, bool A_MN, bool B_MN, bool BATCHED = false>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, int M, int N, int K,
                       const GemmEpilogue& ep, int num_sms, cudaStream_t stream) {
  using C = GemmCfg<BN>;
  auto kern = gemm_bf16_kernel<BN, A_MN, B_MN, BATCHED>;
  static bool configured = false;
  if (!configured) {
    DTG_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     C::SMEM_BYTES));
    configured = true;
  }
  const long long tiles = (long long)((M + BM - 1) / BM) * ((N + BN - 1) / BN) *
                          (BATCHED ? (long long)ep.nb1 * ep.nb2 : 1) *
                          ((!BATCHED && ep.ksplit > 1) ? ep.ksplit : 1);
  const int grid = tiles < num_sms ? (int)tiles : num_sms;
  kern<<<grid, GEMM_THREADS, C::SMEM_BYTES, stream>>>(ta, tb, M, N, K, ep);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

template <int BN, bool A_MN, bool B_MN, int NST = 0>
static int launch_gemm_2cta(const CUtensorMap& ta, const CUtensorMap& tb, int M, int N, int K,
                            const GemmEpilogue& ep, int num_sms, cudaStream_t stream) {
  using C = Gemm2Cfg<BN, NST>;
  auto kern = gemm_bf16_2cta_kernel<BN, A_MN, B_MN, NST>;
  static bool configured = false;
  if (!configured) {
    DTG_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     C::SMEM_BYTES));
    configured = true;
  }
  const int tiles = ((M + 2 * BM - 1) / (2 * BM)) * ((N + BN - 1) / BN);
  const int pairs = num_sms / 2;
  const int grid = 2 * (tiles < pairs ? tiles : pairs);
  kern<<<grid, GEMM_THREADS, C::SMEM_BYTES, stream>>>(ta, tb, M, N, K, ep);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}


template <bool A_MN, bool B_MN>
static int launch_gemm_tall(const CUtensorMap& ta, const CUtensorMap& tb, int M, int N, int K,
                            const GemmEpilogue& ep, int num_sms, cudaStream_t stream) {
  using C = TallCfg;
  auto kern = gemm_bf16_tall_kernel<A_MN, B_MN>;
  static bool configured = false;
  if (!configured) {
    DTG_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     C::SMEM_BYTES));
    configured = true;
  }
  const int tiles = ep.tall_nbig + ep.tall_nsmall;
  const int pairs = num_sms / 2;
  const int grid = 2 * (tiles < pairs ? tiles : pairs);
  kern<<<grid, GEMM_THREADS, C::SMEM_BYTES, stream>>>(ta, tb, M, N, K, ep);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

// Schedule of the tall kernel: how many 512-row m-blocks run as big tiles, and how many rounds of
// small tiles go first to the clusters that got one big tile fewer -- chosen by minimising the
// slowest cluster's load (big tile = 1, small tile = `small_cost`; TallSched is the device mirror).
static void plan_tall(int M, int N, int clusters, float small_cost, GemmEpilogue& ep) {
  const int tn = (N + 255) / 256;
  float best = 1e30f;
  ep.tall_big_mblocks = 0; ep.tall_nbig = 0; ep.tall_extra = 0;
  ep.tall_nsmall = ((M + 255) / 256) * tn;
  for (int nbm = 0; nbm <= M / 512; ++nbm) {
    const int n_big = nbm * tn;
    const int n_small = ((M - nbm * 512 + 255) / 256) * tn;
    const int C = std::min(clusters, n_big + n_small);
    if (C <= 0) continue;
    const int fs = n_big % C, n_short = C - fs;
    for (int extra = 0; extra <= 3; ++extra) {
      float worst = 0.f;
      for (int c = 0; c < C; ++c) {
        const int nb = c < n_big ? (n_big - c + C - 1) / C : 0;
        int ns = 0;
        if (c >= fs)
          for (int ia = 0; ia < extra; ++ia)
            if ((c - fs) + ia * n_short < n_small) ++ns;
        const int j0 = extra * n_short + c;
        if (j0 < n_small) ns += (n_small - j0 + C - 1) / C;
        worst = std::max(worst, (float)nb + small_cost * (float)ns);
      }
      if (worst < best - 1e-4f || (worst < best + 1e-4f && nbm > ep.tall_big_mblocks)) {
        best = worst;
        ep.tall_big_mblocks = nbm; ep.tall_nbig = n_big; ep.tall_nsmall = n_small;
        ep.tall_extra = extra;
      }
    }
  }
}

static int g_num_sms = 0;
static int g_use_2cta = -1;
static int g_2cta_stages = -1;   // DTG_GEMM2_STAGES=5: round-1 depth; default: as many as fit (6)
static unsigned long long g_l2hint = ~0ull;   // DTG_GEMM_L2HINT: 0 none (default), 1 evict_first, 2 evict_last
static int g_use_tall = 1;       // DTG_GEMM_TALL=0: never use the 512 x 256 pair-tile kernel
static int g_tall_min_k = 6144;   // DTG_GEMM_TALL_MIN_K: a big tile's epilogue is exposed (all 512 TMEM
                                  // columns hold its accumulators), so it only pays when the main
                                  // loop is long: measured break-even between K = 4096 and 8192
static float g_tall_small_cost = 0.62f;   // DTG_GEMM_TALL_SMALL_COST: small tile / big tile time
static int g_group = -1;         // DTG_GEMM_GROUP: m-blocks per tile-order group (default 8)

int tensor_map_bf16(const void* ptr, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_inner,
                    uint32_t box_outer, CUtensorMap* out, uint64_t nb1, uint64_t nb2, uint64_t bs1,
                    uint64_t bs2) {
  return get_tensor_map(ptr, inner, outer, ld, box_inner, box_outer, out, nb1, nb2, bs1, bs2);
}
int num_sms() {
  if (g_num_sms == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
      return 0;
  }
  return g_num_sms;
}

}  // namespace dtg

using namespace dtg;

// C-ABI entry point.  See include/dtg_b200.h for the contract.
static int gemm_impl(const void* A, const void* B, int M, int N, int K, long long lda,
                     long long ldb, int a_mn, int b_mn, void* out, void* out2,
                     const float* bias, const void* mask, const void* residual,
                     long long ldo, long long ldmask, float alpha, int mode, int act,
                     int mask_mode, int force_bn, void* stream_, const ConvDesc* cv) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if (M <= 0 || N <= 0 || K <= 0 || A == nullptr || B == nullptr || out == nullptr) {
    dtg_set_last_error("dtg_gemm_bf16", "bad shape or null pointer");
    return ERR_BAD_ARG;
  }
  if ((lda % 8) != 0 || (ldb % 8) != 0 || (reinterpret_cast<uintptr_t>(A) & 15) ||
      (reinterpret_cast<uintptr_t>(B) & 15)) {
    dtg_set_last_error("dtg_gemm_bf16", "operands need 16-byte aligned base and row pitch (TMA)");
    return ERR_BAD_ARG;
  }
  if (g_num_sms == 0) {
    int dev = 0;
    DTG_CUDA_OK(cudaGetDevice(&dev));
    DTG_CUDA_OK(cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev));
  }
  if (g_use_2cta < 0) {
    const char* e = getenv("DTG_GEMM_2CTA");
    g_use_2cta = (e != nullptr && e[0] == '0') ? 0 : 1;
  }
  // force_bn: 0 = choose; 64/128/256 = 1-CTA kernel with that tile width; 512 = the
  // 2-CTA (cta_group::2) kernel with a 256 x 256 cluster tile (the planner may still pick the
  // tall kernel for large K); 768 = the tall pair kernel (512 x 256 tiles + 256-row tail).
  int BN = force_bn;
  bool two_cta = false, force_tall = false;
  if (cv != nullptr && BN >= 512) BN = 0;
  if (BN == 768) { BN = 256; two_cta = true; force_tall = true; }   // 512 x 256 pair tiles
  else if (BN == 512) { BN = 256; two_cta = true; }
  else if (BN == 0) {
    const int tm = (M + BM - 1) / BM;
    if (N > 128 && tm * ((N + 255) / 256) >= (g_num_sms * 3) / 4) BN = 256;
    else if (N > 64) BN = 128;
    else BN = 64;
    // enough 256-row cluster tiles to fill the SM pairs -> pair kernel
    if (BN == 256 && g_use_2cta && M >= 2 * BM && (cv == nullptr || cv->which == 1) &&
        ((M + 2 * BM - 1) / (2 * BM)) * ((N + 255) / 256) >= (g_num_sms / 2) * 3 / 4)
      two_cta = true;
  }
  if (BN != 64 && BN != 128 && BN != 256) {
    dtg_set_last_error("dtg_gemm_bf16", "force_bn must be 0, 64, 128, 256, 512 or 768");
    return ERR_BAD_ARG;
  }
  GemmEpilogue ep;
  ep.out = out; ep.out2 = out2; ep.bias = bias;
  ep.mask = reinterpret_cast<const __nv_bfloat16*>(mask);
  ep.residual = reinterpret_cast<const __nv_bfloat16*>(residual);
  ep.ldo = ldo; ep.ldmask = ldmask; ep.alpha = alpha; ep.mode = mode; ep.act = act;
  ep.nb1 = ep.nb2 = 1; ep.o_bs1 = ep.o_bs2 = 0; ep.ksplit = 1;
  ep.rowv1 = ep.rowv2 = nullptr; ep.stats = nullptr; ep.scale2 = 0.f; ep.causal = 0;
  if (g_2cta_stages < 0) {
    const char* e = getenv("DTG_GEMM2_STAGES");
    g_2cta_stages = e ? atoi(e) : 0;
    e = getenv("DTG_GEMM_L2HINT");
    const int h = e ? atoi(e) : 0;
    g_l2hint = h == 1 ? 0x12F0000000000000ull : (h == 2 ? 0x14F0000000000000ull : 0ull);
    e = getenv("DTG_GEMM_GROUP");
    g_group = (e && atoi(e) > 0) ? atoi(e) : 8;
    e = getenv("DTG_GEMM_TALL");
    g_use_tall = (e && e[0] == '0') ? 0 : 1;
    e = getenv("DTG_GEMM_TALL_MIN_K");
    if (e) g_tall_min_k = atoi(e);
    e = getenv("DTG_GEMM_TALL_SMALL_COST");
    if (e && atof(e) > 0) g_tall_small_cost = (float)atof(e);
  }
  ep.l2hint = g_l2hint; ep.group = g_group;
  ep.conv_which = 0;
  if (cv != nullptr) {
    ep.conv_which = cv->which; ep.cv_C = cv->C; ep.cv_k = cv->k; ep.cv_stride = cv->stride;
    ep.cv_pad = cv->pad;
    ep.cv_Ho = (cv->H + 2 * cv->pad - cv->k) / cv->stride + 1;
    ep.cv_Wo = (cv->W + 2 * cv->pad - cv->k) / cv->stride + 1;
  }
  {
    static int sleep_ns = -1;
    if (sleep_ns < 0) { const char* e = getenv("DTG_EPI_SLEEP_NS"); sleep_ns = e ? atoi(e) : 0; }
    ep.epi_sleep_ns = (unsigned)sleep_ns;
  }
  ep.mask_mode = mask ? mask_mode : MASK_NONE;
  const uintptr_t oa = reinterpret_cast<uintptr_t>(out);
  bool vec = (ldo % 4 == 0) && ((oa & (mode == EPI_BF16 ? 7 : 15)) == 0);
  if (out2) vec = vec && ((reinterpret_cast<uintptr_t>(out2) & 7) == 0);
  if (residual != nullptr && (reinterpret_cast<uintptr_t>(residual) & 7) != 0) vec = false;
  ep.vec = vec ? 1 : 0;
  ep.fancy = (bias != nullptr || act != ACT_NONE || ep.mask_mode != MASK_NONE ||
              residual != nullptr || out2 != nullptr) ? 1 : 0;
  {
    auto a16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    const long long esz = mode == EPI_BF16 ? 2 : 4;
    bool d = a16(out) && (ldo * esz) % 16 == 0;
    if (bias) d = d && a16(bias);
    if (ep.mask_mode != MASK_NONE) d = d && a16(mask) && (ldmask * 2) % 16 == 0;
    if (residual) d = d && a16(residual);
    if (out2) d = d && a16(out2);
    static int allow = -1;
    if (allow < 0) { const char* e = getenv("DTG_GEMM_DIRECT_EPI"); allow = (e && e[0] == '0') ? 0 : 1; }
    ep.direct = (d && allow) ? 1 : 0;
  }
  ep.vec_mask = (mask != nullptr && (ldmask % 4 == 0) &&
                 (reinterpret_cast<uintptr_t>(mask) & 7) == 0) ? 1 : 0;
  // K-major SW128: 8-row groups 1024 B apart (SBO), K advances 32 B inside the atom.
  // MN-major SW128: 64-element MN atoms one TMA box (64 x BK x 2 B) apart (LBO),
  // 8-k-row groups 1024 B apart (SBO), K advances two groups per UMMA_K.
  const uint32_t mn_lbo = 64 * BK * 2, mn_sbo = 1024, mn_kstep = (UMMA_K / 8) * 1024;
  ep.a_lbo = a_mn ? mn_lbo : 16; ep.a_sbo = a_mn ? mn_sbo : 1024;
  ep.a_kstep = a_mn ? mn_kstep : UMMA_K * 2;
  ep.b_lbo = b_mn ? mn_lbo : 16; ep.b_sbo = b_mn ? mn_sbo : 1024;
  ep.b_kstep = b_mn ? mn_kstep : UMMA_K * 2;

  // split-K for local fp32 outputs (never for pushes to a peer: it would multiply the NVLink
  // traffic): pick the split that minimises waves x (k-blocks per unit + fixed unit cost).
  // (EPI_F32_RED is the push into a LOCAL shard -- remote targets use EPI_F32_RED_COAL.)
  if (!two_cta && !ep.fancy && (mode == EPI_F32 || mode == EPI_F32_ACC || mode == EPI_F32_RED) &&
      force_bn >= 0) {
    static int allow_split = -1;
    if (allow_split < 0) { const char* e = getenv("DTG_GEMM_SPLITK"); allow_split = (e && e[0] == '0') ? 0 : 1; }
    const long long tiles = (long long)((M + BM - 1) / BM) * ((N + BN - 1) / BN);
    const int kblocks = (K + BK - 1) / BK;
    if (allow_split && kblocks >= 16) {
      auto cost = [&](int sp) {
        const long long waves = (tiles * sp + g_num_sms - 1) / g_num_sms;
        return waves * ((kblocks + sp - 1) / sp + 6);
      };
      int best = 1;
      long long best_cost = cost(1);
      for (int sp = 2; sp <= 256 && sp * 4 <= kblocks; ++sp) {
        const long long c = cost(sp);
        if (c < best_cost) { best = sp; best_cost = c; }
      }
      if (best > 1 && best_cost * 4 <= cost(1) * 3) {            // worth it only for >= 25 %
        const int kper = (kblocks + best - 1) / best;
        ep.ksplit = (kblocks + kper - 1) / kper;
        if (mode == EPI_F32)
          DTG_CUDA_OK(cudaMemset2DAsync(out, (size_t)ldo * 4, 0, (size_t)N * 4, (size_t)M, stream));
        ep.mode = EPI_F32_RED;
      }
    }
  }

  // Outputs through the TMA unit.  DTG_PUSH_TMA: 0 off, 1 pushes into remote shards [default],
  // 2 also local red.add outputs incl. split-K partials.  DTG_STORE_TMA: 0 register-direct stores,
  // 1 TMA stores, 2 staged + coalesced stores by the warp itself, for bf16 /
  // fp32 stores of interior chunks (the second gelu output, out2, keeps the register-direct path).
  ep.tma_out = TMAOUT_NONE;
  ep.tma_single = (BN == 64) ? 1 : 0;
  {
    static int push_tma = -1, store_tma = -1;
    if (push_tma < 0) {
      const char* e = getenv("DTG_PUSH_TMA"); push_tma = e ? atoi(e) : 1;
      e = getenv("DTG_STORE_TMA"); store_tma = e ? atoi(e) : 1;
    }
    const bool a16o = (reinterpret_cast<uintptr_t>(out) & 15) == 0;
    const bool want_red = (ep.mode == EPI_F32_RED_COAL && push_tma >= 1) ||
                          (ep.mode == EPI_F32_RED && push_tma >= 2);
    int rc2 = OK;
    if (want_red && !ep.fancy && (ldo % 4) == 0 && a16o) {
      rc2 = get_tensor_map(out, (uint64_t)N, (uint64_t)M, (uint64_t)ldo, 32, 32, &ep.tmO, 0, 0, 0, 0,
                           /*f32=*/true);
      ep.mode = EPI_F32_RED_TMA; ep.tma_out = TMAOUT_RED;
    } else if (store_tma >= 1 && ep.direct && out2 == nullptr && ep.mode == EPI_F32 && (ldo % 4) == 0) {
      rc2 = get_tensor_map(out, (uint64_t)N, (uint64_t)M, (uint64_t)ldo, 32, 32, &ep.tmO, 0, 0, 0, 0,
                           /*f32=*/true);
      ep.tma_out = store_tma >= 2 ? COALOUT_F32 : TMAOUT_F32;
    } else if (store_tma >= 1 && ep.direct && out2 == nullptr && ep.mode == EPI_BF16 && (ldo % 8) == 0) {
      rc2 = get_tensor_map(out, (uint64_t)N, (uint64_t)M, (uint64_t)ldo, 32, 32, &ep.tmO, 0, 0, 0, 0,
                           /*f32=*/false, /*sw64=*/true);
      ep.tma_out = store_tma >= 2 ? COALOUT_BF16 : TMAOUT_BF16;
    }
    if (rc2 != OK) return rc2;
  }

  CUtensorMap ta, tb;
  int rc;
  if (cv != nullptr && cv->which == 1) rc = get_im2col_map(*cv, BM, &ta);
  else if (!a_mn) rc = get_tensor_map(A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, BK, BM, &ta);
  else       rc = get_tensor_map(A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, 64, BK, &ta);
  if (rc != OK) return rc;
  if (cv != nullptr && cv->which == 2) rc = get_im2col_map(*cv, BK, &tb);
  else if (!b_mn) rc = get_tensor_map(B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, BK,
                                 (uint32_t)(two_cta ? BN / 2 : BN), &tb);
  else       rc = get_tensor_map(B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, 64, BK, &tb);
  if (rc != OK) return rc;

#define DTG_DISPATCH(BN_)                                                                       \
  if (!a_mn && !b_mn) return launch_gemm<BN_, false, false>(ta, tb, M, N, K, ep, g_num_sms, stream); \
  if (!a_mn && b_mn)  return launch_gemm<BN_, false, true>(ta, tb, M, N, K, ep, g_num_sms, stream);  \
  if (a_mn && b_mn)   return launch_gemm<BN_, true, true>(ta, tb, M, N, K, ep, g_num_sms, stream);   \
  return launch_gemm<BN_, true, false>(ta, tb, M, N, K, ep, g_num_sms, stream);
  if (two_cta && cv == nullptr && M >= 4 * BM && (force_tall || (g_use_tall && K >= g_tall_min_k))) {
    // 512 x 256 pair tiles (the SM's operand intake, not the tensor core, bounds the 256 x 256
    // pair tile), finished by double-buffered 256-row tiles
    const int pairs = g_num_sms / 2;
    plan_tall(M, N, pairs, g_tall_small_cost, ep);
    if (ep.tall_big_mblocks > 0) {
      if (!a_mn && !b_mn) return launch_gemm_tall<false, false>(ta, tb, M, N, K, ep, g_num_sms, stream);
      if (!a_mn && b_mn)  return launch_gemm_tall<false, true>(ta, tb, M, N, K, ep, g_num_sms, stream);
      if (a_mn && b_mn)   return launch_gemm_tall<true, true>(ta, tb, M, N, K, ep, g_num_sms, stream);
      return launch_gemm_tall<true, false>(ta, tb, M, N, K, ep, g_num_sms, stream);
    }
  }
  if (two_cta && g_2cta_stages == 5) {     // round-1 pipeline depth, kept for A/B runs
    if (!a_mn && !b_mn) return launch_gemm_2cta<256, false, false, 5>(ta, tb, M, N, K, ep, g_num_sms, stream);
    if (!a_mn && b_mn)  return launch_gemm_2cta<256, false, true, 5>(ta, tb, M, N, K, ep, g_num_sms, stream);
    if (a_mn && b_mn)   return launch_gemm_2cta<256, true, true, 5>(ta, tb, M, N, K, ep, g_num_sms, stream);
    return launch_gemm_2cta<256, true, false, 5>(ta, tb, M, N, K, ep, g_num_sms, stream);
  }
  if (two_cta) {
    if (!a_mn && !b_mn) return launch_gemm_2cta<256, false, false>(ta, tb, M, N, K, ep, g_num_sms, stream);
    if (!a_mn && b_mn)  return launch_gemm_2cta<256, false, true>(ta, tb, M, N, K, ep, g_num_sms, stream);
    if (a_mn && b_mn)   return launch_gemm_2cta<256, true, true>(ta, tb, M, N, K, ep, g_num_sms, stream);
    return launch_gemm_2cta<256, true, false>(ta, tb, M, N, K, ep, g_num_sms, stream);
  }
  if (BN == 256) { DTG_DISPATCH(256) }
  if (BN == 128) { DTG_DISPATCH(128) }
  DTG_DISPATCH(64)
#undef DTG_DISPATCH
}


extern "C" int dtg_gemm_bf16(const void* A, const void* B, int M, int N, int K, long long lda,
                             long long ldb, int a_mn, int b_mn, void* out, void* out2,
                             const float* bias, const void* mask, const void* residual,
                             long long ldo, long long ldmask, float alpha, int mode, int act,
                             int mask_mode, int force_bn, void* stream_) {
  return gemm_impl(A, B, M, N, K, lda, ldb, a_mn, b_mn, out, out2, bias, mask, residual, ldo, ldmask,
                   alpha, mode, act, mask_mode, force_bn, stream_, nullptr);
}

// Implicit-GEMM convolution (k x k, NHWC bf16, C % 64 == 0): the patch matrix is never written.
//   which = 1 (forward):  out[N*Ho*Wo, Cout] = im2col(x) . Wt[Cout, k*k*C]^T   (other = Wt, K-major)
//   which = 2 (wgrad):    out[Cout, k*k*C]  (+)= alpha * dY[N*Ho*Wo, Cout]^T . im2col(x)  (other = dY)
// The im2col operand is gathered by TMA in im2col mode straight into the swizzled stage
// (UTMALDG.4D.IM2COL); everything else -- epilogues, split-K, sinks -- is dtg_gemm_bf16's.
extern "C" int dtg_conv_gemm_bf16(const void* x, int cN, int cH, int cW, int cC, int k, int stride,
                                  int pad, int which, const void* other, long long ld_other,
                                  int Cout, void* out, long long ldo, const float* bias,
                                  float alpha, int mode, int force_bn, void* stream_) {
  if (x == nullptr || other == nullptr || out == nullptr || (which != 1 && which != 2) ||
      cC % 64 != 0 || k < 1 || stride < 1 || (reinterpret_cast<uintptr_t>(x) & 15) != 0) {
    dtg_set_last_error("dtg_conv_gemm_bf16", "need NHWC bf16 input with C % 64 == 0, which in {1, 2}");
    return ERR_BAD_ARG;
  }
  const int Ho = (cH + 2 * pad - k) / stride + 1, Wo = (cW + 2 * pad - k) / stride + 1;
  if (Ho < 1 || Wo < 1 || pad > 127 || k - 1 - pad > 127) {
    dtg_set_last_error("dtg_conv_gemm_bf16", "unsupported geometry");
    return ERR_BAD_ARG;
  }
  ConvDesc cv{x, cN, cH, cW, cC, k, stride, pad, which};
  const long long rows = (long long)cN * Ho * Wo;
  const int Kc = k * k * cC;
  if (rows > 0x7fffffffLL) { dtg_set_last_error("dtg_conv_gemm_bf16", "too many pixels"); return ERR_TOO_LARGE; }
  if (which == 1)     // A = im2col(x) [rows, Kc] (K-major), B = weights [Cout, Kc] (K-major)
    return gemm_impl(x, other, (int)rows, Cout, Kc, cC, ld_other, 0, 0, out, nullptr, bias, nullptr,
                     nullptr, ldo, 0, alpha, mode, ACT_NONE, MASK_NONE, force_bn, stream_, &cv);
  // A = dY [rows, Cout] used as dY^T (MN-major), B = im2col(x) [rows, Kc] (MN-major)
  return gemm_impl(other, x, Cout, Kc, (int)rows, ld_other, cC, 1, 1, out, nullptr, nullptr, nullptr,
                   nullptr, ldo, 0, alpha, mode, ACT_NONE, MASK_NONE, force_bn, stream_, &cv);
}

// Batched form: nb1 * nb2 problems of identical shape, operand/out batch strides in
// elements.  Plain alpha-scaled outputs only (modes 0..3); used for attention, where a
// problem is one (head, batch) pair and the strides walk the packed QKV / score buffers.
static int gemm_batched_impl(const void* A, const void* B, int M, int N, int K,
                                     long long lda, long long ldb, int a_mn, int b_mn, void* out,
                                     long long ldo, float alpha, int mode, int nb1, int nb2,
                                     long long a_bs1, long long a_bs2, long long b_bs1,
                                     long long b_bs2, long long o_bs1, long long o_bs2,
                                     int force_bn, void* stream_, const float* rowv1,
                                     const float* rowv2, float* stats, const void* pmat,
                                     float scale2, int causal) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if (M <= 0 || N <= 0 || K <= 0 || nb1 <= 0 || nb2 <= 0 || !A || !B || mode < 0 || mode > 7 ||
      (!out && mode != EPI_ATTN_STATS)) {
    dtg_set_last_error("dtg_gemm_bf16_batched", "bad shape, mode or null pointer");
    return ERR_BAD_ARG;
  }
  if ((lda % 8) || (ldb % 8) || (a_bs1 % 8) || (a_bs2 % 8) || (b_bs1 % 8) || (b_bs2 % 8) ||
      (reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15)) {
    dtg_set_last_error("dtg_gemm_bf16_batched",
                       "operand bases, row pitches and batch strides must be 16-byte multiples");
    return ERR_BAD_ARG;
  }
  if (g_num_sms == 0) {
    int dev = 0;
    DTG_CUDA_OK(cudaGetDevice(&dev));
    DTG_CUDA_OK(cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev));
  }
  if (mode >= EPI_ATTN_STATS) force_bn = 128;     // stats are laid out per 128-key tile
  int BN = force_bn ? force_bn : (N > 64 ? 128 : 64);
  if (BN != 64 && BN != 128) {
    dtg_set_last_error("dtg_gemm_bf16_batched", "force_bn must be 0, 64 or 128");
    return ERR_BAD_ARG;
  }
  GemmEpilogue ep;
  std::memset(&ep, 0, sizeof(ep));
  ep.out = out; ep.ldo = ldo; ep.alpha = alpha; ep.mode = mode;
  const bool bf_out = mode == EPI_BF16 || mode == EPI_ATTN_P || mode == EPI_ATTN_DS;
  const int esz = bf_out ? 2 : 4;
  const int align = bf_out ? 8 : 16;
  if (mode >= EPI_ATTN_STATS) {
    if ((mode == EPI_ATTN_STATS && !stats) || (mode == EPI_ATTN_P && (!rowv1 || !rowv2)) ||
        (mode == EPI_ATTN_DS && (!rowv1 || !pmat)) || (N % 8) != 0) {
      dtg_set_last_error("dtg_gemm_bf16_batched_ex", "attention mode is missing an operand");
      return ERR_BAD_ARG;
    }
  }
  ep.rowv1 = rowv1; ep.rowv2 = rowv2; ep.stats = stats; ep.scale2 = scale2; ep.causal = causal;
  ep.mask = reinterpret_cast<const __nv_bfloat16*>(pmat);
  ep.vec = ((ldo * esz) % align == 0 && (reinterpret_cast<uintptr_t>(out) % align) == 0 &&
            (o_bs1 * esz) % align == 0 && (o_bs2 * esz) % align == 0) ? 1 : 0;
  ep.nb1 = nb1; ep.nb2 = nb2; ep.o_bs1 = o_bs1; ep.o_bs2 = o_bs2;
  ep.direct = ((reinterpret_cast<uintptr_t>(out) & 15) == 0 && (ldo * esz) % 16 == 0 &&
               (o_bs1 * esz) % 16 == 0 && (o_bs2 * esz) % 16 == 0) ? 1 : 0;
  const uint32_t mn_lbo = 64 * BK * 2, mn_sbo = 1024, mn_kstep = (UMMA_K / 8) * 1024;
  ep.a_lbo = a_mn ? mn_lbo : 16; ep.a_sbo = a_mn ? mn_sbo : 1024;
  ep.a_kstep = a_mn ? mn_kstep : UMMA_K * 2;
  ep.b_lbo = b_mn ? mn_lbo : 16; ep.b_sbo = b_mn ? mn_sbo : 1024;
  ep.b_kstep = b_mn ? mn_kstep : UMMA_K * 2;
  CUtensorMap ta, tb;
  int rc;
  if (!a_mn) rc = get_tensor_map(A, K, M, lda, BK, BM, &ta, nb1, nb2, a_bs1, a_bs2);
  else       rc = get_tensor_map(A, M, K, lda, 64, BK, &ta, nb1, nb2, a_bs1, a_bs2);
  if (rc != OK) return rc;
  if (!b_mn) rc = get_tensor_map(B, K, N, ldb, BK, (uint32_t)BN, &tb, nb1, nb2, b_bs1, b_bs2);
  else       rc = get_tensor_map(B, N, K, ldb, 64, BK, &tb, nb1, nb2, b_bs1, b_bs2);
  if (rc != OK) return rc;
#define DTG_DISPATCH_B(BN_)                                                                          \
  if (!a_mn && !b_mn) return launch_gemm<BN_, false, false, true>(ta, tb, M, N, K, ep, g_num_sms, stream); \
  if (!a_mn && b_mn)  return launch_gemm<BN_, false, true, true>(ta, tb, M, N, K, ep, g_num_sms, stream);  \
  if (a_mn && b_mn)   return launch_gemm<BN_, true, true, true>(ta, tb, M, N, K, ep, g_num_sms, stream);   \
  return launch_gemm<BN_, true, false, true>(ta, tb, M, N, K, ep, g_num_sms, stream);
  if (BN == 128) { DTG_DISPATCH_B(128) }
  DTG_DISPATCH_B(64)
#undef DTG_DISPATCH_B
}

extern "C" int dtg_gemm_bf16_batched(const void* A, const void* B, int M, int N, int K,
                                     long long lda, long long ldb, int a_mn, int b_mn, void* out,
                                     long long ldo, float alpha, int mode, int nb1, int nb2,
                                     long long a_bs1, long long a_bs2, long long b_bs1,
                                     long long b_bs2, long long o_bs1, long long o_bs2,
                                     int force_bn, void* stream) {
  if (mode > 4) {
    dtg_set_last_error("dtg_gemm_bf16_batched", "modes 5..7 need dtg_gemm_bf16_batched_ex");
    return ERR_BAD_ARG;
  }
  return gemm_batched_impl(A, B, M, N, K, lda, ldb, a_mn, b_mn, out, ldo, alpha, mode, nb1, nb2,
                           a_bs1, a_bs2, b_bs1, b_bs2, o_bs1, o_bs2, force_bn, stream, nullptr,
                           nullptr, nullptr, nullptr, 0.f, 0);
}

// Batched GEMM with the attention epilogues (modes 5..7) and causal tile skipping.
extern "C" int dtg_gemm_bf16_batched_ex(const void* A, const void* B, int M, int N, int K,
                                        long long lda, long long ldb, int a_mn, int b_mn,
                                        void* out, long long ldo, float alpha, int mode, int nb1,
                                        int nb2, long long a_bs1, long long a_bs2, long long b_bs1,
                                        long long b_bs2, long long o_bs1, long long o_bs2,
                                        const float* rowv1, const float* rowv2, float* stats,
                                        const void* pmat, float scale2, int causal, void* stream) {
  return gemm_batched_impl(A, B, M, N, K, lda, ldb, a_mn, b_mn, out, ldo, alpha, mode, nb1, nb2,
                           a_bs1, a_bs2, b_bs1, b_bs2, o_bs1, o_bs2, 0, stream, rowv1, rowv2,
                           stats, pmat, scale2, causal);
}

extern "C" void dtg_gemm_clear_tensor_map_cache() {
  std::lock_guard<std::mutex> g(g_map_mu);
  g_maps.clear();
}

// Host replay of the tall kernel's static schedule: plan_tall() for (M, N, clusters), then every
// cluster walks the same TallSched / tile_coords the device does.  Checks that each 256 x 256
// block of the output is produced exactly once, that tile indices stay in range and that the two
// accumulator halves alternate on small tiles (the MMA / epilogue hand-off relies on it).
// info[0..5] = big m-blocks, big tiles, small tiles, extra rounds, clusters used, 1000 x the
// slowest cluster's load in big-tile units.  Returns OK, or ERR_BAD_ARG with the reason in
// dtg_last_error().  Runs without a GPU (tests/test_gemm_schedule.py).
extern "C" int dtg_tall_schedule_check(int M, int N, int clusters, float small_cost, int group,
                                       int* info) {
  if (M <= 0 || N <= 0 || clusters <= 0 || group <= 0) {
    dtg_set_last_error("dtg_tall_schedule_check", "non-positive argument");
    return ERR_BAD_ARG;
  }
  GemmEpilogue ep{};
  ep.group = group;
  plan_tall(M, N, clusters, small_cost, ep);
  const int tn = (N + 255) / 256;
  const int rows256 = (M + 255) / 256;
  const int big_rows = ep.tall_big_mblocks * 512;
  const int small_m = (M - big_rows + 255) / 256;
  const int tiles = ep.tall_nbig + ep.tall_nsmall;
  const int C = clusters < tiles ? clusters : tiles;
  std::vector<int> cover((size_t)rows256 * tn, 0);
  float worst = 0.f;
  char why[160];
  for (int c = 0; c < C; ++c) {
    TallSched sch(ep, c, C);
    int kind, idx, half, smalls = 0, bigs = 0;
    while (sch.next(kind, idx, half)) {
      int mb, nb;
      if (kind) {
        if (idx < 0 || idx >= ep.tall_nbig || smalls > 0) {
          snprintf(why, sizeof(why), "cluster %d: big tile %d out of range or after a small one", c, idx);
          dtg_set_last_error("dtg_tall_schedule_check", why);
          return ERR_BAD_ARG;
        }
        tile_coords(idx, ep.tall_big_mblocks, tn, group > 1 ? group / 2 : 1, mb, nb);
        if (mb < 0 || mb >= ep.tall_big_mblocks || nb < 0 || nb >= tn) {
          snprintf(why, sizeof(why), "big tile %d -> block (%d, %d) outside the grid", idx, mb, nb);
          dtg_set_last_error("dtg_tall_schedule_check", why);
          return ERR_BAD_ARG;
        }
        ++cover[(size_t)(2 * mb) * tn + nb];
        ++cover[(size_t)(2 * mb + 1) * tn + nb];
        ++bigs;
      } else {
        if (idx < 0 || idx >= ep.tall_nsmall || half != (smalls & 1)) {
          snprintf(why, sizeof(why), "cluster %d: small tile %d out of range or half %d out of turn", c, idx, half);
          dtg_set_last_error("dtg_tall_schedule_check", why);
          return ERR_BAD_ARG;
        }
        tile_coords(idx, small_m, tn, group, mb, nb);
        if (mb < 0 || mb >= small_m || nb < 0 || nb >= tn) {
          snprintf(why, sizeof(why), "small tile %d -> block (%d, %d) outside the grid", idx, mb, nb);
          dtg_set_last_error("dtg_tall_schedule_check", why);
          return ERR_BAD_ARG;
        }
        ++cover[(size_t)(big_rows / 256 + mb) * tn + nb];
        ++smalls;
      }
    }
    const float load = (float)bigs + small_cost * (float)smalls;
    worst = load > worst ? load : worst;
  }
  for (size_t i = 0; i < cover.size(); ++i)
    if (cover[i] != 1) {
      snprintf(why, sizeof(why), "output block (%d, %d) produced %d times", (int)(i / tn), (int)(i % tn), cover[i]);
      dtg_set_last_error("dtg_tall_schedule_check", why);
      return ERR_BAD_ARG;
    }
  if (info != nullptr) {
    info[0] = ep.tall_big_mblocks; info[1] = ep.tall_nbig; info[2] = ep.tall_nsmall;
    info[3] = ep.tall_extra; info[4] = C; info[5] = (int)(worst * 1000.f + 0.5f);
  }
  return OK;
}

// The grouped tile order every persistent GEMM kernel uses (tile_coords): a bijection between
// [0, tiles_m * tiles_n) and the (m-block, n-block) grid for any group size.  Host check.
extern "C" int dtg_tile_order_check(int tiles_m, int tiles_n, int group) {
  if (tiles_m <= 0 || tiles_n <= 0 || group <= 0) {
    dtg_set_last_error("dtg_tile_order_check", "non-positive argument");
    return ERR_BAD_ARG;
  }
  std::vector<char> seen((size_t)tiles_m * tiles_n, 0);
  for (int t = 0; t < tiles_m * tiles_n; ++t) {
    int mb, nb;
    tile_coords(t, tiles_m, tiles_n, group, mb, nb);
    if (mb < 0 || mb >= tiles_m || nb < 0 || nb >= tiles_n || seen[(size_t)mb * tiles_n + nb]) {
      char why[128];
      snprintf(why, sizeof(why), "tile %d -> (%d, %d): outside the grid or visited twice", t, mb, nb);
      dtg_set_last_error("dtg_tile_order_check", why);
      return ERR_BAD_ARG;
    }
    seen[(size_t)mb * tiles_n + nb] = 1;
  }
  return OK;
}
