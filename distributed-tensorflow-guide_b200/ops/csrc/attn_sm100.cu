// Fused attention forward for sm_100a (head_dim 64): one persistent kernel computes, for a
// block of 128 queries of one (batch, head),
//   phase A:  S_j = Q K_j^T  -> exact row max m and sum-exp l   (tcgen05 -> TMEM -> registers)
//   phase B:  S_j again, P_j = exp2(S_j*c - m)/l in bf16 -> shared memory (the 128B-swizzled
//             K-major layout tcgen05 reads) -> O += P_j V_j in TMEM, and P_j -> HBM by TMA store
//             for the backward pass.
// Scores never exist in global memory and no accumulator is ever rescaled (two exact passes
// over the keys instead of an online softmax; the second QK^T is ~free on the tensor core).
//
// Warp roles (352 threads): warps 0-7 softmax (warp w owns TMEM lane quadrant w%4 and the
// 64-column half w/4 of each 128x128 score tile), warp 8 TMA producer, warp 9 MMA issuer,
// warp 10 P storer.  Pipelines: Q x2, K/V ring x3, S (TMEM) x2, P (smem) x2, O (TMEM) x2.
//
// Replaces (SURVEY 2.5 / section 8 W1-W2 for the transformer family): the batched score
// GEMM, the softmax kernel and the PV GEMM of the unfused path in models/transformer_native.
#include "common.cuh"

namespace dtg {

constexpr int AT_THREADS = 352;
constexpr int AT_ST = 3;                     // K/V ring depth
constexpr int AT_TILE_BYTES = 128 * 64 * 2;  // one 128 x 64 bf16 operand tile
constexpr int AT_P_BYTES = 128 * 128 * 2;
constexpr uint32_t AT_TMEM_COLS = 512;
constexpr uint32_t AT_TMEM_O = 256;          // S: columns [0,256), O: [256,384)

struct AttnParams {
  int B, H, T, causal, n_t, items, store_p;
  float scale2;                 // softmax scale * log2(e)
  float* rowm;                  // [B*H*T]  row max (log2 domain)
  float* rowinvl;               // [B*H*T]  1 / sum exp
  __nv_bfloat16* ctx;           // [B*T, ctx_ld]; head h occupies columns [h*64, h*64+64)
  long long ctx_ld;
};

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;\n" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void at_tma_load(void* smem_dst, const CUtensorMap* m, uint64_t* bar,
                                            int32_t c0, int32_t c1, int32_t c2, int32_t c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];\n" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void at_tma_store(const CUtensorMap* m, const void* smem_src, int32_t c0,
                                             int32_t c1, int32_t c2, int32_t c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];\n" ::"l"(
          reinterpret_cast<uint64_t>(m)),
      "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;\n" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() {
  asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory");
}
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory"); }
__device__ __forceinline__ void named_bar_sync(int id, int n) {
  asm volatile("bar.sync %0, %1;\n" ::"r"(id), "r"(n) : "memory");
}

struct AttnItem {
  int qt, b, h, nj;
};
// Work item order: heaviest query tiles first (causal: tile qt needs qt+1 key tiles).
__device__ __forceinline__ AttnItem attn_item(const AttnParams& p, int it) {
  AttnItem w;
  const int bh = p.B * p.H;
  w.qt = p.n_t - 1 - it / bh;
  const int r = it % bh;
  w.b = r / p.H;
  w.h = r % p.H;
  w.nj = p.causal ? w.qt + 1 : p.n_t;
  return w;
}

__global__ void __launch_bounds__(AT_THREADS, 1)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmP,
                const __grid_constant__ AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* sQ = smem;                                   // 2 x 16 KB
  uint8_t* sK = sQ + 2 * AT_TILE_BYTES;                 // ring, K-major [128 keys][64]
  uint8_t* sV = sK + AT_ST * AT_TILE_BYTES;             // ring, [128 keys][64] = MN-major B
  uint8_t* sP = sV + AT_ST * AT_TILE_BYTES;             // 2 x (two [128][64] K-major panels)
  float* sStat = reinterpret_cast<float*>(sP + 2 * AT_P_BYTES);   // [2 items][2 halves][128][2]
  float* sSum = sStat + 2 * 2 * 128 * 2;                          // [2 items][2 halves][128]
  uint64_t* q_full = reinterpret_cast<uint64_t*>(sSum + 2 * 2 * 128);
  uint64_t* q_empty = q_full + 2;
  uint64_t* kv_full = q_empty + 2;
  uint64_t* kv_empty = kv_full + AT_ST;
  uint64_t* s_full = kv_empty + AT_ST;
  uint64_t* s_empty = s_full + 2;
  uint64_t* p_full = s_empty + 2;
  uint64_t* p_empty = p_full + 2;
  uint64_t* o_full = p_empty + 2;
  uint64_t* o_empty = o_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    if (p.store_p) tma_prefetch_desc(&tmP);
  }
  if (warp == 9 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&q_full[i], 1);
      mbar_init(&q_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&s_empty[i], 8);
      mbar_init(&p_full[i], 8);
      mbar_init(&p_empty[i], 1 + (p.store_p ? 1 : 0));
      mbar_init(&o_full[i], 1);
      mbar_init(&o_empty[i], 8);
    }
    for (int i = 0; i < AT_ST; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(tmem_slot, AT_TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

  if (warp == 8) {
    // ============================ TMA producer ============================
    if (lane == 0) {
      uint32_t ic = 0, kc = 0;
      for (int it = blockIdx.x; it < p.items; it += gridDim.x, ++ic) {
        const AttnItem w = attn_item(p, it);
        const int qb = ic & 1;
        mbar_wait(&q_empty[qb], ((ic >> 1) & 1) ^ 1);
        mbar_arrive_expect_tx(&q_full[qb], AT_TILE_BYTES);
        at_tma_load(sQ + qb * AT_TILE_BYTES, &tmQ, &q_full[qb], 0, w.qt * 128, w.h, w.b);
        for (int phase = 0; phase < 2; ++phase) {
          for (int j = 0; j < w.nj; ++j, ++kc) {
            const int st = kc % AT_ST;
            mbar_wait(&kv_empty[st], ((kc / AT_ST) & 1) ^ 1);
            mbar_arrive_expect_tx(&kv_full[st], phase ? 2 * AT_TILE_BYTES : AT_TILE_BYTES);
            at_tma_load(sK + st * AT_TILE_BYTES, &tmK, &kv_full[st], 0, j * 128, w.h, w.b);
            if (phase) at_tma_load(sV + st * AT_TILE_BYTES, &tmV, &kv_full[st], 0, j * 128, w.h, w.b);
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 9) {
    // ============================ MMA issuer ============================
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(128, 128, false, false);
      constexpr uint32_t idesc_o = make_idesc_bf16(128, 64, false, true);
      uint32_t ic = 0, kc = 0, sc = 0, pc = 0;
      auto issue_qk = [&](int qb, int st, int sb) {
        const uint32_t a = smem_u32(sQ + qb * AT_TILE_BYTES), b = smem_u32(sK + st * AT_TILE_BYTES);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16_ss(tmem_base + sb * 128, make_smem_desc_sw128(a + k * 32, 16, 1024),
                       make_smem_desc_sw128(b + k * 32, 16, 1024), idesc_s, k ? 1u : 0u);
      };
      for (int it = blockIdx.x; it < p.items; it += gridDim.x, ++ic) {
        const AttnItem w = attn_item(p, it);
        const int qb = ic & 1, ob = ic & 1;
        const uint32_t iph = (ic >> 1) & 1;
        mbar_wait(&q_full[qb], iph);
        for (int j = 0; j < w.nj; ++j, ++kc, ++sc) {            // phase A: scores only
          const int st = kc % AT_ST, sb = sc & 1;
          mbar_wait(&kv_full[st], (kc / AT_ST) & 1);
          mbar_wait(&s_empty[sb], ((sc >> 1) & 1) ^ 1);
          tc_fence_after();
          issue_qk(qb, st, sb);
          umma_commit(&s_full[sb]);
          umma_commit(&kv_empty[st]);
        }
        for (int j = 0; j <= w.nj; ++j) {                       // phase B, PV one tile behind QK
          if (j < w.nj) {
            const uint32_t kj = kc + j;
            const int st = kj % AT_ST, sb = sc & 1;
            mbar_wait(&kv_full[st], (kj / AT_ST) & 1);
            mbar_wait(&s_empty[sb], ((sc >> 1) & 1) ^ 1);
            tc_fence_after();
            issue_qk(qb, st, sb);
            umma_commit(&s_full[sb]);
            if (j == w.nj - 1) umma_commit(&q_empty[qb]);
            ++sc;
          }
          if (j >= 1) {
            const int jj = j - 1;
            const uint32_t kj = kc + jj;
            const int st = kj % AT_ST, pb = pc & 1;
            if (jj == 0) mbar_wait(&o_empty[ob], iph ^ 1);
            mbar_wait(&p_full[pb], (pc >> 1) & 1);
            tc_fence_after();
            const uint32_t a = smem_u32(sP + pb * AT_P_BYTES), b = smem_u32(sV + st * AT_TILE_BYTES);
#pragma unroll
            for (int ks = 0; ks < 8; ++ks)
              umma_bf16_ss(tmem_base + AT_TMEM_O + ob * 64,
                           make_smem_desc_sw128(a + (ks >> 2) * AT_TILE_BYTES + (ks & 3) * 32, 16, 1024),
                           make_smem_desc_sw128(b + ks * 2048, AT_TILE_BYTES, 1024), idesc_o,
                           (jj | ks) ? 1u : 0u);
            umma_commit(&p_empty[pb]);
            umma_commit(&kv_empty[st]);
            if (jj == w.nj - 1) umma_commit(&o_full[ob]);
            ++pc;
          }
        }
        kc += w.nj;
      }
    }
    __syncwarp();
  } else if (warp == 10) {
    // ============================ P storer ============================
    if (lane == 0 && p.store_p) {
      uint32_t pc = 0;
      for (int it = blockIdx.x; it < p.items; it += gridDim.x) {
        const AttnItem w = attn_item(p, it);
        for (int j = 0; j < w.nj; ++j, ++pc) {
          const int pb = pc & 1;
          mbar_wait(&p_full[pb], (pc >> 1) & 1);
          at_tma_store(&tmP, sP + pb * AT_P_BYTES, j * 128, w.qt * 128, w.h, w.b);
          if (j * 128 + 64 < p.T)
            at_tma_store(&tmP, sP + pb * AT_P_BYTES + AT_TILE_BYTES, j * 128 + 64, w.qt * 128, w.h, w.b);
          bulk_commit();
          bulk_wait_read0();
          mbar_arrive(&p_empty[pb]);
        }
      }
      bulk_wait0();
    }
    __syncwarp();
  } else {
    // ============================ softmax warps 0..7 ============================
    const int q = warp & 3, hf = warp >> 2;
    const int rt = q * 32 + lane;                        // query row inside the tile
    const uint32_t tlane = uint32_t(q * 32) << 16;
    const float c2 = p.scale2;
    uint32_t ic = 0, sc = 0, pc = 0;
    for (int it = blockIdx.x; it < p.items; it += gridDim.x, ++ic) {
      const AttnItem w = attn_item(p, it);
      const int row = w.qt * 128 + rt;
      const bool row_ok = row < p.T;
      const long long grow = ((long long)w.b * p.H + w.h) * p.T + row;
      const int climit = p.causal ? min(row, p.T - 1) : p.T - 1;   // last visible key
      uint32_t r[64];
      auto load_scores = [&](int j) {
        const int sb = sc & 1;
        mbar_wait(&s_full[sb], (sc >> 1) & 1);
        tc_fence_after();
        const uint32_t t = tmem_base + tlane + sb * 128 + hf * 64;
        tmem_ld_32x32(t, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
        tmem_ld_32x32(t + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&s_empty[sb]);      // TMEM tile is in registers now
        ++sc;
        const int col0 = j * 128 + hf * 64;
        if ((p.causal && j == w.qt) || col0 + 64 > p.T) {
#pragma unroll
          for (int c = 0; c < 64; ++c)
            if (col0 + c > climit) r[c] = 0xff800000u;   // -inf: exp2 gives an exact 0
        }
      };
      // ---------------- phase A: exact row statistics ----------------
      // P must be normalised before it is stored (store_p); otherwise phase A finds the max
      // only, phase B sums the exponentials it computes anyway and O is scaled by 1/l at the
      // end -- half the MUFU work.
      const bool norm_first = p.store_p != 0;
      float m = -INFINITY, l = 0.f;
      for (int j = 0; j < w.nj; ++j) {
        load_scores(j);
        float cm = __uint_as_float(r[0]);
#pragma unroll
        for (int c = 1; c < 64; ++c) cm = fmaxf(cm, __uint_as_float(r[c]));
        const float mn = fmaxf(m, cm * c2);
        if (!norm_first) {
          m = mn;                           // max only: the sum is accumulated in phase B
        } else if (mn > -INFINITY) {
          float a0 = 0.f, a1 = 0.f;
#pragma unroll
          for (int c = 0; c < 64; c += 2) {
            a0 += ex2_approx(fmaf(__uint_as_float(r[c]), c2, -mn));
            a1 += ex2_approx(fmaf(__uint_as_float(r[c + 1]), c2, -mn));
          }
          l = l * ex2_approx(m - mn) + (a0 + a1);
          m = mn;
        }
      }
      float* ss = sStat + (ic & 1) * 512;
      ss[(hf * 128 + rt) * 2] = m;
      ss[(hf * 128 + rt) * 2 + 1] = l;
      named_bar_sync(1, 256);
      const float m0 = ss[rt * 2], l0 = ss[rt * 2 + 1];
      const float m1 = ss[(128 + rt) * 2], l1 = ss[(128 + rt) * 2 + 1];
      const float mt = fmaxf(m0, m1);
      float lt = 0.f;
      if (mt > -INFINITY) lt = l0 * ex2_approx(m0 - mt) + l1 * ex2_approx(m1 - mt);
      float il = lt > 0.f ? 1.f / lt : 0.f;
      const float mneg = mt > -INFINITY ? -mt : 0.f;
      float lsum0 = 0.f, lsum1 = 0.f;
      // ---------------- phase B: P -> smem (-> PV MMA, -> HBM) ----------------
      for (int j = 0; j < w.nj; ++j) {
        load_scores(j);
        uint32_t pk[32];
        if (norm_first) {
#pragma unroll
          for (int c = 0; c < 32; ++c)
            pk[c] = pack_bf16x2(ex2_approx(fmaf(__uint_as_float(r[2 * c]), c2, mneg)) * il,
                                ex2_approx(fmaf(__uint_as_float(r[2 * c + 1]), c2, mneg)) * il);
        } else {
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            const float e0 = ex2_approx(fmaf(__uint_as_float(r[2 * c]), c2, mneg));
            const float e1 = ex2_approx(fmaf(__uint_as_float(r[2 * c + 1]), c2, mneg));
            lsum0 += e0;
            lsum1 += e1;
            pk[c] = pack_bf16x2(e0, e1);
          }
        }
        const int pb = pc & 1;
        mbar_wait(&p_empty[pb], ((pc >> 1) & 1) ^ 1);
        uint8_t* prow = sP + pb * AT_P_BYTES + hf * AT_TILE_BYTES + rt * 128;
#pragma unroll
        for (int ch = 0; ch < 8; ++ch)
          *reinterpret_cast<uint4*>(prow + ((ch ^ (rt & 7)) << 4)) =
              make_uint4(pk[4 * ch], pk[4 * ch + 1], pk[4 * ch + 2], pk[4 * ch + 3]);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[pb]);
        ++pc;
      }
      float oscale = 1.f;
      if (!norm_first) {
        float* sl = sSum + (ic & 1) * 256;
        sl[hf * 128 + rt] = lsum0 + lsum1;
        named_bar_sync(1, 256);
        const float ltot = sl[rt] + sl[128 + rt];
        il = ltot > 0.f ? 1.f / ltot : 0.f;
        oscale = il;
      }
      if (hf == 0 && row_ok) {
        p.rowm[grow] = mt;
        p.rowinvl[grow] = il;
      }
      // ---------------- O epilogue ----------------
      const int ob = ic & 1;
      mbar_wait(&o_full[ob], (ic >> 1) & 1);
      tc_fence_after();
      uint32_t o[32];
      tmem_ld_32x32(tmem_base + tlane + AT_TMEM_O + ob * 64 + hf * 32, o);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&o_empty[ob]);
      if (row_ok) {
        uint4* dst = reinterpret_cast<uint4*>(p.ctx + ((long long)w.b * p.T + row) * p.ctx_ld +
                                              w.h * 64 + hf * 32);
#pragma unroll
        for (int g = 0; g < 4; ++g)
          dst[g] = make_uint4(
              pack_bf16x2(__uint_as_float(o[8 * g]) * oscale, __uint_as_float(o[8 * g + 1]) * oscale),
              pack_bf16x2(__uint_as_float(o[8 * g + 2]) * oscale, __uint_as_float(o[8 * g + 3]) * oscale),
              pack_bf16x2(__uint_as_float(o[8 * g + 4]) * oscale, __uint_as_float(o[8 * g + 5]) * oscale),
              pack_bf16x2(__uint_as_float(o[8 * g + 6]) * oscale, __uint_as_float(o[8 * g + 7]) * oscale));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, AT_TMEM_COLS);
}

constexpr size_t AT_SMEM = 1024 + 2 * AT_TILE_BYTES + 2 * AT_ST * AT_TILE_BYTES + 2 * AT_P_BYTES +
                           2 * 2 * 128 * 2 * 4 + 2 * 2 * 128 * 4 + 256;
}  // namespace dtg

// q, k, v: [B, T, H, 64] views with row pitch `ld`, head stride `hs`, batch stride `bs` (elements).
// P (optional, may be null): [B, H, T, T] bf16 probabilities for the backward pass.
extern "C" int dtg_attn_fwd(const void* q, const void* k, const void* v, long long ld, long long hs,
                            long long bs, int B, int H, int T, int hd, void* P, void* ctx,
                            long long ctx_ld, float* rowm, float* rowinvl, float scale, int causal,
                            void* stream_) {
  using namespace dtg;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if (hd != 64) {
    dtg_set_last_error("dtg_attn_fwd", "fused attention supports head_dim 64 only");
    return ERR_UNSUPPORTED;
  }
  if (!q || !k || !v || !ctx || !rowm || !rowinvl || B <= 0 || H <= 0 || T <= 0 || (ld % 8) ||
      (hs % 8) || (bs % 8) || (ctx_ld % 8) || (reinterpret_cast<uintptr_t>(ctx) & 15) ||
      (P && (T % 8))) {
    dtg_set_last_error("dtg_attn_fwd", "null pointer, bad shape or unaligned pitch");
    return ERR_BAD_ARG;
  }
  CUtensorMap tq, tk, tv, tp;
  int rc;
  if ((rc = tensor_map_bf16(q, 64, T, ld, 64, 128, &tq, H, B, hs, bs)) != OK) return rc;
  if ((rc = tensor_map_bf16(k, 64, T, ld, 64, 128, &tk, H, B, hs, bs)) != OK) return rc;
  if ((rc = tensor_map_bf16(v, 64, T, ld, 64, 128, &tv, H, B, hs, bs)) != OK) return rc;
  tp = tq;
  if (P && (rc = tensor_map_bf16(P, T, T, T, 64, 128, &tp, H, B, (uint64_t)T * T,
                                 (uint64_t)H * T * T)) != OK)
    return rc;
  AttnParams p;
  p.B = B; p.H = H; p.T = T; p.causal = causal ? 1 : 0;
  p.n_t = (T + 127) / 128;
  p.items = B * H * p.n_t;
  p.store_p = P ? 1 : 0;
  p.scale2 = scale * 1.4426950408889634f;
  p.rowm = rowm; p.rowinvl = rowinvl;
  p.ctx = reinterpret_cast<__nv_bfloat16*>(ctx); p.ctx_ld = ctx_ld;
  static bool attr_set = false;
  if (!attr_set) {
    DTG_CUDA_OK(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)AT_SMEM));
    attr_set = true;
  }
  const int sms = num_sms();
  const int grid = p.items < sms ? p.items : sms;
  attn_fwd_kernel<<<grid, AT_THREADS, AT_SMEM, stream>>>(tq, tk, tv, tp, p);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

// =====================================================================================
// Backward.  One CTA owns a block of 128 keys (K_j, V_j resident in smem, dK_j / dV_j in
// TMEM) and walks the query blocks i that see it.  Per (i, j) pair:
//   S = Q_i K_j^T, dP = dO_i V_j^T            (tcgen05 -> TMEM)
//   P = exp2(S*c - m_i) / l_i  (recomputed from the saved row statistics; P is never read
//   from HBM), dS = scale * P * (dP - D_i)    (softmax warps, registers)
//   P, dS -> smem (bf16, 128B-swizzled) -> dV_j += P^T dO_i, dK_j += dS^T Q_i (MN-major A)
//   dS -> HBM by TMA store: dQ = dS K is the batched GEMM that follows (it sums over j).
// =====================================================================================
namespace dtg {

constexpr int AB_RING = 2;
constexpr uint32_t AB_TM_DP = 256, AB_TM_DV = 384, AB_TM_DK = 448;

struct AttnBwdParams {
  int B, H, T, causal, n_t, items;
  float scale2, scale;
  const float* rowm;
  const float* rowinvl;
  const float* D;
  __nv_bfloat16* dk;
  __nv_bfloat16* dv;
  long long ld, hs, bs;          // element strides of the dk / dv views [B, T, H, 64]
};
struct AttnBwdItem {
  int j, b, h, i0, ni;
};
__device__ __forceinline__ AttnBwdItem attn_bwd_item(const AttnBwdParams& p, int it) {
  AttnBwdItem w;
  const int bh = p.B * p.H;
  w.j = it / bh;                                   // causal: low j = most query blocks first
  const int r = it % bh;
  w.b = r / p.H;
  w.h = r % p.H;
  w.i0 = p.causal ? w.j : 0;
  w.ni = p.n_t - w.i0;
  return w;
}

__global__ void __launch_bounds__(AT_THREADS, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                const __grid_constant__ CUtensorMap tmdS, const __grid_constant__ AttnBwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* sK = smem;                                   // 2 x 16 KB (per item, double-buffered)
  uint8_t* sV = sK + 2 * AT_TILE_BYTES;
  uint8_t* sQ = sV + 2 * AT_TILE_BYTES;                 // ring
  uint8_t* sdO = sQ + AB_RING * AT_TILE_BYTES;
  uint8_t* sPm = sdO + AB_RING * AT_TILE_BYTES;         // 32 KB, [128 q][128 keys] two panels
  uint8_t* sdS = sPm + AT_P_BYTES;
  uint64_t* kv_full = reinterpret_cast<uint64_t*>(sdS + AT_P_BYTES);
  uint64_t* kv_empty = kv_full + 2;
  uint64_t* qd_full = kv_empty + 2;
  uint64_t* qd_empty = qd_full + AB_RING;
  uint64_t* s_full = qd_empty + AB_RING;
  uint64_t* s_empty = s_full + 2;
  uint64_t* dp_full = s_empty + 2;
  uint64_t* dp_empty = dp_full + 1;
  uint64_t* pds_full = dp_empty + 1;
  uint64_t* pds_empty = pds_full + 1;
  uint64_t* acc_full = pds_empty + 1;
  uint64_t* acc_empty = acc_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmdO);
    tma_prefetch_desc(&tmdS);
  }
  if (warp == 9 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&s_empty[i], 8);
    }
    for (int i = 0; i < AB_RING; ++i) {
      mbar_init(&qd_full[i], 1);
      mbar_init(&qd_empty[i], 1);
    }
    mbar_init(dp_full, 1);
    mbar_init(dp_empty, 8);
    mbar_init(pds_full, 8);
    mbar_init(pds_empty, 2);          // MMA commit + dS storer
    mbar_init(acc_full, 1);
    mbar_init(acc_empty, 8);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(tmem_slot, AT_TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

  if (warp == 8) {
    // ============================ TMA producer ============================
    if (lane == 0) {
      uint32_t ic = 0, rc = 0;
      for (int it = blockIdx.x; it < p.items; it += gridDim.x, ++ic) {
        const AttnBwdItem w = attn_bwd_item(p, it);
        const int kvb = ic & 1;
        mbar_wait(&kv_empty[kvb], ((ic >> 1) & 1) ^ 1);
        mbar_arrive_expect_tx(&kv_full[kvb], 2 * AT_TILE_BYTES);
        at_tma_load(sK + kvb * AT_TILE_BYTES, &tmK, &kv_full[kvb], 0, w.j * 128, w.h, w.b);
        at_tma_load(sV + kvb * AT_TILE_BYTES, &tmV, &kv_full[kvb], 0, w.j * 128, w.h, w.b);
        for (int x = 0; x < w.ni; ++x, ++rc) {
          const int st = rc % AB_RING;
          mbar_wait(&qd_empty[st], ((rc / AB_RING) & 1) ^ 1);
          mbar_arrive_expect_tx(&qd_full[st], 2 * AT_TILE_BYTES);
          at_tma_load(sQ + st * AT_TILE_BYTES, &tmQ, &qd_full[st], 0, (w.i0 + x) * 128, w.h, w.b);
          at_tma_load(sdO + st * AT_TILE_BYTES, &tmdO, &qd_full[st], 0, (w.i0 + x) * 128, w.h, w.b);
        }
      }
    }
    __syncwarp();
  } else if (warp == 9) {
    // ============================ MMA issuer ============================
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(128, 128, false, false);
      constexpr uint32_t idesc_g = make_idesc_bf16(128, 64, true, true);
      uint32_t ic = 0, rc = 0, sc = 0, dc = 0, pc = 0;
      for (int it = blockIdx.x; it < p.items; it += gridDim.x, ++ic) {
        const AttnBwdItem w = attn_bwd_item(p, it);
        const int kvb = ic & 1;
        mbar_wait(&kv_full[kvb], (ic >> 1) & 1);
        const uint32_t kaddr = smem_u32(sK + kvb * AT_TILE_BYTES);
        const uint32_t vaddr = smem_u32(sV + kvb * AT_TILE_BYTES);
        for (int x = 0; x <= w.ni; ++x) {
          if (x < w.ni) {
            const uint32_t rx = rc + x;
            const int st = rx % AB_RING, sb = sc & 1;
            mbar_wait(&qd_full[st], (rx / AB_RING) & 1);
            const uint32_t qaddr = smem_u32(sQ + st * AT_TILE_BYTES);
            const uint32_t oaddr = smem_u32(sdO + st * AT_TILE_BYTES);
            mbar_wait(&s_empty[sb], ((sc >> 1) & 1) ^ 1);
            tc_fence_after();
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_bf16_ss(tmem_base + sb * 128, make_smem_desc_sw128(qaddr + k * 32, 16, 1024),
                           make_smem_desc_sw128(kaddr + k * 32, 16, 1024), idesc_s, k ? 1u : 0u);
            umma_commit(&s_full[sb]);
            ++sc;
            mbar_wait(dp_empty, (dc & 1) ^ 1);
            tc_fence_after();
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_bf16_ss(tmem_base + AB_TM_DP, make_smem_desc_sw128(oaddr + k * 32, 16, 1024),
                           make_smem_desc_sw128(vaddr + k * 32, 16, 1024), idesc_s, k ? 1u : 0u);
            umma_commit(dp_full);
            ++dc;
          }
          if (x >= 1) {
            const int xx = x - 1;
            const int st = (rc + xx) % AB_RING;
            if (xx == 0) mbar_wait(acc_empty, (ic & 1) ^ 1);
            mbar_wait(pds_full, pc & 1);
            tc_fence_after();
            const uint32_t pa = smem_u32(sPm), da = smem_u32(sdS);
            const uint32_t qaddr = smem_u32(sQ + st * AT_TILE_BYTES);
            const uint32_t oaddr = smem_u32(sdO + st * AT_TILE_BYTES);
#pragma unroll
            for (int ks = 0; ks < 8; ++ks)            // dV_j += P^T dO_i   (K = 128 queries)
              umma_bf16_ss(tmem_base + AB_TM_DV,
                           make_smem_desc_sw128(pa + ks * 2048, AT_TILE_BYTES, 1024),
                           make_smem_desc_sw128(oaddr + ks * 2048, AT_TILE_BYTES, 1024), idesc_g,
                           (xx | ks) ? 1u : 0u);
#pragma unroll
            for (int ks = 0; ks < 8; ++ks)            // dK_j += dS^T Q_i
              umma_bf16_ss(tmem_base + AB_TM_DK,
                           make_smem_desc_sw128(da + ks * 2048, AT_TILE_BYTES, 1024),
                           make_smem_desc_sw128(qaddr + ks * 2048, AT_TILE_BYTES, 1024), idesc_g,
                           (xx | ks) ? 1u : 0u);
            umma_commit(pds_empty);
            umma_commit(&qd_empty[st]);
            if (xx == w.ni - 1) {
              umma_commit(acc_full);
              umma_commit(&kv_empty[kvb]);
            }
            ++pc;
          }
        }
        rc += w.ni;
      }
    }
    __syncwarp();
  } else if (warp == 10) {
    // ============================ dS storer ============================
    if (lane == 0) {
      uint32_t pc = 0;
      for (int it = blockIdx.x; it < p.items; it += gridDim.x) {
        const AttnBwdItem w = attn_bwd_item(p, it);
        for (int x = 0; x < w.ni; ++x, ++pc) {
          mbar_wait(pds_full, pc & 1);
          at_tma_store(&tmdS, sdS, w.j * 128, (w.i0 + x) * 128, w.h, w.b);
          if (w.j * 128 + 64 < p.T)
            at_tma_store(&tmdS, sdS + AT_TILE_BYTES, w.j * 128 + 64, (w.i0 + x) * 128, w.h, w.b);
          bulk_commit();
          bulk_wait_read0();
          mbar_arrive(pds_empty);
        }
      }
      bulk_wait0();
    }
    __syncwarp();
  } else {
    // ============================ softmax warps 0..7 ============================
    const int q = warp & 3, hf = warp >> 2;
    const int rt = q * 32 + lane;
    const uint32_t tlane = uint32_t(q * 32) << 16;
    const float c2 = p.scale2, sc_ = p.scale;
    uint32_t ic = 0, sc = 0, dc = 0, pc = 0;
    for (int it = blockIdx.x; it < p.items; it += gridDim.x, ++ic) {
      const AttnBwdItem w = attn_bwd_item(p, it);
      const long long gbase = ((long long)w.b * p.H + w.h) * p.T;
      for (int x = 0; x < w.ni; ++x) {
        const int i = w.i0 + x;
        const int row = i * 128 + rt;
        const bool row_ok = row < p.T;
        const float mneg = row_ok ? -p.rowm[gbase + row] : 0.f;
        const float il = row_ok ? p.rowinvl[gbase + row] : 0.f;
        const float Dr = row_ok ? p.D[gbase + row] : 0.f;
        const int climit = p.causal ? min(row, p.T - 1) : p.T - 1;
        const bool edge = (p.causal && i == w.j) || (w.j * 128 + hf * 64 + 64 > p.T);
        // dP tile -> registers
        uint32_t dp[64];
        mbar_wait(dp_full, dc & 1);
        tc_fence_after();
        tmem_ld_32x32(tmem_base + tlane + AB_TM_DP + hf * 64, *reinterpret_cast<uint32_t(*)[32]>(&dp[0]));
        tmem_ld_32x32(tmem_base + tlane + AB_TM_DP + hf * 64 + 32,
                      *reinterpret_cast<uint32_t(*)[32]>(&dp[32]));
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(dp_empty);
        ++dc;
        const int sb = sc & 1;
        mbar_wait(&s_full[sb], (sc >> 1) & 1);
        tc_fence_after();
        uint8_t* prow = sPm + hf * AT_TILE_BYTES + rt * 128;
        uint8_t* drow = sdS + hf * AT_TILE_BYTES + rt * 128;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t s[32];
          tmem_ld_32x32(tmem_base + tlane + sb * 128 + hf * 64 + half * 32, s);
          tmem_ld_wait();
          if (half == 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_empty[sb]);
          }
          if (edge) {
            const int col0 = w.j * 128 + hf * 64 + half * 32;
#pragma unroll
            for (int c = 0; c < 32; ++c)
              if (col0 + c > climit) s[c] = 0xff800000u;
          }
          uint32_t pk[16], dk_[16];
#pragma unroll
          for (int c = 0; c < 16; ++c) {
            const float p0 = ex2_approx(fmaf(__uint_as_float(s[2 * c]), c2, mneg)) * il;
            const float p1 = ex2_approx(fmaf(__uint_as_float(s[2 * c + 1]), c2, mneg)) * il;
            const float d0 = p0 * sc_ * (__uint_as_float(dp[half * 32 + 2 * c]) - Dr);
            const float d1 = p1 * sc_ * (__uint_as_float(dp[half * 32 + 2 * c + 1]) - Dr);
            pk[c] = pack_bf16x2(p0, p1);
            dk_[c] = pack_bf16x2(d0, d1);
          }
          if (half == 0) mbar_wait(pds_empty, (pc & 1) ^ 1);
#pragma unroll
          for (int ch = 0; ch < 4; ++ch) {
            const int off = ((half * 4 + ch) ^ (rt & 7)) << 4;
            *reinterpret_cast<uint4*>(prow + off) =
                make_uint4(pk[4 * ch], pk[4 * ch + 1], pk[4 * ch + 2], pk[4 * ch + 3]);
            *reinterpret_cast<uint4*>(drow + off) =
                make_uint4(dk_[4 * ch], dk_[4 * ch + 1], dk_[4 * ch + 2], dk_[4 * ch + 3]);
          }
        }
        ++sc;
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(pds_full);
        ++pc;
      }
      // ---------------- dV_j, dK_j epilogue ----------------
      mbar_wait(acc_full, ic & 1);
      tc_fence_after();
      uint32_t gv[32], gk[32];
      tmem_ld_32x32(tmem_base + tlane + AB_TM_DV + hf * 32, gv);
      tmem_ld_32x32(tmem_base + tlane + AB_TM_DK + hf * 32, gk);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty);
      const int key = w.j * 128 + rt;
      if (key < p.T) {
        const long long off = (long long)w.b * p.bs + (long long)key * p.ld + (long long)w.h * p.hs + hf * 32;
        uint4* dv4 = reinterpret_cast<uint4*>(p.dv + off);
        uint4* dk4 = reinterpret_cast<uint4*>(p.dk + off);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          dv4[g] = make_uint4(
              pack_bf16x2(__uint_as_float(gv[8 * g]), __uint_as_float(gv[8 * g + 1])),
              pack_bf16x2(__uint_as_float(gv[8 * g + 2]), __uint_as_float(gv[8 * g + 3])),
              pack_bf16x2(__uint_as_float(gv[8 * g + 4]), __uint_as_float(gv[8 * g + 5])),
              pack_bf16x2(__uint_as_float(gv[8 * g + 6]), __uint_as_float(gv[8 * g + 7])));
          dk4[g] = make_uint4(
              pack_bf16x2(__uint_as_float(gk[8 * g]), __uint_as_float(gk[8 * g + 1])),
              pack_bf16x2(__uint_as_float(gk[8 * g + 2]), __uint_as_float(gk[8 * g + 3])),
              pack_bf16x2(__uint_as_float(gk[8 * g + 4]), __uint_as_float(gk[8 * g + 5])),
              pack_bf16x2(__uint_as_float(gk[8 * g + 6]), __uint_as_float(gk[8 * g + 7])));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, AT_TMEM_COLS);
}

constexpr size_t AB_SMEM = 1024 + 4 * AT_TILE_BYTES + 2 * AB_RING * AT_TILE_BYTES + 2 * AT_P_BYTES + 256;
}  // namespace dtg

// dK, dV and dS (= scale * P * (dO V^T - D)) for head_dim 64.  q/k/v/dk/dv: [B, T, H, 64] views
// with strides (ld, hs, bs); dO: [B, T, H, 64] view with its own strides; dS: [B, H, T, T]
// bf16, zero-initialised when causal (tiles above the diagonal are not written).
extern "C" int dtg_attn_bwd(const void* q, const void* k, const void* v, void* dk, void* dv,
                            long long ld, long long hs, long long bs, const void* dO,
                            long long do_ld, long long do_hs, long long do_bs, int B, int H, int T,
                            int hd, void* dS, const float* rowm, const float* rowinvl,
                            const float* D, float scale, int causal, void* stream_) {
  using namespace dtg;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if (hd != 64) {
    dtg_set_last_error("dtg_attn_bwd", "fused attention supports head_dim 64 only");
    return ERR_UNSUPPORTED;
  }
  if (!q || !k || !v || !dk || !dv || !dO || !dS || !rowm || !rowinvl || !D || B <= 0 || H <= 0 ||
      T <= 0 || (ld % 8) || (hs % 8) || (bs % 8) || (do_ld % 8) || (do_hs % 8) || (do_bs % 8) ||
      (T % 8) || (reinterpret_cast<uintptr_t>(dk) & 15) || (reinterpret_cast<uintptr_t>(dv) & 15)) {
    dtg_set_last_error("dtg_attn_bwd", "null pointer, bad shape or unaligned pitch");
    return ERR_BAD_ARG;
  }
  CUtensorMap tq, tk, tv, to, ts;
  int rc;
  if ((rc = tensor_map_bf16(q, 64, T, ld, 64, 128, &tq, H, B, hs, bs)) != OK) return rc;
  if ((rc = tensor_map_bf16(k, 64, T, ld, 64, 128, &tk, H, B, hs, bs)) != OK) return rc;
  if ((rc = tensor_map_bf16(v, 64, T, ld, 64, 128, &tv, H, B, hs, bs)) != OK) return rc;
  if ((rc = tensor_map_bf16(dO, 64, T, do_ld, 64, 128, &to, H, B, do_hs, do_bs)) != OK) return rc;
  if ((rc = tensor_map_bf16(dS, T, T, T, 64, 128, &ts, H, B, (uint64_t)T * T,
                            (uint64_t)H * T * T)) != OK)
    return rc;
  AttnBwdParams p;
  p.B = B; p.H = H; p.T = T; p.causal = causal ? 1 : 0;
  p.n_t = (T + 127) / 128;
  p.items = B * H * p.n_t;
  p.scale = scale;
  p.scale2 = scale * 1.4426950408889634f;
  p.rowm = rowm; p.rowinvl = rowinvl; p.D = D;
  p.dk = reinterpret_cast<__nv_bfloat16*>(dk);
  p.dv = reinterpret_cast<__nv_bfloat16*>(dv);
  p.ld = ld; p.hs = hs; p.bs = bs;
  static bool attr_set = false;
  if (!attr_set) {
    DTG_CUDA_OK(cudaFuncSetAttribute(attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)AB_SMEM));
    attr_set = true;
  }
  const int sms = num_sms();
  const int grid = p.items < sms ? p.items : sms;
  attn_bwd_kernel<<<grid, AT_THREADS, AB_SMEM, stream>>>(tq, tk, tv, to, ts, p);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}
