// Transformer support kernels for sm_100a (SURVEY section 8, W12 and the non-GEMM half of
// W13): LayerNorm fwd/bwd, token+position embedding fwd/bwd, and the attention softmax
// fwd/bwd (causal or bidirectional).  The GEMM-shaped parts of attention (QK^T, PV and
// the four backward products) run on the tcgen05 GEMM of gemm_sm100.cu in its batched
// form; these kernels are the row-wise, memory-bound glue between them.  GELU and its
// derivative live in the GEMM epilogue.
//
// Gradient outputs that belong to parameters (dgamma, dbeta, dwte, dwpe) are always
// *reduction adds* of alpha * grad into their target, like bias_grad: the same kernel
// then serves a zeroed gradient buffer, a window accumulator, or a push into a ps shard
// in peer memory (alpha = -lr).
#include "common.cuh"

namespace dtg {

__device__ __forceinline__ void bf16x4_to_float(const uint2 raw, float (&f)[4]) {
  const __nv_bfloat162 lo = *reinterpret_cast<const __nv_bfloat162*>(&raw.x);
  const __nv_bfloat162 hi = *reinterpret_cast<const __nv_bfloat162*>(&raw.y);
  f[0] = __low2float(lo); f[1] = __high2float(lo);
  f[2] = __low2float(hi); f[3] = __high2float(hi);
}

// ------------------------------------------------------------------ LayerNorm
// One warp per row; cols % 4 == 0.  y = (x - mean) * rstd * gamma + beta.
constexpr int LN_MAX_PER_LANE = 32;   // cols <= 32 lanes * 4 * 8 = 4096 (cols/128 vectors per lane)

__global__ void __launch_bounds__(256)
layernorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma,
                     const float* __restrict__ beta, __nv_bfloat16* __restrict__ y,
                     float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows, int cols,
                     float eps) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  if (row >= rows) return;
  const __nv_bfloat16* xr = x + (long long)row * cols;
  const int nvec = cols >> 2;
  float v[LN_MAX_PER_LANE];
  float s = 0.f;
  int cnt = 0;
  for (int i = lane; i < nvec; i += 32, ++cnt) {
    float f[4];
    bf16x4_to_float(reinterpret_cast<const uint2*>(xr)[i], f);
#pragma unroll
    for (int e = 0; e < 4; ++e) { v[cnt * 4 + e] = f[e]; s += f[e]; }
  }
  const float mean = warp_sum(s) / cols;
  float q = 0.f;
  for (int j = 0; j < cnt * 4; ++j) { const float d = v[j] - mean; q += d * d; }
  const float rstd = rsqrtf(warp_sum(q) / cols + eps);
  cnt = 0;
  __nv_bfloat16* yr = y + (long long)row * cols;
  for (int i = lane; i < nvec; i += 32, ++cnt) {
    const float4 g = reinterpret_cast<const float4*>(gamma)[i];
    const float4 b = reinterpret_cast<const float4*>(beta)[i];
    const float o0 = (v[cnt * 4 + 0] - mean) * rstd * g.x + b.x;
    const float o1 = (v[cnt * 4 + 1] - mean) * rstd * g.y + b.y;
    const float o2 = (v[cnt * 4 + 2] - mean) * rstd * g.z + b.z;
    const float o3 = (v[cnt * 4 + 3] - mean) * rstd * g.w + b.w;
    reinterpret_cast<uint2*>(yr)[i] = make_uint2(pack_bf16x2(o0, o1), pack_bf16x2(o2, o3));
  }
  if (lane == 0) {
    if (mean_out) mean_out[row] = mean;
    if (rstd_out) rstd_out[row] = rstd;
  }
}

// dx = rstd * (g*dy - mean(g*dy) - xhat * mean(g*dy*xhat)) (+ dx_add);
// dgamma += alpha * sum_rows dy*xhat; dbeta += alpha * sum_rows dy.
// Block = 8 warps, each warp walks rows (grid-stride); per-lane partial dgamma/dbeta for
// its columns are reduced across warps in smem and leave as one red.add per column.
__global__ void __launch_bounds__(256)
layernorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                     const float* __restrict__ gamma, const float* __restrict__ mean,
                     const float* __restrict__ rstd, const __nv_bfloat16* __restrict__ dx_add,
                     __nv_bfloat16* __restrict__ dx, float* __restrict__ dgamma,
                     float* __restrict__ dbeta, int rows, int cols, float alpha) {
  extern __shared__ float sred[];   // [2][cols]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = cols >> 2;
  for (int i = threadIdx.x; i < 2 * cols; i += blockDim.x) sred[i] = 0.f;
  __syncthreads();
  float dg[LN_MAX_PER_LANE], db[LN_MAX_PER_LANE];
#pragma unroll
  for (int j = 0; j < LN_MAX_PER_LANE; ++j) { dg[j] = 0.f; db[j] = 0.f; }
  for (int row = blockIdx.x * 8 + warp; row < rows; row += gridDim.x * 8) {
    const __nv_bfloat16* xr = x + (long long)row * cols;
    const __nv_bfloat16* dyr = dy + (long long)row * cols;
    const float mu = mean[row], rs = rstd[row];
    float xh[LN_MAX_PER_LANE], gdy[LN_MAX_PER_LANE];
    float s1 = 0.f, s2 = 0.f;
    int cnt = 0;
    for (int i = lane; i < nvec; i += 32, ++cnt) {
      float fx[4], fd[4];
      bf16x4_to_float(reinterpret_cast<const uint2*>(xr)[i], fx);
      bf16x4_to_float(reinterpret_cast<const uint2*>(dyr)[i], fd);
      const float4 g = reinterpret_cast<const float4*>(gamma)[i];
      const float gg[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float h = (fx[e] - mu) * rs;
        const float t = gg[e] * fd[e];
        xh[cnt * 4 + e] = h;
        gdy[cnt * 4 + e] = t;
        s1 += t;
        s2 += t * h;
        dg[cnt * 4 + e] += fd[e] * h;
        db[cnt * 4 + e] += fd[e];
      }
    }
    s1 = warp_sum(s1) / cols;
    s2 = warp_sum(s2) / cols;
    cnt = 0;
    __nv_bfloat16* dxr = dx + (long long)row * cols;
    for (int i = lane; i < nvec; i += 32, ++cnt) {
      float o[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) o[e] = rs * (gdy[cnt * 4 + e] - s1 - xh[cnt * 4 + e] * s2);
      if (dx_add != nullptr) {
        float a[4];
        bf16x4_to_float(reinterpret_cast<const uint2*>(dx_add + (long long)row * cols)[i], a);
#pragma unroll
        for (int e = 0; e < 4; ++e) o[e] += a[e];
      }
      reinterpret_cast<uint2*>(dxr)[i] = make_uint2(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]));
    }
  }
  if (dgamma != nullptr || dbeta != nullptr) {
    int cnt = 0;
    for (int i = lane; i < nvec; i += 32, ++cnt) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        atomicAdd(&sred[i * 4 + e], dg[cnt * 4 + e]);
        atomicAdd(&sred[cols + i * 4 + e], db[cnt * 4 + e]);
      }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < cols; c += blockDim.x) {
      if (dgamma) red_add_f32(dgamma + c, alpha * sred[c]);
      if (dbeta) red_add_f32(dbeta + c, alpha * sred[cols + c]);
    }
  }
}

// Fast path for cols = 256 * V: every lane owns V 16-byte vectors of the row, all loads of a
// row are in flight before the first use, x / dy stay packed in registers and xhat is
// recomputed for the second pass instead of being kept in fp32.
__device__ __forceinline__ void bf16x8_to_float(const uint4& u, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) { f[2 * i] = __low2float(h[i]); f[2 * i + 1] = __high2float(h[i]); }
}

template <int V>
__global__ void __launch_bounds__(256, 1)
layernorm_bwd_fast_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                          const float* __restrict__ gamma, const float* __restrict__ mean,
                          const float* __restrict__ rstd, const __nv_bfloat16* __restrict__ dx_add,
                          __nv_bfloat16* __restrict__ dx, float* __restrict__ dgamma,
                          float* __restrict__ dbeta, int rows, float alpha) {
  constexpr int COLS = 256 * V;
  __shared__ float sred[2 * COLS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 2 * COLS; i += 256) sred[i] = 0.f;
  __syncthreads();
  float dg[V * 8], db[V * 8];
  // gamma is re-read per row (8 KB, L1-resident) instead of pinning 8V registers
  auto load_gamma = [&](int v, float (&g)[8]) {
    const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma) + (v * 32 + lane) * 2);
    const float4 g1 = __ldg(reinterpret_cast<const float4*>(gamma) + (v * 32 + lane) * 2 + 1);
    g[0] = g0.x; g[1] = g0.y; g[2] = g0.z; g[3] = g0.w;
    g[4] = g1.x; g[5] = g1.y; g[6] = g1.z; g[7] = g1.w;
  };
#pragma unroll
  for (int j = 0; j < V * 8; ++j) { dg[j] = 0.f; db[j] = 0.f; }
  for (int row = blockIdx.x * 8 + warp; row < rows; row += gridDim.x * 8) {
    const uint4* xr = reinterpret_cast<const uint4*>(x + (long long)row * COLS);
    const uint4* dyr = reinterpret_cast<const uint4*>(dy + (long long)row * COLS);
    const uint4* ar = reinterpret_cast<const uint4*>(dx_add + (long long)row * COLS);
    uint4 xv[V], dv[V], av[V];
#pragma unroll
    for (int v = 0; v < V; ++v) { xv[v] = xr[v * 32 + lane]; dv[v] = dyr[v * 32 + lane]; }
    if (dx_add != nullptr) {
#pragma unroll
      for (int v = 0; v < V; ++v) av[v] = ar[v * 32 + lane];
    }
    const float mu = mean[row], rs = rstd[row];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int v = 0; v < V; ++v) {
      float fx[8], fd[8], gam[8];
      bf16x8_to_float(xv[v], fx);
      bf16x8_to_float(dv[v], fd);
      load_gamma(v, gam);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float h = (fx[e] - mu) * rs;
        const float t = gam[e] * fd[e];
        s1 += t;
        s2 += t * h;
        dg[v * 8 + e] += fd[e] * h;
        db[v * 8 + e] += fd[e];
      }
    }
    s1 = warp_sum(s1) * (1.f / COLS);
    s2 = warp_sum(s2) * (1.f / COLS);
    uint4* dxr = reinterpret_cast<uint4*>(dx + (long long)row * COLS);
#pragma unroll
    for (int v = 0; v < V; ++v) {
      float fx[8], fd[8], o[8], gam[8];
      bf16x8_to_float(xv[v], fx);
      bf16x8_to_float(dv[v], fd);
      load_gamma(v, gam);
#pragma unroll
      for (int e = 0; e < 8; ++e)
        o[e] = rs * (gam[e] * fd[e] - s1 - (fx[e] - mu) * rs * s2);
      if (dx_add != nullptr) {
        float a[8];
        bf16x8_to_float(av[v], a);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] += a[e];
      }
      dxr[v * 32 + lane] = make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]),
                                      pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
    }
  }
  if (dgamma != nullptr || dbeta != nullptr) {
#pragma unroll
    for (int v = 0; v < V; ++v)
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int c = (v * 32 + lane) * 8 + e;
        atomicAdd(&sred[c], dg[v * 8 + e]);
        atomicAdd(&sred[COLS + c], db[v * 8 + e]);
      }
    __syncthreads();
    for (int c = threadIdx.x; c < COLS; c += 256) {
      if (dgamma) red_add_f32(dgamma + c, alpha * sred[c]);
      if (dbeta) red_add_f32(dbeta + c, alpha * sred[COLS + c]);
    }
  }
}

// ------------------------------------------------------------------ embedding
// out[r, :] = wte[ids[r], :] + wpe[r % seq, :]   (wpe may be null)
__global__ void __launch_bounds__(256)
embedding_fwd_kernel(const long long* __restrict__ ids, const __nv_bfloat16* __restrict__ wte,
                     const __nv_bfloat16* __restrict__ wpe, __nv_bfloat16* __restrict__ out,
                     int rows, int seq, int dim) {
  const int nvec = dim >> 2;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < (long long)rows * nvec;
       t += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(t / nvec), i = (int)(t % nvec);
    float a[4], b[4] = {0.f, 0.f, 0.f, 0.f};
    bf16x4_to_float(reinterpret_cast<const uint2*>(wte + ids[r] * dim)[i], a);
    if (wpe) bf16x4_to_float(reinterpret_cast<const uint2*>(wpe + (long long)(r % seq) * dim)[i], b);
    reinterpret_cast<uint2*>(out + (long long)r * dim)[i] =
        make_uint2(pack_bf16x2(a[0] + b[0], a[1] + b[1]), pack_bf16x2(a[2] + b[2], a[3] + b[3]));
  }
}

// dwte[ids[r], :] += alpha * dout[r, :];  dwpe[r % seq, :] += alpha * dout[r, :]
__global__ void __launch_bounds__(256)
embedding_bwd_kernel(const long long* __restrict__ ids, const __nv_bfloat16* __restrict__ dout,
                     float* __restrict__ dwte, float* __restrict__ dwpe, int rows, int seq, int dim,
                     float alpha) {
  const int nvec = dim >> 2;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < (long long)rows * nvec;
       t += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(t / nvec), i = (int)(t % nvec);
    float d[4];
    bf16x4_to_float(reinterpret_cast<const uint2*>(dout + (long long)r * dim)[i], d);
    if (dwte)
      red_add_v4_f32(dwte + ids[r] * dim + 4 * i, alpha * d[0], alpha * d[1], alpha * d[2],
                     alpha * d[3]);
    if (dwpe)
      red_add_v4_f32(dwpe + (long long)(r % seq) * dim + 4 * i, alpha * d[0], alpha * d[1],
                     alpha * d[2], alpha * d[3]);
  }
}

// ----------------------------------------------------- attention softmax fwd/bwd
// S: fp32 [nrows_total, T] (rows = (b, h, q)), P: bf16 same shape.  Row r has query index
// q = r % T; causal keeps keys k <= q.  One warp per row, T <= 4096, T % 4 == 0.
__global__ void __launch_bounds__(256)
attn_softmax_fwd_kernel(const float* __restrict__ S, __nv_bfloat16* __restrict__ P, long long nrows,
                        int T, float scale, int causal) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * 8 + warp;
  if (row >= nrows) return;
  const int q = (int)(row % T);
  const int limit = causal ? q + 1 : T;
  const float* sr = S + row * T;
  const int nvec = T >> 2;
  float v[LN_MAX_PER_LANE];
  float mx = -INFINITY;
  int cnt = 0;
  for (int i = lane; i < nvec; i += 32, ++cnt) {
    if (4 * i < limit) {
      const float4 f = reinterpret_cast<const float4*>(sr)[i];
      const float ff[4] = {f.x, f.y, f.z, f.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float t = (4 * i + e < limit) ? ff[e] * scale : -INFINITY;
        v[cnt * 4 + e] = t;
        mx = fmaxf(mx, t);
      }
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e) v[cnt * 4 + e] = -INFINITY;
    }
  }
  mx = warp_max(mx);
  float sum = 0.f;
  for (int j = 0; j < cnt * 4; ++j) { const float e = __expf(v[j] - mx); v[j] = e; sum += e; }
  const float inv = 1.f / warp_sum(sum);
  cnt = 0;
  __nv_bfloat16* pr = P + row * T;
  for (int i = lane; i < nvec; i += 32, ++cnt)
    reinterpret_cast<uint2*>(pr)[i] =
        make_uint2(pack_bf16x2(v[cnt * 4] * inv, v[cnt * 4 + 1] * inv),
                   pack_bf16x2(v[cnt * 4 + 2] * inv, v[cnt * 4 + 3] * inv));
}

// dS = scale * P * (dP - sum_k P*dP), written bf16 over... a separate buffer dS.
__global__ void __launch_bounds__(256)
attn_softmax_bwd_kernel(const __nv_bfloat16* __restrict__ P, const float* __restrict__ dP,
                        __nv_bfloat16* __restrict__ dS, long long nrows, int T, float scale) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * 8 + warp;
  if (row >= nrows) return;
  const int nvec = T >> 2;
  float p[LN_MAX_PER_LANE], d[LN_MAX_PER_LANE];
  float dot = 0.f;
  int cnt = 0;
  for (int i = lane; i < nvec; i += 32, ++cnt) {
    float fp[4];
    bf16x4_to_float(reinterpret_cast<const uint2*>(P + row * T)[i], fp);
    const float4 f = reinterpret_cast<const float4*>(dP + row * T)[i];
    const float fd[4] = {f.x, f.y, f.z, f.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) { p[cnt * 4 + e] = fp[e]; d[cnt * 4 + e] = fd[e]; dot += fp[e] * fd[e]; }
  }
  dot = warp_sum(dot);
  cnt = 0;
  for (int i = lane; i < nvec; i += 32, ++cnt) {
    float o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) o[e] = scale * p[cnt * 4 + e] * (d[cnt * 4 + e] - dot);
    reinterpret_cast<uint2*>(dS + row * T)[i] =
        make_uint2(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]));
  }
}

// elementwise add of two bf16 tensors (residual joins in backward)
__global__ void __launch_bounds__(512)
add_bf16_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b,
                __nv_bfloat16* __restrict__ out, long long n4) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
       i += (long long)gridDim.x * blockDim.x) {
    float x[4], y[4];
    bf16x4_to_float(reinterpret_cast<const uint2*>(a)[i], x);
    bf16x4_to_float(reinterpret_cast<const uint2*>(b)[i], y);
    reinterpret_cast<uint2*>(out)[i] =
        make_uint2(pack_bf16x2(x[0] + y[0], x[1] + y[1]), pack_bf16x2(x[2] + y[2], x[3] + y[3]));
  }
}

}  // namespace dtg

using namespace dtg;

static bool al(const void* p, int a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; }

extern "C" int dtg_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y,
                                 float* mean, float* rstd, int rows, int cols, float eps,
                                 void* stream) {
  if (rows <= 0 || cols <= 0 || (cols & 3) || cols > 128 * LN_MAX_PER_LANE || !al(x, 8) ||
      !al(y, 8) || !al(gamma, 16) || !al(beta, 16)) {
    dtg_set_last_error("dtg_layernorm_fwd", "cols must be a multiple of 4 (<= 4096), buffers aligned");
    return ERR_BAD_ARG;
  }
  layernorm_fwd_kernel<<<(rows + 7) / 8, 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, gamma, beta, (__nv_bfloat16*)y, mean, rstd, rows, cols, eps);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_layernorm_bwd(const void* dy, const void* x, const float* gamma,
                                 const float* mean, const float* rstd, const void* dx_add, void* dx,
                                 float* dgamma, float* dbeta, int rows, int cols, float alpha,
                                 void* stream) {
  if (rows <= 0 || cols <= 0 || (cols & 3) || cols > 128 * LN_MAX_PER_LANE) {
    dtg_set_last_error("dtg_layernorm_bwd", "cols must be a multiple of 4 (<= 4096)");
    return ERR_BAD_ARG;
  }
  int blocks = (rows + 7) / 8;
  if (blocks > 148 * 2) blocks = 148 * 2;
  const bool al16 = ((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(x) |
                      reinterpret_cast<uintptr_t>(dx) | reinterpret_cast<uintptr_t>(dx_add) |
                      reinterpret_cast<uintptr_t>(gamma)) & 15) == 0;
#define DTG_LN_FAST(V_)                                                                        \
  if (al16 && cols == 256 * V_) {                                                              \
    if (blocks > 148) blocks = 148;                                                            \
    layernorm_bwd_fast_kernel<V_><<<blocks, 256, 0, (cudaStream_t)stream>>>(                   \
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, gamma, mean, rstd,                  \
        (const __nv_bfloat16*)dx_add, (__nv_bfloat16*)dx, dgamma, dbeta, rows, alpha);         \
    DTG_CUDA_OK(cudaGetLastError());                                                           \
    return OK;                                                                                 \
  }
  DTG_LN_FAST(2) DTG_LN_FAST(3) DTG_LN_FAST(4) DTG_LN_FAST(5)
#undef DTG_LN_FAST
  layernorm_bwd_kernel<<<blocks, 256, 2 * cols * sizeof(float), (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, gamma, mean, rstd,
      (const __nv_bfloat16*)dx_add, (__nv_bfloat16*)dx, dgamma, dbeta, rows, cols, alpha);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_embedding_fwd(const long long* ids, const void* wte, const void* wpe, void* out,
                                 int rows, int seq_len, int dim, void* stream) {
  if (rows <= 0 || (dim & 3)) {
    dtg_set_last_error("dtg_embedding_fwd", "dim must be a multiple of 4");
    return ERR_BAD_ARG;
  }
  long long work = (long long)rows * (dim >> 2);
  int blocks = (int)((work + 255) / 256 > 148 * 8 ? 148 * 8 : (work + 255) / 256);
  embedding_fwd_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(
      ids, (const __nv_bfloat16*)wte, (const __nv_bfloat16*)wpe, (__nv_bfloat16*)out, rows,
      seq_len, dim);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_embedding_bwd(const long long* ids, const void* dout, float* dwte, float* dwpe,
                                 int rows, int seq_len, int dim, float alpha, void* stream) {
  if (rows <= 0 || (dim & 3)) {
    dtg_set_last_error("dtg_embedding_bwd", "dim must be a multiple of 4");
    return ERR_BAD_ARG;
  }
  long long work = (long long)rows * (dim >> 2);
  int blocks = (int)((work + 255) / 256 > 148 * 8 ? 148 * 8 : (work + 255) / 256);
  embedding_bwd_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(
      ids, (const __nv_bfloat16*)dout, dwte, dwpe, rows, seq_len, dim, alpha);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_attn_softmax_fwd(const float* S, void* P, long long nrows, int T, float scale,
                                    int causal, void* stream) {
  if (nrows <= 0 || (T & 3) || T > 128 * LN_MAX_PER_LANE) {
    dtg_set_last_error("dtg_attn_softmax_fwd", "T must be a multiple of 4 (<= 4096)");
    return ERR_BAD_ARG;
  }
  attn_softmax_fwd_kernel<<<(unsigned)((nrows + 7) / 8), 256, 0, (cudaStream_t)stream>>>(
      S, (__nv_bfloat16*)P, nrows, T, scale, causal);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_attn_softmax_bwd(const void* P, const float* dP, void* dS, long long nrows, int T,
                                    float scale, void* stream) {
  if (nrows <= 0 || (T & 3) || T > 128 * LN_MAX_PER_LANE) {
    dtg_set_last_error("dtg_attn_softmax_bwd", "T must be a multiple of 4 (<= 4096)");
    return ERR_BAD_ARG;
  }
  attn_softmax_bwd_kernel<<<(unsigned)((nrows + 7) / 8), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)P, dP, (__nv_bfloat16*)dS, nrows, T, scale);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_add_bf16(const void* a, const void* b, void* out, long long n, void* stream) {
  if (n <= 0 || (n & 3)) {
    dtg_set_last_error("dtg_add_bf16", "n must be a positive multiple of 4");
    return ERR_BAD_ARG;
  }
  long long n4 = n >> 2;
  int blocks = (int)((n4 + 511) / 512 > 148 * 8 ? 148 * 8 : (n4 + 511) / 512);
  add_bf16_kernel<<<blocks, 512, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)a, (const __nv_bfloat16*)b, (__nv_bfloat16*)out, n4);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

// ------------------------------------------------ attention: statistics merge, row dots
namespace dtg {
// stats [nrows][tiles_n][2] (running max, sum-exp in the log2 domain, per 128-key tile)
// -> m[row], 1/l[row].  Causal rows only merge the tiles that contain visible keys.
__global__ void __launch_bounds__(256)
attn_stats_merge_kernel(const float* __restrict__ stats, float* __restrict__ rowm,
                        float* __restrict__ rowinvl, long long nrows, int tiles_n, int T, int BN,
                        int causal) {
  const long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= nrows) return;
  const int q = (int)(row % T);
  const int nt = causal ? min(tiles_n, q / BN + 1) : tiles_n;
  const float* sp = stats + row * tiles_n * 2;
  float m = -INFINITY;
  for (int t = 0; t < nt; ++t) m = fmaxf(m, sp[2 * t]);
  float l = 0.f;
  for (int t = 0; t < nt; ++t) l += sp[2 * t + 1] * exp2f(sp[2 * t] - m);
  rowm[row] = m;
  rowinvl[row] = 1.f / l;
}

// D[b, h, t] = sum_c dO[b*T+t, h*hd+c] * O[b*T+t, h*hd+c]   (one warp per (row, head))
__global__ void __launch_bounds__(256)
attn_rowdot_kernel(const __nv_bfloat16* __restrict__ dO, const __nv_bfloat16* __restrict__ O,
                   float* __restrict__ D, int B, int T, int H, int hd) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long idx = (long long)blockIdx.x * 8 + warp;       // over B*T*H
  if (idx >= (long long)B * T * H) return;
  const int h = (int)(idx % H);
  const long long r = idx / H;                                   // b*T + t
  const long long base = r * (long long)H * hd + (long long)h * hd;
  float s = 0.f;
  for (int c = lane * 2; c < hd; c += 64) {
    const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(dO + base + c);
    const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162*>(O + base + c);
    s += __low2float(a) * __low2float(b) + __high2float(a) * __high2float(b);
  }
  s = warp_sum(s);
  if (lane == 0) {
    const long long b_ = r / T, t = r % T;
    D[(b_ * H + h) * T + t] = s;
  }
}
}  // namespace dtg

extern "C" int dtg_attn_stats_merge(const float* stats, float* rowm, float* rowinvl,
                                    long long nrows, int tiles_n, int T, int causal, void* stream) {
  if (nrows <= 0 || tiles_n <= 0) return dtg::OK;
  dtg::attn_stats_merge_kernel<<<(unsigned)((nrows + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      stats, rowm, rowinvl, nrows, tiles_n, T, 128, causal);
  DTG_CUDA_OK(cudaGetLastError());
  return dtg::OK;
}

extern "C" int dtg_attn_rowdot(const void* dO, const void* O, float* D, int B, int T, int H, int hd,
                               void* stream) {
  if ((hd & 1) != 0) {
    dtg_set_last_error("dtg_attn_rowdot", "head_dim must be even");
    return dtg::ERR_BAD_ARG;
  }
  const long long n = (long long)B * T * H;
  dtg::attn_rowdot_kernel<<<(unsigned)((n + 7) / 8), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dO, (const __nv_bfloat16*)O, D, B, T, H, hd);
  DTG_CUDA_OK(cudaGetLastError());
  return dtg::OK;
}
