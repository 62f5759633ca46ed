// Convolution-network support kernels for sm_100a: everything a
// ResNet needs around the tcgen05 GEMM.  Activations are NHWC bf16, so a convolution is
//   im2col (gather, this file) -> GEMM [N*Ho*Wo, kh*kw*Cin] x [Cout, kh*kw*Cin]^T
// with 1x1/stride-1 convolutions skipping the gather entirely; the data gradient is
//   GEMM dY x W -> col2im (gather form: each input pixel sums the taps that read it);
// the weight gradient is the MN-major GEMM dY^T x col with the usual gradient sinks.
// BatchNorm (batch statistics), ReLU, the residual add, max/average pooling are the
// row-wise kernels below.  Parameter gradients (gamma, beta) leave as reduction adds
// of alpha * grad, like every other parameter gradient in this library.
#include "common.cuh"
#include <cstdlib>

namespace dtg {

__device__ __forceinline__ void bf16x8_to_float(const uint4 raw, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
  for (int i = 0; i < 4; ++i) { f[2 * i] = __low2float(h[i]); f[2 * i + 1] = __high2float(h[i]); }
}
__device__ __forceinline__ uint4 float_to_bf16x8(const float (&f)[8]) {
  return make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]),
                    pack_bf16x2(f[6], f[7]));
}

struct ConvGeom {
  int N, H, W, C;        // input NHWC
  int Ho, Wo;            // output spatial
  int kh, kw, stride, pad;
};

// col[(n,ho,wo), (r,s,c)] = x[n, ho*stride - pad + r, wo*stride - pad + s, c]  (0 outside)
// C % 8 == 0: one thread moves 8 channels (16 B).
__global__ void __launch_bounds__(256)
im2col_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ col, ConvGeom g) {
  const int cv = g.C >> 3;
  const long long total = (long long)g.N * g.Ho * g.Wo * g.kh * g.kw * cv;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    int c8 = (int)(t % cv);
    long long u = t / cv;
    const int s = (int)(u % g.kw); u /= g.kw;
    const int r = (int)(u % g.kh); u /= g.kh;
    const int wo = (int)(u % g.Wo); u /= g.Wo;
    const int ho = (int)(u % g.Ho);
    const int n = (int)(u / g.Ho);
    const int hi = ho * g.stride - g.pad + r, wi = wo * g.stride - g.pad + s;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (hi >= 0 && hi < g.H && wi >= 0 && wi < g.W)
      v = *reinterpret_cast<const uint4*>(x + (((long long)n * g.H + hi) * g.W + wi) * g.C + c8 * 8);
    reinterpret_cast<uint4*>(col)[t] = v;
  }
}

// dx[n,hi,wi,c] = sum over taps (r,s) with (hi+pad-r) % stride == 0 ... of
//                 dcol[(n,ho,wo),(r,s,c)]   (+ dx_add)
__global__ void __launch_bounds__(256)
col2im_kernel(const __nv_bfloat16* __restrict__ dcol, const __nv_bfloat16* __restrict__ dx_add,
              __nv_bfloat16* __restrict__ dx, ConvGeom g) {
  const int cv = g.C >> 3;
  const long long total = (long long)g.N * g.H * g.W * cv;
  const long long krow = (long long)g.kh * g.kw * g.C;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int c8 = (int)(t % cv);
    long long u = t / cv;
    const int wi = (int)(u % g.W); u /= g.W;
    const int hi = (int)(u % g.H);
    const int n = (int)(u / g.H);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int r = 0; r < g.kh; ++r) {
      const int hn = hi + g.pad - r;
      if (hn < 0 || hn % g.stride) continue;
      const int ho = hn / g.stride;
      if (ho >= g.Ho) continue;
      for (int s = 0; s < g.kw; ++s) {
        const int wn = wi + g.pad - s;
        if (wn < 0 || wn % g.stride) continue;
        const int wo = wn / g.stride;
        if (wo >= g.Wo) continue;
        float f[8];
        bf16x8_to_float(*reinterpret_cast<const uint4*>(
                            dcol + (((long long)n * g.Ho + ho) * g.Wo + wo) * krow +
                            ((long long)r * g.kw + s) * g.C + c8 * 8), f);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] += f[e];
      }
    }
    if (dx_add != nullptr) {
      float f[8];
      bf16x8_to_float(reinterpret_cast<const uint4*>(dx_add)[t], f);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] += f[e];
    }
    reinterpret_cast<uint4*>(dx)[t] = float_to_bf16x8(acc);
  }
}

// ------------------------------------------------------------------ BatchNorm
// Column statistics over rows of a [rows, C] bf16 matrix: sums[c] += x, sumsq[c] += x^2
// (fp32, zeroed by the caller).  Block = 64 channels x rows_per_block rows.
__global__ void __launch_bounds__(256)
bn_stats_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ sums,
                float* __restrict__ sumsq, long long rows, int C, int rows_per_block) {
  __shared__ float r1[8][65], r2[8][65];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 64 + lane * 2;
  const long long r0 = (long long)blockIdx.y * rows_per_block;
  const long long re = min(rows, r0 + rows_per_block);
  float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
  if (c < C) {
    for (long long r = r0 + warp; r < re; r += 8) {
      const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(x + r * C + c);
      const float v0 = __low2float(v), v1 = __high2float(v);
      a0 += v0; a1 += v1; b0 += v0 * v0; b1 += v1 * v1;
    }
  }
  r1[warp][lane * 2] = a0; r1[warp][lane * 2 + 1] = a1;
  r2[warp][lane * 2] = b0; r2[warp][lane * 2 + 1] = b1;
  __syncthreads();
  if (threadIdx.x < 64 && blockIdx.x * 64 + threadIdx.x < C) {
    float s = 0.f, q = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) { s += r1[w][threadIdx.x]; q += r2[w][threadIdx.x]; }
    atomicAdd(sums + blockIdx.x * 64 + threadIdx.x, s);
    atomicAdd(sumsq + blockIdx.x * 64 + threadIdx.x, q);
  }
}

// y = relu?( (x - mean) * rstd * gamma + beta (+ residual) ); mean/rstd from sums.
__global__ void __launch_bounds__(256)
bn_apply_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ sums,
                const float* __restrict__ sumsq, const float* __restrict__ gamma,
                const float* __restrict__ beta, const __nv_bfloat16* __restrict__ residual,
                __nv_bfloat16* __restrict__ y, long long rows, int C, float eps, int relu) {
  const int cv = C >> 3;
  const long long total = rows * cv;
  const float inv_n = 1.f / (float)rows;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int c0 = (int)(t % cv) * 8;
    float f[8], o[8];
    bf16x8_to_float(reinterpret_cast<const uint4*>(x)[t], f);
    float rs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (residual) bf16x8_to_float(reinterpret_cast<const uint4*>(residual)[t], rs);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float m = sums[c0 + e] * inv_n;
      const float var = fmaxf(sumsq[c0 + e] * inv_n - m * m, 0.f);
      float v = (f[e] - m) * rsqrtf(var + eps) * gamma[c0 + e] + beta[c0 + e] + rs[e];
      o[e] = relu ? fmaxf(v, 0.f) : v;
    }
    reinterpret_cast<uint4*>(y)[t] = float_to_bf16x8(o);
  }
}

// Backward reductions: with g = dy * (y > 0 if relu), xhat = (x - mean) * rstd:
//   dsum[c] += g, dsumx[c] += g * xhat.   (fp32 scratch, zeroed by the caller)
__global__ void __launch_bounds__(256)
bn_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                     const __nv_bfloat16* __restrict__ y, const float* __restrict__ sums,
                     const float* __restrict__ sumsq, float* __restrict__ dsum,
                     float* __restrict__ dsumx, long long rows, int C, float eps, int relu,
                     int rows_per_block) {
  __shared__ float r1[8][65], r2[8][65];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 64 + lane * 2;
  const long long r0 = (long long)blockIdx.y * rows_per_block;
  const long long re = min(rows, r0 + rows_per_block);
  float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
  if (c < C) {
    const float inv_n = 1.f / (float)rows;
    const float m0 = sums[c] * inv_n, m1 = sums[c + 1] * inv_n;
    const float s0 = rsqrtf(fmaxf(sumsq[c] * inv_n - m0 * m0, 0.f) + eps);
    const float s1 = rsqrtf(fmaxf(sumsq[c + 1] * inv_n - m1 * m1, 0.f) + eps);
    for (long long r = r0 + warp; r < re; r += 8) {
      const __nv_bfloat162 d = *reinterpret_cast<const __nv_bfloat162*>(dy + r * C + c);
      const __nv_bfloat162 xv = *reinterpret_cast<const __nv_bfloat162*>(x + r * C + c);
      float g0 = __low2float(d), g1 = __high2float(d);
      if (relu) {
        const __nv_bfloat162 yv = *reinterpret_cast<const __nv_bfloat162*>(y + r * C + c);
        if (!(__low2float(yv) > 0.f)) g0 = 0.f;
        if (!(__high2float(yv) > 0.f)) g1 = 0.f;
      }
      a0 += g0; a1 += g1;
      b0 += g0 * (__low2float(xv) - m0) * s0;
      b1 += g1 * (__high2float(xv) - m1) * s1;
    }
  }
  r1[warp][lane * 2] = a0; r1[warp][lane * 2 + 1] = a1;
  r2[warp][lane * 2] = b0; r2[warp][lane * 2 + 1] = b1;
  __syncthreads();
  if (threadIdx.x < 64 && blockIdx.x * 64 + threadIdx.x < C) {
    float s = 0.f, q = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) { s += r1[w][threadIdx.x]; q += r2[w][threadIdx.x]; }
    atomicAdd(dsum + blockIdx.x * 64 + threadIdx.x, s);
    atomicAdd(dsumx + blockIdx.x * 64 + threadIdx.x, q);
  }
}

// dx = gamma * rstd * (g - dsum/n - xhat * dsumx/n); optionally dres = g (gradient of the
// residual branch, same masked value).  dgamma += alpha * dsumx, dbeta += alpha * dsum.
__global__ void __launch_bounds__(256)
bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                    const __nv_bfloat16* __restrict__ y, const float* __restrict__ sums,
                    const float* __restrict__ sumsq, const float* __restrict__ gamma,
                    const float* __restrict__ dsum, const float* __restrict__ dsumx,
                    __nv_bfloat16* __restrict__ dx, __nv_bfloat16* __restrict__ dres,
                    float* __restrict__ dgamma, float* __restrict__ dbeta, long long rows, int C,
                    float eps, int relu, float alpha) {
  const int cv = C >> 3;
  const long long total = rows * cv;
  const float inv_n = 1.f / (float)rows;
  if (blockIdx.x == 0) {
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      if (dgamma) red_add_f32(dgamma + c, alpha * dsumx[c]);
      if (dbeta) red_add_f32(dbeta + c, alpha * dsum[c]);
    }
  }
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int c0 = (int)(t % cv) * 8;
    float g[8], xv[8], yv[8], o[8];
    bf16x8_to_float(reinterpret_cast<const uint4*>(dy)[t], g);
    bf16x8_to_float(reinterpret_cast<const uint4*>(x)[t], xv);
    if (relu) {
      bf16x8_to_float(reinterpret_cast<const uint4*>(y)[t], yv);
#pragma unroll
      for (int e = 0; e < 8; ++e) if (!(yv[e] > 0.f)) g[e] = 0.f;
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float m = sums[c0 + e] * inv_n;
      const float rstd = rsqrtf(fmaxf(sumsq[c0 + e] * inv_n - m * m, 0.f) + eps);
      const float xh = (xv[e] - m) * rstd;
      o[e] = gamma[c0 + e] * rstd * (g[e] - dsum[c0 + e] * inv_n - xh * dsumx[c0 + e] * inv_n);
    }
    reinterpret_cast<uint4*>(dx)[t] = float_to_bf16x8(o);
    if (dres) reinterpret_cast<uint4*>(dres)[t] = float_to_bf16x8(g);
  }
}

// ---------------------------------------------------------------------------
// Bandwidth forms of the four BatchNorm kernels (used whenever C/8 divides 256, i.e. every
// ResNet width).  A thread owns ONE 8-channel vector for the whole kernel (its column never
// changes: 256 threads = (256 / cv) rows x cv vectors, and every stride is a multiple of that),
// so the per-channel constants (mean, rstd, gamma*rstd, ...) live in registers instead of being
// re-derived from five global arrays for every 16 bytes moved, and consecutive threads read
// consecutive 16-byte vectors.  Four rows are in flight per thread and stream.  ResNet-50 spends
// half of its step in these kernels (profiles/resnet50_launches.md); the per-element forms above
// ran at 0.26-0.49 of the HBM copy rate.
// ---------------------------------------------------------------------------
constexpr int BN_UNROLL = 4;

__device__ __forceinline__ void bn_block_reduce16(float (&acc)[16], float* smem, float* out_a,
                                                  float* out_b, int cv, int C) {
  // smem: [16][256]; channel c = 8 * cvec + e sums the threads rsub * cv + cvec
#pragma unroll
  for (int k = 0; k < 16; ++k) smem[k * 256 + threadIdx.x] = acc[k];
  __syncthreads();
  const int rpi = 256 / cv;
  for (int ch = threadIdx.x; ch < C; ch += 256) {
    const int cvec = ch >> 3, e = ch & 7;
    float a = 0.f, b = 0.f;
    for (int r = 0; r < rpi; ++r) {
      a += smem[e * 256 + r * cv + cvec];
      b += smem[(8 + e) * 256 + r * cv + cvec];
    }
    atomicAdd(out_a + ch, a);
    atomicAdd(out_b + ch, b);
  }
}

__global__ void __launch_bounds__(256)
bn_stats_fast_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ sums,
                     float* __restrict__ sumsq, long long rows, int C) {
  __shared__ float smem[16 * 256];
  const int cv = C >> 3, rpi = 256 / cv;
  const int cvec = threadIdx.x % cv, rsub = threadIdx.x / cv;
  const long long rstride = (long long)gridDim.x * rpi;
  const uint4* xp = reinterpret_cast<const uint4*>(x);
  float acc[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) acc[k] = 0.f;
  for (long long r = (long long)blockIdx.x * rpi + rsub; r < rows; r += BN_UNROLL * rstride) {
    uint4 v[BN_UNROLL];
#pragma unroll
    for (int u = 0; u < BN_UNROLL; ++u)
      if (r + u * rstride < rows) v[u] = xp[(r + u * rstride) * cv + cvec];
#pragma unroll
    for (int u = 0; u < BN_UNROLL; ++u)
      if (r + u * rstride < rows) {
        float f[8];
        bf16x8_to_float(v[u], f);
#pragma unroll
        for (int e = 0; e < 8; ++e) { acc[e] += f[e]; acc[8 + e] += f[e] * f[e]; }
      }
  }
  bn_block_reduce16(acc, smem, sums, sumsq, cv, C);
}

__global__ void __launch_bounds__(256)
bn_apply_fast_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ sums,
                     const float* __restrict__ sumsq, const float* __restrict__ gamma,
                     const float* __restrict__ beta, const __nv_bfloat16* __restrict__ residual,
                     __nv_bfloat16* __restrict__ y, long long rows, int C, float eps, int relu) {
  const int cv = C >> 3, rpi = 256 / cv;
  const int cvec = threadIdx.x % cv, rsub = threadIdx.x / cv;
  const long long rstride = (long long)gridDim.x * rpi;
  const float inv_n = 1.f / (float)rows;
  float sc[8], sh[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int c = cvec * 8 + e;
    const float m = sums[c] * inv_n;
    const float rstd = rsqrtf(fmaxf(sumsq[c] * inv_n - m * m, 0.f) + eps);
    sc[e] = rstd * gamma[c];
    sh[e] = beta[c] - m * sc[e];
  }
  const uint4* xp = reinterpret_cast<const uint4*>(x);
  const uint4* rp = reinterpret_cast<const uint4*>(residual);
  uint4* yp = reinterpret_cast<uint4*>(y);
  for (long long r = (long long)blockIdx.x * rpi + rsub; r < rows; r += BN_UNROLL * rstride) {
    uint4 v[BN_UNROLL], rs[BN_UNROLL];
#pragma unroll
    for (int u = 0; u < BN_UNROLL; ++u)
      if (r + u * rstride < rows) {
        const long long t = (r + u * rstride) * cv + cvec;
        v[u] = xp[t];
        if (residual) rs[u] = rp[t];
      }
#pragma unroll
    for (int u = 0; u < BN_UNROLL; ++u)
      if (r + u * rstride < rows) {
        float f[8], o[8], q[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        bf16x8_to_float(v[u], f);
        if (residual) bf16x8_to_float(rs[u], q);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          // (x - m) * rstd * gamma + beta folded into one fma per element
          const float val = f[e] * sc[e] + sh[e] + q[e];
          o[e] = relu ? fmaxf(val, 0.f) : val;
        }
        yp[(r + u * rstride) * cv + cvec] = float_to_bf16x8(o);
      }
  }
}

__global__ void __launch_bounds__(256)
bn_bwd_reduce_fast_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                          const __nv_bfloat16* __restrict__ y, const float* __restrict__ sums,
                          const float* __restrict__ sumsq, float* __restrict__ dsum,
                          float* __restrict__ dsumx, long long rows, int C, float eps, int relu) {
  __shared__ float smem[16 * 256];
  const int cv = C >> 3, rpi = 256 / cv;
  const int cvec = threadIdx.x % cv, rsub = threadIdx.x / cv;
  const long long rstride = (long long)gridDim.x * rpi;
  const float inv_n = 1.f / (float)rows;
  float mean[8], rstd[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int c = cvec * 8 + e;
    mean[e] = sums[c] * inv_n;
    rstd[e] = rsqrtf(fmaxf(sumsq[c] * inv_n - mean[e] * mean[e], 0.f) + eps);
  }
  const uint4* dp = reinterpret_cast<const uint4*>(dy);
  const uint4* xp = reinterpret_cast<const uint4*>(x);
  const uint4* yp = reinterpret_cast<const uint4*>(y);
  float acc[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) acc[k] = 0.f;
  for (long long r = (long long)blockIdx.x * rpi + rsub; r < rows; r += BN_UNROLL * rstride) {
    uint4 vd[BN_UNROLL], vx[BN_UNROLL], vy[BN_UNROLL];
#pragma unroll
    for (int u = 0; u < BN_UNROLL; ++u)
      if (r + u * rstride < rows) {
        const long long t = (r + u * rstride) * cv + cvec;
        vd[u] = dp[t]; vx[u] = xp[t];
        if (relu) vy[u] = yp[t];
      }
#pragma unroll
    for (int u = 0; u < BN_UNROLL; ++u)
      if (r + u * rstride < rows) {
        float g[8], xv[8], yv[8];
        bf16x8_to_float(vd[u], g);
        bf16x8_to_float(vx[u], xv);
        if (relu) {
          bf16x8_to_float(vy[u], yv);
#pragma unroll
          for (int e = 0; e < 8; ++e) if (!(yv[e] > 0.f)) g[e] = 0.f;
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          acc[e] += g[e];
          acc[8 + e] += g[e] * (xv[e] - mean[e]) * rstd[e];
        }
      }
  }
  bn_block_reduce16(acc, smem, dsum, dsumx, cv, C);
}

__global__ void __launch_bounds__(256)
bn_bwd_apply_fast_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                         const __nv_bfloat16* __restrict__ y, const float* __restrict__ sums,
                         const float* __restrict__ sumsq, const float* __restrict__ gamma,
                         const float* __restrict__ dsum, const float* __restrict__ dsumx,
                         __nv_bfloat16* __restrict__ dx, __nv_bfloat16* __restrict__ dres,
                         float* __restrict__ dgamma, float* __restrict__ dbeta, long long rows,
                         int C, float eps, int relu, float alpha) {
  const int cv = C >> 3, rpi = 256 / cv;
  const int cvec = threadIdx.x % cv, rsub = threadIdx.x / cv;
  const long long rstride = (long long)gridDim.x * rpi;
  const float inv_n = 1.f / (float)rows;
  if (blockIdx.x == 0) {
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      if (dgamma) red_add_f32(dgamma + c, alpha * dsumx[c]);
      if (dbeta) red_add_f32(dbeta + c, alpha * dsum[c]);
    }
  }
  // dx = k1 * (g - ds - xhat * dsx), xhat = (x - mean) * rstd, folded per channel into
  // dx = A * g + B * x + K  (three constants in registers instead of five)
  float A[8], B[8], K[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int c = cvec * 8 + e;
    const float mean = sums[c] * inv_n;
    const float rstd = rsqrtf(fmaxf(sumsq[c] * inv_n - mean * mean, 0.f) + eps);
    A[e] = gamma[c] * rstd;
    B[e] = -A[e] * rstd * (dsumx[c] * inv_n);
    K[e] = -A[e] * (dsum[c] * inv_n) - B[e] * mean;
  }
  const uint4* dp = reinterpret_cast<const uint4*>(dy);
  const uint4* xp = reinterpret_cast<const uint4*>(x);
  const uint4* yp = reinterpret_cast<const uint4*>(y);
  uint4* op = reinterpret_cast<uint4*>(dx);
  uint4* rp = reinterpret_cast<uint4*>(dres);
  for (long long r = (long long)blockIdx.x * rpi + rsub; r < rows; r += BN_UNROLL * rstride) {
    uint4 vd[BN_UNROLL], vx[BN_UNROLL], vy[BN_UNROLL];
#pragma unroll
    for (int u = 0; u < BN_UNROLL; ++u)
      if (r + u * rstride < rows) {
        const long long t = (r + u * rstride) * cv + cvec;
        vd[u] = dp[t]; vx[u] = xp[t];
        if (relu) vy[u] = yp[t];
      }
#pragma unroll
    for (int u = 0; u < BN_UNROLL; ++u)
      if (r + u * rstride < rows) {
        const long long t = (r + u * rstride) * cv + cvec;
        float g[8], xv[8], yv[8], o[8];
        bf16x8_to_float(vd[u], g);
        bf16x8_to_float(vx[u], xv);
        if (relu) {
          bf16x8_to_float(vy[u], yv);
#pragma unroll
          for (int e = 0; e < 8; ++e) if (!(yv[e] > 0.f)) g[e] = 0.f;
        }
#pragma unroll
        for (int e = 0; e < 8; ++e)
          o[e] = A[e] * g[e] + B[e] * xv[e] + K[e];
        op[t] = float_to_bf16x8(o);
        if (dres) rp[t] = float_to_bf16x8(g);
      }
  }
}

// -------------------------------------------------------------------- pooling
__global__ void __launch_bounds__(256)
maxpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, ConvGeom g) {
  const int cv = g.C >> 3;
  const long long total = (long long)g.N * g.Ho * g.Wo * cv;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int c8 = (int)(t % cv);
    long long u = t / cv;
    const int wo = (int)(u % g.Wo); u /= g.Wo;
    const int ho = (int)(u % g.Ho);
    const int n = (int)(u / g.Ho);
    float m[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) m[e] = -INFINITY;
    for (int r = 0; r < g.kh; ++r)
      for (int s = 0; s < g.kw; ++s) {
        const int hi = ho * g.stride - g.pad + r, wi = wo * g.stride - g.pad + s;
        if (hi < 0 || hi >= g.H || wi < 0 || wi >= g.W) continue;
        float f[8];
        bf16x8_to_float(*reinterpret_cast<const uint4*>(
                            x + (((long long)n * g.H + hi) * g.W + wi) * g.C + c8 * 8), f);
#pragma unroll
        for (int e = 0; e < 8; ++e) m[e] = fmaxf(m[e], f[e]);
      }
    reinterpret_cast<uint4*>(y)[t] = float_to_bf16x8(m);
  }
}

// dx[pixel] = sum over windows containing it of dy[window] * (x[pixel] == y[window])
__global__ void __launch_bounds__(256)
maxpool_bwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ y,
                   const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx, ConvGeom g) {
  const int cv = g.C >> 3;
  const long long total = (long long)g.N * g.H * g.W * cv;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int c8 = (int)(t % cv);
    long long u = t / cv;
    const int wi = (int)(u % g.W); u /= g.W;
    const int hi = (int)(u % g.H);
    const int n = (int)(u / g.H);
    float xv[8], acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    bf16x8_to_float(reinterpret_cast<const uint4*>(x)[t], xv);
    for (int r = 0; r < g.kh; ++r) {
      const int hn = hi + g.pad - r;
      if (hn < 0 || hn % g.stride) continue;
      const int ho = hn / g.stride;
      if (ho >= g.Ho) continue;
      for (int s = 0; s < g.kw; ++s) {
        const int wn = wi + g.pad - s;
        if (wn < 0 || wn % g.stride) continue;
        const int wo = wn / g.stride;
        if (wo >= g.Wo) continue;
        const long long o = (((long long)n * g.Ho + ho) * g.Wo + wo) * g.C + c8 * 8;
        float yv[8], dv[8];
        bf16x8_to_float(*reinterpret_cast<const uint4*>(y + o), yv);
        bf16x8_to_float(*reinterpret_cast<const uint4*>(dy + o), dv);
#pragma unroll
        for (int e = 0; e < 8; ++e) if (xv[e] == yv[e]) acc[e] += dv[e];
      }
    }
    reinterpret_cast<uint4*>(dx)[t] = float_to_bf16x8(acc);
  }
}

// global average pool: y[n, c] = mean_{hw} x[n, hw, c];  bwd: dx[n, hw, c] = dy[n, c] / HW
__global__ void __launch_bounds__(256)
avgpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, int N, int HW,
                   int C) {
  const long long total = (long long)N * C;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(t % C);
    const int n = (int)(t / C);
    float s = 0.f;
    for (int p = 0; p < HW; ++p) s += __bfloat162float(x[((long long)n * HW + p) * C + c]);
    y[t] = __float2bfloat16(s / HW);
  }
}
__global__ void __launch_bounds__(256)
avgpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx, int N, int HW,
                   int C) {
  const long long total = (long long)N * HW * C;
  const float inv = 1.f / HW;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(t % C);
    const int n = (int)(t / ((long long)HW * C));
    dx[t] = __float2bfloat16(__bfloat162float(dy[(long long)n * C + c]) * inv);
  }
}

static int grid_for(long long work, int threads) {
  long long b = (work + threads - 1) / threads;
  if (b < 1) b = 1;
  if (b > 148 * 16) b = 148 * 16;
  return (int)b;
}

}  // namespace dtg

using namespace dtg;

static ConvGeom geom(int N, int H, int W, int C, int kh, int kw, int stride, int pad) {
  ConvGeom g;
  g.N = N; g.H = H; g.W = W; g.C = C; g.kh = kh; g.kw = kw; g.stride = stride; g.pad = pad;
  g.Ho = (H + 2 * pad - kh) / stride + 1;
  g.Wo = (W + 2 * pad - kw) / stride + 1;
  return g;
}

#define DTG_NEED_C8(name)                                                          \
  if (C <= 0 || (C & 7)) {                                                         \
    dtg_set_last_error(name, "channel count must be a positive multiple of 8");   \
    return ERR_BAD_ARG;                                                            \
  }

extern "C" int dtg_im2col(const void* x, void* col, int N, int H, int W, int C, int kh, int kw,
                          int stride, int pad, void* stream) {
  DTG_NEED_C8("dtg_im2col")
  ConvGeom g = geom(N, H, W, C, kh, kw, stride, pad);
  const long long work = (long long)N * g.Ho * g.Wo * kh * kw * (C >> 3);
  im2col_kernel<<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, (__nv_bfloat16*)col, g);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_col2im(const void* dcol, const void* dx_add, void* dx, int N, int H, int W, int C,
                          int kh, int kw, int stride, int pad, void* stream) {
  DTG_NEED_C8("dtg_col2im")
  ConvGeom g = geom(N, H, W, C, kh, kw, stride, pad);
  const long long work = (long long)N * H * W * (C >> 3);
  col2im_kernel<<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dcol, (const __nv_bfloat16*)dx_add, (__nv_bfloat16*)dx, g);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

static bool bn_fast_ok(int C) {
  static int allow = -1;
  if (allow < 0) { const char* e = getenv("DTG_BN_FAST"); allow = (e && e[0] == '0') ? 0 : 1; }
  const int cv = C >> 3;
  return allow && cv >= 1 && cv <= 256 && (256 % cv) == 0;
}
static int bn_fast_grid(long long rows, int C) {
  const int rpi = 256 / (C >> 3);
  const long long want = (rows + (long long)rpi * dtg::BN_UNROLL - 1) / ((long long)rpi * dtg::BN_UNROLL);
  const long long cap = 148 * 8;
  return (int)(want < 1 ? 1 : (want < cap ? want : cap));
}

extern "C" int dtg_bn_stats(const void* x, float* sums, float* sumsq, long long rows, int C,
                            void* stream) {
  DTG_NEED_C8("dtg_bn_stats")
  if (bn_fast_ok(C)) {
    bn_stats_fast_kernel<<<bn_fast_grid(rows, C), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)x, sums, sumsq, rows, C);
    DTG_CUDA_OK(cudaGetLastError());
    return OK;
  }
  const int rpb = 512;
  dim3 grid((C + 63) / 64, (unsigned)((rows + rpb - 1) / rpb));
  bn_stats_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, sums, sumsq,
                                                          rows, C, rpb);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_bn_apply(const void* x, const float* sums, const float* sumsq,
                            const float* gamma, const float* beta, const void* residual, void* y,
                            long long rows, int C, float eps, int relu, void* stream) {
  DTG_NEED_C8("dtg_bn_apply")
  if (bn_fast_ok(C)) {
    bn_apply_fast_kernel<<<bn_fast_grid(rows, C), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)x, sums, sumsq, gamma, beta, (const __nv_bfloat16*)residual,
        (__nv_bfloat16*)y, rows, C, eps, relu);
    DTG_CUDA_OK(cudaGetLastError());
    return OK;
  }
  bn_apply_kernel<<<grid_for(rows * (C >> 3), 256), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, sums, sumsq, gamma, beta, (const __nv_bfloat16*)residual,
      (__nv_bfloat16*)y, rows, C, eps, relu);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_bn_bwd(const void* dy, const void* x, const void* y, const float* sums,
                          const float* sumsq, const float* gamma, float* dsum, float* dsumx,
                          void* dx, void* dres, float* dgamma, float* dbeta, long long rows, int C,
                          float eps, int relu, float alpha, void* stream) {
  DTG_NEED_C8("dtg_bn_bwd")
  cudaStream_t st = (cudaStream_t)stream;
  DTG_CUDA_OK(cudaMemsetAsync(dsum, 0, C * sizeof(float), st));
  DTG_CUDA_OK(cudaMemsetAsync(dsumx, 0, C * sizeof(float), st));
  if (bn_fast_ok(C)) {
    const int grid = bn_fast_grid(rows, C);
    bn_bwd_reduce_fast_kernel<<<grid, 256, 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x,
                                                    (const __nv_bfloat16*)y, sums, sumsq, dsum, dsumx,
                                                    rows, C, eps, relu);
    bn_bwd_apply_fast_kernel<<<grid, 256, 0, st>>>(
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)y, sums, sumsq,
        gamma, dsum, dsumx, (__nv_bfloat16*)dx, (__nv_bfloat16*)dres, dgamma, dbeta, rows, C, eps,
        relu, alpha);
    DTG_CUDA_OK(cudaGetLastError());
    return OK;
  }
  const int rpb = 512;
  dim3 grid((C + 63) / 64, (unsigned)((rows + rpb - 1) / rpb));
  bn_bwd_reduce_kernel<<<grid, 256, 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x,
                                             (const __nv_bfloat16*)y, sums, sumsq, dsum, dsumx,
                                             rows, C, eps, relu, rpb);
  bn_bwd_apply_kernel<<<grid_for(rows * (C >> 3), 256), 256, 0, st>>>(
      (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)y, sums, sumsq,
      gamma, dsum, dsumx, (__nv_bfloat16*)dx, (__nv_bfloat16*)dres, dgamma, dbeta, rows, C, eps,
      relu, alpha);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_maxpool_fwd(const void* x, void* y, int N, int H, int W, int C, int k,
                               int stride, int pad, void* stream) {
  DTG_NEED_C8("dtg_maxpool_fwd")
  ConvGeom g = geom(N, H, W, C, k, k, stride, pad);
  maxpool_fwd_kernel<<<grid_for((long long)N * g.Ho * g.Wo * (C >> 3), 256), 256, 0,
                       (cudaStream_t)stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, g);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_maxpool_bwd(const void* x, const void* y, const void* dy, void* dx, int N, int H,
                               int W, int C, int k, int stride, int pad, void* stream) {
  DTG_NEED_C8("dtg_maxpool_bwd")
  ConvGeom g = geom(N, H, W, C, k, k, stride, pad);
  maxpool_bwd_kernel<<<grid_for((long long)N * H * W * (C >> 3), 256), 256, 0,
                       (cudaStream_t)stream>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)y,
                                               (const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, g);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_avgpool_fwd(const void* x, void* y, int N, int HW, int C, void* stream) {
  avgpool_fwd_kernel<<<grid_for((long long)N * C, 256), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, (__nv_bfloat16*)y, N, HW, C);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream) {
  avgpool_bwd_kernel<<<grid_for((long long)N * HW * C, 256), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, N, HW, C);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}
