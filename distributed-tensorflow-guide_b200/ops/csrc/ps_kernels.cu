// Parameter-server data-plane kernels for sm_100a (work list
// W5-W8, W11).  "ps" memory is an HBM shard that may live on a peer GPU and is mapped
// into this process over NVLink (parallel/fabric.py): the kernels below address it
// with ordinary global pointers, pushes are fire-and-forget fp32 reductions
// (red.global.add, 16 B vectors), pulls are streaming peer loads fused with the
// fp32 -> bf16 conversion the tensor-core GEMMs need.
#include "common.cuh"
#include <algorithm>

namespace dtg {

constexpr int EW_THREADS = 512;
// The kernels that move parameters between GPUs (pull, push) run on side streams WHILE the
// persistent GEMM kernels hold every SM with one 320-thread CTA (168 registers per thread, all
// but 2 KB of shared memory).  What is left on an SM is 11 776 registers and one block slot: a
// 512-thread block never fits next to a GEMM CTA (it would only run in the gaps between GEMM
// kernels, i.e. not overlap at all), a 128-thread block with <= 88 registers does.
constexpr int COMM_THREADS = 128;

__device__ __forceinline__ long long gtid() {
  return (long long)blockIdx.x * blockDim.x + threadIdx.x;
}
__device__ __forceinline__ long long gstride() { return (long long)gridDim.x * blockDim.x; }

// ---------------------------------------------------------------- W7: pull
// dst_bf16[i] = bf16(src[i]); optionally dst_f32[i] = src[i].  src is the ps shard
// (peer or local).  4 independent 16 B loads in flight per thread.
__global__ void __launch_bounds__(COMM_THREADS)
pull_convert_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst_bf16,
                    float* __restrict__ dst_f32, long long n) {
  const long long n4 = n >> 2;
  const float4* s4 = reinterpret_cast<const float4*>(src);
  const long long stride = gstride();
  long long i = gtid();
  for (; i + 3 * stride < n4; i += 4 * stride) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = ld_v4_relaxed_sys(s4 + i + u * stride);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long j = i + u * stride;
      if (dst_bf16)
        reinterpret_cast<uint2*>(dst_bf16)[j] =
            make_uint2(pack_bf16x2(v[u].x, v[u].y), pack_bf16x2(v[u].z, v[u].w));
      if (dst_f32) reinterpret_cast<float4*>(dst_f32)[j] = v[u];
    }
  }
  for (; i < n4; i += stride) {
    float4 v = ld_v4_relaxed_sys(s4 + i);
    if (dst_bf16)
      reinterpret_cast<uint2*>(dst_bf16)[i] =
          make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
    if (dst_f32) reinterpret_cast<float4*>(dst_f32)[i] = v;
  }
  for (long long t = (n4 << 2) + gtid(); t < n; t += stride) {
    float v = src[t];
    if (dst_bf16) dst_bf16[t] = __float2bfloat16(v);
    if (dst_f32) dst_f32[t] = v;
  }
}

// ------------------------------------------------- K3 / W6: push, SGD on the ps
// w_ps[i] += alpha * g[i]   (alpha = -lr, or -lr/T for ADAG's window mean).  Unlocked.
__global__ void __launch_bounds__(COMM_THREADS)
push_axpy_kernel(float* __restrict__ w_ps, const float* __restrict__ g, float alpha,
                 long long n) {
  const long long n4 = n >> 2;
  const long long stride = gstride();
  for (long long i = gtid(); i < n4; i += stride) {
    float4 v = ld_nc_v4(reinterpret_cast<const float4*>(g) + i);
    red_add_v4_f32(w_ps + 4 * i, alpha * v.x, alpha * v.y, alpha * v.z, alpha * v.w);
  }
  for (long long t = (n4 << 2) + gtid(); t < n; t += stride) red_add_f32(w_ps + t, alpha * g[t]);
}

// ------------------------------------------- K4 / W5: push, Adagrad on the ps
// acc_ps += g^2 (atomic, returns the old value); w_ps -= lr * g * rsqrt(acc_new).
// Lock-free: two workers interleaving see each other's accumulator increments, the
// weight update itself is a commutative add.
__global__ void __launch_bounds__(COMM_THREADS)
push_adagrad_kernel(float* __restrict__ w_ps, float* __restrict__ acc_ps,
                    const float* __restrict__ g, float lr, float gscale, long long n) {
  const long long stride = gstride();
  for (long long i = gtid(); i < n; i += stride) {
    const float gv = gscale * g[i];
    const float old = atom_add_f32_sys(acc_ps + i, gv * gv);
    red_add_f32(w_ps + i, -lr * gv * rsqrtf(old + gv * gv));
  }
}

// ------------------------------------- K7+K8 / W8: fused local step + window
// One pass over flat buffers: window += g; Adagrad/SGD on the fp32 local params;
// refresh the bf16 copy the next forward GEMM reads.
__global__ void __launch_bounds__(EW_THREADS)
local_step_kernel(float* __restrict__ p, __nv_bfloat16* __restrict__ p_bf16,
                  float* __restrict__ slot, float* __restrict__ window,
                  const float* __restrict__ g, float lr, int adagrad, long long n) {
  const long long n4 = n >> 2;
  const long long stride = gstride();
  for (long long i = gtid(); i < n4; i += stride) {
    float4 gv = reinterpret_cast<const float4*>(g)[i];
    float4 pv = reinterpret_cast<float4*>(p)[i];
    if (window) {
      float4 w = reinterpret_cast<float4*>(window)[i];
      w.x += gv.x; w.y += gv.y; w.z += gv.z; w.w += gv.w;
      reinterpret_cast<float4*>(window)[i] = w;
    }
    if (adagrad) {
      float4 a = reinterpret_cast<float4*>(slot)[i];
      a.x += gv.x * gv.x; a.y += gv.y * gv.y; a.z += gv.z * gv.z; a.w += gv.w * gv.w;
      reinterpret_cast<float4*>(slot)[i] = a;
      pv.x -= lr * gv.x * rsqrtf(a.x); pv.y -= lr * gv.y * rsqrtf(a.y);
      pv.z -= lr * gv.z * rsqrtf(a.z); pv.w -= lr * gv.w * rsqrtf(a.w);
    } else {
      pv.x -= lr * gv.x; pv.y -= lr * gv.y; pv.z -= lr * gv.z; pv.w -= lr * gv.w;
    }
    reinterpret_cast<float4*>(p)[i] = pv;
    if (p_bf16)
      reinterpret_cast<uint2*>(p_bf16)[i] =
          make_uint2(pack_bf16x2(pv.x, pv.y), pack_bf16x2(pv.z, pv.w));
  }
  for (long long t = (n4 << 2) + gtid(); t < n; t += stride) {
    float gv = g[t], pv = p[t];
    if (window) window[t] += gv;
    if (adagrad) { float a = slot[t] + gv * gv; slot[t] = a; pv -= lr * gv * rsqrtf(a); }
    else pv -= lr * gv;
    p[t] = pv;
    if (p_bf16) p_bf16[t] = __float2bfloat16(pv);
  }
}

// --------------------------------------------- W11: softmax cross-entropy fwd+bwd
// logits fp32 [B, ldl] (first C columns valid), labels int64.  One warp per row.
// dlogits bf16 [B, ldd] = (softmax - onehot) * scale, padded columns written as 0.
// loss_sum += sum_rows(-log p[label]) * scale  (one atomic per block).
__global__ void __launch_bounds__(256)
softmax_xent_kernel(const float* __restrict__ logits, const long long* __restrict__ labels,
                    __nv_bfloat16* __restrict__ dlogits, float* __restrict__ loss_sum, int B,
                    int C, int ldl, int ldd, float scale) {
  __shared__ float s_loss[8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  float my_loss = 0.f;
  if (row < B) {
    const float* lp = logits + (long long)row * ldl;
    float mx = -INFINITY;
    for (int c = lane; c < C; c += 32) mx = fmaxf(mx, lp[c]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int c = lane; c < C; c += 32) sum += __expf(lp[c] - mx);
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    const int y = (int)labels[row];
    __nv_bfloat16* dp = dlogits ? dlogits + (long long)row * ldd : nullptr;
    for (int c = lane; c < ldd; c += 32) {
      float d = 0.f;
      if (c < C) d = (__expf(lp[c] - mx) * inv - (c == y ? 1.f : 0.f)) * scale;
      if (dp) dp[c] = __float2bfloat16(d);
    }
    if (lane == 0) my_loss = (logf(sum) + mx - lp[y]) * scale;
  }
  if (lane == 0) s_loss[warp] = my_loss;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += s_loss[w];
    if (loss_sum) atomicAdd(loss_sum, t);
  }
}

// bf16 logits, loss and gradient with ONE read of HBM: a 512-thread block owns a row, keeps it
// in shared memory (up to 65 536 columns = 128 KB), reduces max and sum-exp across the block and
// overwrites the row with scale * (softmax - onehot).  4 bytes of HBM
// traffic per logit instead of the 14 of fp32 logits read three times plus a separate bf16
// gradient (the LM head of GPT-2 medium: 8192 x 50 304).  Columns in [C, ld) are padding: they
// are ignored on the way in and zeroed on the way out.
constexpr int XENT_THREADS = 512;
__device__ __forceinline__ float xent_block_reduce(float v, bool is_max, float* s_red, float* s_bcast) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = is_max ? warp_max(v) : warp_sum(v);
  __syncthreads();
  if (lane == 0) s_red[warp] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = s_red[0];
    for (int w = 1; w < XENT_THREADS / 32; ++w) t = is_max ? fmaxf(t, s_red[w]) : t + s_red[w];
    *s_bcast = t;
  }
  __syncthreads();
  return *s_bcast;
}

__global__ void __launch_bounds__(XENT_THREADS)
softmax_xent_bf16_inplace_kernel(__nv_bfloat16* __restrict__ logits,
                                 const long long* __restrict__ labels, float* __restrict__ loss_sum,
                                 int C, int ld, float scale) {
  extern __shared__ uint4 s_row[];          // the whole row, bf16, read from HBM exactly once
  __shared__ float s_red[XENT_THREADS / 32];
  __shared__ float s_bcast;
  const int tid = threadIdx.x;
  const long long row = blockIdx.x;
  __nv_bfloat16* rowp = logits + row * ld;
  uint4* rp = reinterpret_cast<uint4*>(rowp);
  const int nvec = ld >> 3;
  const int full_vecs = C >> 3;             // vectors with all 8 columns real; at most one is partial
  const float LOG2E = 1.4426950408889634f;
  const int y = (int)labels[row];
  const float x_label = __bfloat162float(rowp[y]);     // read before the row is overwritten
  float mx = -INFINITY;
#pragma unroll 4
  for (int j = tid; j < nvec; j += XENT_THREADS) {
    uint4 v = rp[j];
    if (j >= full_vecs) {                   // padding columns become -inf once, here
      __nv_bfloat16* e = reinterpret_cast<__nv_bfloat16*>(&v);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (j * 8 + k >= C) e[k] = __float2bfloat16(-INFINITY);
    }
    s_row[j] = v;
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 f = __bfloat1622float2(h[k]);
      mx = fmaxf(mx, fmaxf(f.x, f.y));
    }
  }
  mx = xent_block_reduce(mx, true, s_red, &s_bcast);
  const float mxs = mx * LOG2E;
  float sum = 0.f;
#pragma unroll 4
  for (int j = tid; j < nvec; j += XENT_THREADS) {
    const uint4 v = s_row[j];
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 f = __bfloat1622float2(h[k]);
      sum += exp2f(fmaf(f.x, LOG2E, -mxs)) + exp2f(fmaf(f.y, LOG2E, -mxs));   // exp2(-inf) = 0: padding
    }
  }
  sum = xent_block_reduce(sum, false, s_red, &s_bcast);
  if (tid == 0 && loss_sum) atomicAdd(loss_sum, (logf(sum) + mx - x_label) * scale);
  const float k1 = scale / sum;
#pragma unroll 4
  for (int j = tid; j < nvec; j += XENT_THREADS) {
    const uint4 v = s_row[j];
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
    uint32_t o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 f = __bfloat1622float2(h[k]);
      o[k] = pack_bf16x2(k1 * exp2f(fmaf(f.x, LOG2E, -mxs)), k1 * exp2f(fmaf(f.y, LOG2E, -mxs)));
    }
    rp[j] = make_uint4(o[0], o[1], o[2], o[3]);
  }
  // the one column that also carries -scale: fixed up by one thread after the row is written
  __syncthreads();
  if (tid == 0) rowp[y] = __float2bfloat16(k1 * exp2f(fmaf(x_label, LOG2E, -mxs)) - scale);
}

// ------------------------------------------------ bias gradient (column sums)
// target[n] (+)= alpha * sum_m dY[m, n]; always a reduction add so the same kernel
// serves "grad buffer" (zeroed by the caller), "window accumulator" and "push to ps"
// (alpha = -lr, target in peer HBM).
__global__ void __launch_bounds__(256)
bias_grad_kernel(const __nv_bfloat16* __restrict__ dY, float* __restrict__ target, int M, int N,
                 long long ldy, float alpha, int rows_per_block) {
  // block = 64 columns (one bf16x2 per lane) x rows_per_block rows; 8 warps stride the rows
  __shared__ float red[8][65];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 64 + lane * 2;
  const int r0 = blockIdx.y * rows_per_block;
  const int r1 = min(M, r0 + rows_per_block);
  float s0 = 0.f, s1 = 0.f;
  const bool pair_ok = (c + 1 < N) && ((ldy & 1) == 0) &&
                       ((reinterpret_cast<uintptr_t>(dY) & 3) == 0);
  if (pair_ok) {
    int r = r0 + warp;
    for (; r + 24 < r1; r += 32) {       // 4 independent loads in flight
      __nv_bfloat162 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        v[u] = *reinterpret_cast<const __nv_bfloat162*>(dY + (long long)(r + 8 * u) * ldy + c);
#pragma unroll
      for (int u = 0; u < 4; ++u) { s0 += __low2float(v[u]); s1 += __high2float(v[u]); }
    }
    for (; r < r1; r += 8) {
      const __nv_bfloat162 v =
          *reinterpret_cast<const __nv_bfloat162*>(dY + (long long)r * ldy + c);
      s0 += __low2float(v); s1 += __high2float(v);
    }
  } else if (c < N) {
    for (int r = r0 + warp; r < r1; r += 8) {
      s0 += __bfloat162float(dY[(long long)r * ldy + c]);
      if (c + 1 < N) s1 += __bfloat162float(dY[(long long)r * ldy + c + 1]);
    }
  }
  red[warp][lane * 2] = s0;
  red[warp][lane * 2 + 1] = s1;
  __syncthreads();
  if (threadIdx.x < 64) {
    const int cc = blockIdx.x * 64 + threadIdx.x;
    if (cc < N) {
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) t += red[w][threadIdx.x];
      red_add_f32(target + cc, alpha * t);
    }
  }
}

// ------------------------------------------------------------------ utilities
__global__ void __launch_bounds__(EW_THREADS)
cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n) {
  const long long stride = gstride();
  for (long long i = gtid(); i < n; i += stride) dst[i] = __float2bfloat16(src[i]);
}

__global__ void counter_add_kernel(unsigned long long* ctr, unsigned long long delta) {
  // K14: global_step lives in the ps shard; atom at system scope (peer memory).
  if (threadIdx.x == 0 && blockIdx.x == 0)
    asm volatile("red.relaxed.sys.global.add.u64 [%0], %1;\n" ::"l"(ctr), "l"(delta) : "memory");
}

// Narrow heads (<= 16 classes in a 16-float row, e.g. the padded 10-way MLP head): one
// thread per row, the whole row in 4 float4 registers.
__global__ void __launch_bounds__(128)
softmax_xent_row16_kernel(const float* __restrict__ logits, const long long* __restrict__ labels,
                          __nv_bfloat16* __restrict__ dlogits, float* __restrict__ loss_sum, int B,
                          int C, long long ldd, float scale) {
  __shared__ float s_loss[4];
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  float my = 0.f;
  if (row < B) {
    float v[16];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float4 f = reinterpret_cast<const float4*>(logits + (long long)row * 16)[j];
      v[4 * j] = f.x; v[4 * j + 1] = f.y; v[4 * j + 2] = f.z; v[4 * j + 3] = f.w;
    }
    const int y = (int)labels[row];
    float mx = -INFINITY, ly = 0.f;
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      if (c >= C) v[c] = -INFINITY;
      mx = fmaxf(mx, v[c]);
      if (c == y) ly = v[c];
    }
    float sum = 0.f;
#pragma unroll
    for (int c = 0; c < 16; ++c) { v[c] = __expf(v[c] - mx); sum += v[c]; }
    const float inv = 1.f / sum;
    if (dlogits) {
      uint32_t pk[8];
#pragma unroll
      for (int c = 0; c < 16; c += 2) {
        const float d0 = (v[c] * inv - (c == y ? 1.f : 0.f)) * scale;
        const float d1 = (v[c + 1] * inv - (c + 1 == y ? 1.f : 0.f)) * scale;
        pk[c >> 1] = pack_bf16x2(c < C ? d0 : 0.f, c + 1 < C ? d1 : 0.f);
      }
      uint4* dp = reinterpret_cast<uint4*>(dlogits + (long long)row * ldd);
      dp[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      dp[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
    }
    my = (logf(sum) + mx - ly) * scale;
  }
  my = warp_sum(my);
  if ((threadIdx.x & 31) == 0) s_loss[threadIdx.x >> 5] = my;
  __syncthreads();
  if (threadIdx.x == 0 && loss_sum) atomicAdd(loss_sum, s_loss[0] + s_loss[1] + s_loss[2] + s_loss[3]);
}

static int ew_grid(long long n, int per_thread, int max_blocks, int threads = EW_THREADS) {
  long long b = (n / per_thread + threads - 1) / threads;
  if (b < 1) b = 1;
  if (b > max_blocks) b = max_blocks;
  return (int)b;
}

}  // namespace dtg

using namespace dtg;

// fp32 refresh of a few small ranges (biases, norm parameters) in one launch: blockIdx.y = range.
struct PullRanges {
  long long lo[16], hi[16];
  int n;
};
__global__ void __launch_bounds__(256)
pull_ranges_kernel(const float* __restrict__ src, float* __restrict__ dst, const PullRanges r) {
  const long long lo = r.lo[blockIdx.y], hi = r.hi[blockIdx.y];
  for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi;
       i += (long long)gridDim.x * blockDim.x) {
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];\n" : "=f"(v) : "l"(src + i) : "memory");
    dst[i] = v;
  }
}

// Zero many small fp32 tensors with ONE launch: table[2*i] = address, table[2*i+1] = element
// count (device memory, built once by the caller).  blockIdx.y = tensor.  A training step of the
// transformer zeroes ~300 bias / LayerNorm gradient vectors; one fill kernel each cost ~1 ms.
__global__ void __launch_bounds__(256)
zero_many_kernel(const unsigned long long* __restrict__ table) {
  float* p = reinterpret_cast<float*>(table[2 * blockIdx.y]);
  const long long n = (long long)table[2 * blockIdx.y + 1];
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x)
    p[i] = 0.f;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

extern "C" int dtg_pull_convert(const float* src, void* dst_bf16, float* dst_f32, long long n,
                                void* stream) {
  if (n <= 0) return OK;
  if (!aligned16(src) || (dst_bf16 && (reinterpret_cast<uintptr_t>(dst_bf16) & 7)) ||
      (dst_f32 && !aligned16(dst_f32))) {
    dtg_set_last_error("dtg_pull_convert", "buffers must be 16-byte aligned");
    return ERR_BAD_ARG;
  }
  pull_convert_kernel<<<ew_grid(n, 16, 148 * 16, COMM_THREADS), COMM_THREADS, 0, (cudaStream_t)stream>>>(
      src, reinterpret_cast<__nv_bfloat16*>(dst_bf16), dst_f32, n);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_zero_many(const void* table_dev, int count, long long longest, void* stream) {
  if (count <= 0) return OK;
  long long bx = (longest + 255) / 256;
  if (bx < 1) bx = 1;
  if (bx > 64) bx = 64;
  for (int y0 = 0; y0 < count; y0 += 65535) {
    const int ny = std::min(65535, count - y0);
    zero_many_kernel<<<dim3((unsigned)bx, (unsigned)ny), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const unsigned long long*>(table_dev) + 2 * (size_t)y0);
  }
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_pull_ranges(const float* src, float* dst_f32, const long long* lo,
                               const long long* hi, int n, void* stream) {
  if (n <= 0) return OK;
  if (n > 16) {
    dtg_set_last_error("dtg_pull_ranges", "at most 16 ranges per call");
    return ERR_TOO_LARGE;
  }
  PullRanges r;
  long long longest = 0;
  for (int i = 0; i < n; ++i) {
    r.lo[i] = lo[i]; r.hi[i] = hi[i];
    longest = std::max(longest, hi[i] - lo[i]);
  }
  r.n = n;
  dim3 grid((unsigned)std::min<long long>(64, (longest + 255) / 256 > 0 ? (longest + 255) / 256 : 1), n);
  pull_ranges_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(src, dst_f32, r);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_push_axpy(float* w_ps, const float* g, float alpha, long long n, void* stream) {
  if (n <= 0) return OK;
  if (!aligned16(w_ps) || !aligned16(g)) {
    dtg_set_last_error("dtg_push_axpy", "buffers must be 16-byte aligned");
    return ERR_BAD_ARG;
  }
  push_axpy_kernel<<<ew_grid(n, 8, 148 * 16, COMM_THREADS), COMM_THREADS, 0, (cudaStream_t)stream>>>(
      w_ps, g, alpha, n);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_push_adagrad(float* w_ps, float* acc_ps, const float* g, float lr, float gscale,
                                long long n, void* stream) {
  if (n <= 0) return OK;
  push_adagrad_kernel<<<ew_grid(n, 4, 148 * 16, COMM_THREADS), COMM_THREADS, 0, (cudaStream_t)stream>>>(
      w_ps, acc_ps, g, lr, gscale, n);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_local_step(float* p, void* p_bf16, float* slot, float* window, const float* g,
                              float lr, int adagrad, long long n, void* stream) {
  if (n <= 0) return OK;
  if (adagrad && slot == nullptr) {
    dtg_set_last_error("dtg_local_step", "adagrad needs a slot buffer");
    return ERR_BAD_ARG;
  }
  if (!aligned16(p) || !aligned16(g) || (window && !aligned16(window)) ||
      (slot && !aligned16(slot)) || (p_bf16 && (reinterpret_cast<uintptr_t>(p_bf16) & 7))) {
    dtg_set_last_error("dtg_local_step", "buffers must be 16-byte aligned");
    return ERR_BAD_ARG;
  }
  local_step_kernel<<<ew_grid(n, 8, 148 * 4), EW_THREADS, 0, (cudaStream_t)stream>>>(
      p, reinterpret_cast<__nv_bfloat16*>(p_bf16), slot, window, g, lr, adagrad, n);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_softmax_xent(const float* logits, const long long* labels, void* dlogits,
                                float* loss_sum, int B, int C, int ldl, int ldd, float scale,
                                void* stream) {
  if (B <= 0 || C <= 0 || ldl < C || (dlogits && ldd < C)) {
    dtg_set_last_error("dtg_softmax_xent", "bad shape");
    return ERR_BAD_ARG;
  }
  if (C <= 16 && ldl == 16 && (dlogits == nullptr || (ldd >= 16 && ldd % 8 == 0)) &&
      (reinterpret_cast<uintptr_t>(logits) & 15) == 0 &&
      (reinterpret_cast<uintptr_t>(dlogits) & 15) == 0) {
    // 16-wide rows: only columns [0,16) of dlogits are written (the padded head width)
    softmax_xent_row16_kernel<<<(B + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        logits, labels, reinterpret_cast<__nv_bfloat16*>(dlogits), loss_sum, B, C, ldd, scale);
    DTG_CUDA_OK(cudaGetLastError());
    return OK;
  }
  softmax_xent_kernel<<<(B + 7) / 8, 256, 0, (cudaStream_t)stream>>>(
      logits, labels, reinterpret_cast<__nv_bfloat16*>(dlogits), loss_sum, B, C, ldl,
      dlogits ? ldd : 0, scale);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_softmax_xent_bf16(void* logits_inout, const long long* labels, float* loss_sum,
                                     int B, int C, long long ld, float scale, void* stream) {
  if (B <= 0 || C <= 0) return OK;
  if ((ld & 7) != 0 || ld < C || (reinterpret_cast<uintptr_t>(logits_inout) & 15) != 0 ||
      ld > 65536) {
    dtg_set_last_error("dtg_softmax_xent_bf16",
                       "need a 16-byte aligned bf16 buffer, row pitch a multiple of 8 and <= 65536");
    return ERR_BAD_ARG;
  }
  __nv_bfloat16* lp = reinterpret_cast<__nv_bfloat16*>(logits_inout);
  cudaStream_t st = (cudaStream_t)stream;
  const int smem = (int)(ld * 2);
  static bool configured = false;
  if (!configured) {
    DTG_CUDA_OK(cudaFuncSetAttribute(softmax_xent_bf16_inplace_kernel,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, 65536 * 2));
    configured = true;
  }
  softmax_xent_bf16_inplace_kernel<<<B, XENT_THREADS, smem, st>>>(lp, labels, loss_sum, C, (int)ld, scale);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_bias_grad(const void* dY, float* target, int M, int N, long long ldy,
                             float alpha, void* stream) {
  if (M <= 0 || N <= 0) return OK;
  const int rows_per_block = 256;
  dim3 grid((N + 63) / 64, (M + rows_per_block - 1) / rows_per_block);
  bias_grad_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(dY), target, M, N, ldy, alpha, rows_per_block);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_cast_f32_bf16(const float* src, void* dst, long long n, void* stream) {
  if (n <= 0) return OK;
  cast_f32_bf16_kernel<<<ew_grid(n, 4, 148 * 8), EW_THREADS, 0, (cudaStream_t)stream>>>(
      src, reinterpret_cast<__nv_bfloat16*>(dst), n);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_counter_add(void* ctr, unsigned long long delta, void* stream) {
  counter_add_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<unsigned long long*>(ctr), delta);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}
