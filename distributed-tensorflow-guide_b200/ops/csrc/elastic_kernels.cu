// Data-plane kernels of the schemes the reference lists as TODO (README.md:40-42): Dynamic
// SGD, AEASGD, AEAMSGD.  All work on the flat fp32 buffers of parallel/flat.py; the ps side
// (`centre`, `w_ps`, the step word) may be a peer GPU's HBM reached over NVLink, so those
// accesses are .sys-scope relaxed loads and red.add -- lock-free like every other push here.
#include "common.cuh"

namespace dtg {
namespace {
constexpr int EL_THREADS = 256;
__device__ __forceinline__ long long gtid() { return (long long)blockIdx.x * blockDim.x + threadIdx.x; }
__device__ __forceinline__ long long gstride() { return (long long)gridDim.x * blockDim.x; }
inline int el_grid(long long n, int per_thread) {
  long long b = (n / per_thread + EL_THREADS - 1) / EL_THREADS;
  if (b < 1) b = 1;
  if (b > 148 * 4) b = 148 * 4;
  return (int)b;
}
inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
}  // namespace

// Elastic exchange (EASGD eq. 5-6) in ONE pass, pull and push fused:
//   e = alpha * (x - centre);  x -= e;  centre += e  (red.add: other workers' exchanges commute)
// and the bf16 mirror the next forward GEMM reads is refreshed.
__global__ void __launch_bounds__(EL_THREADS)
easgd_elastic_kernel(float* __restrict__ centre, float* __restrict__ x,
                     __nv_bfloat16* __restrict__ x16, const float* __restrict__ vel, float delta,
                     float alpha, long long n) {
  const long long n4 = n >> 2, stride = gstride();
  for (long long i = gtid(); i < n4; i += stride) {
    const float4 c = ld_v4_relaxed_sys(reinterpret_cast<const float4*>(centre) + i);
    float4 v = reinterpret_cast<float4*>(x)[i];
    const float e0 = alpha * (v.x - c.x), e1 = alpha * (v.y - c.y), e2 = alpha * (v.z - c.z),
                e3 = alpha * (v.w - c.w);
    v.x -= e0; v.y -= e1; v.z -= e2; v.w -= e3;
    reinterpret_cast<float4*>(x)[i] = v;
    red_add_v4_f32(centre + 4 * i, e0, e1, e2, e3);
    if (x16 != nullptr) {
      if (vel != nullptr) {       // EAMSGD: the mirror holds the Nesterov look-ahead x + delta v
        const float4 m = reinterpret_cast<const float4*>(vel)[i];
        v.x += delta * m.x; v.y += delta * m.y; v.z += delta * m.z; v.w += delta * m.w;
      }
      reinterpret_cast<uint2*>(x16)[i] = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
    }
  }
  for (long long t = (n4 << 2) + gtid(); t < n; t += stride) {
    const float c = centre[t];
    const float e = alpha * (x[t] - c);
    x[t] -= e;
    red_add_f32(centre + t, e);
    if (x16 != nullptr) x16[t] = __float2bfloat16(x[t] + (vel != nullptr ? delta * vel[t] : 0.f));
  }
}

// Nesterov momentum step of EAMSGD: v = delta*v - lr*g;  x += v;  mirror = bf16(x + delta*v)
// (the next gradient is evaluated at the look-ahead point).  delta = 0 is plain SGD.
__global__ void __launch_bounds__(EL_THREADS)
momentum_step_kernel(float* __restrict__ x, float* __restrict__ v, const float* __restrict__ g,
                     __nv_bfloat16* __restrict__ x16, float lr, float delta, long long n) {
  const long long n4 = n >> 2, stride = gstride();
  for (long long i = gtid(); i < n4; i += stride) {
    const float4 gv = ld_nc_v4(reinterpret_cast<const float4*>(g) + i);
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float4 xv = reinterpret_cast<float4*>(x)[i];
    vv.x = delta * vv.x - lr * gv.x; vv.y = delta * vv.y - lr * gv.y;
    vv.z = delta * vv.z - lr * gv.z; vv.w = delta * vv.w - lr * gv.w;
    xv.x += vv.x; xv.y += vv.y; xv.z += vv.z; xv.w += vv.w;
    reinterpret_cast<float4*>(v)[i] = vv;
    reinterpret_cast<float4*>(x)[i] = xv;
    if (x16 != nullptr)
      reinterpret_cast<uint2*>(x16)[i] =
          make_uint2(pack_bf16x2(xv.x + delta * vv.x, xv.y + delta * vv.y),
                     pack_bf16x2(xv.z + delta * vv.z, xv.w + delta * vv.w));
  }
  for (long long t = (n4 << 2) + gtid(); t < n; t += stride) {
    const float nv = delta * v[t] - lr * g[t];
    v[t] = nv;
    x[t] += nv;
    if (x16 != nullptr) x16[t] = __float2bfloat16(x[t] + delta * nv);
  }
}

// DynSGD push: w_ps += (alpha / staleness) * g with staleness = step_now - step_seen + 1 read
// on the device from the ps-resident step word (no host round trip); reports the staleness.
__global__ void __launch_bounds__(EL_THREADS)
push_axpy_dyn_kernel(float* __restrict__ w_ps, const float* __restrict__ g, float alpha,
                     const long long* __restrict__ step_word, const long long* __restrict__ seen,
                     long long* __restrict__ staleness_out, long long n) {
  long long st = *reinterpret_cast<const volatile long long*>(step_word) - *seen + 1;
  if (st < 1) st = 1;
  if (staleness_out != nullptr && gtid() == 0) *staleness_out = st;
  const float a = alpha / (float)st;
  const long long n4 = n >> 2, stride = gstride();
  for (long long i = gtid(); i < n4; i += stride) {
    const float4 v = ld_nc_v4(reinterpret_cast<const float4*>(g) + i);
    red_add_v4_f32(w_ps + 4 * i, a * v.x, a * v.y, a * v.z, a * v.w);
  }
  for (long long t = (n4 << 2) + gtid(); t < n; t += stride) red_add_f32(w_ps + t, a * g[t]);
}
}  // namespace dtg

using namespace dtg;

extern "C" int dtg_easgd_elastic(float* centre, float* x, void* x_bf16, const float* vel,
                                 float delta, float alpha, long long n, void* stream) {
  if (n <= 0) return OK;
  if (!al16(centre) || !al16(x) || (vel && !al16(vel)) ||
      (x_bf16 && (reinterpret_cast<uintptr_t>(x_bf16) & 7))) {
    dtg_set_last_error("dtg_easgd_elastic", "buffers must be 16-byte aligned");
    return ERR_BAD_ARG;
  }
  easgd_elastic_kernel<<<el_grid(n, 8), EL_THREADS, 0, (cudaStream_t)stream>>>(
      centre, x, reinterpret_cast<__nv_bfloat16*>(x_bf16), vel, delta, alpha, n);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_momentum_step(float* x, float* v, const float* g, void* x_bf16, float lr,
                                 float delta, long long n, void* stream) {
  if (n <= 0) return OK;
  if (!al16(x) || !al16(v) || !al16(g) || (x_bf16 && (reinterpret_cast<uintptr_t>(x_bf16) & 7))) {
    dtg_set_last_error("dtg_momentum_step", "buffers must be 16-byte aligned");
    return ERR_BAD_ARG;
  }
  momentum_step_kernel<<<el_grid(n, 8), EL_THREADS, 0, (cudaStream_t)stream>>>(
      x, v, g, reinterpret_cast<__nv_bfloat16*>(x_bf16), lr, delta, n);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_push_axpy_dyn(float* w_ps, const float* g, float alpha, const long long* step_word,
                                 const long long* seen, long long* staleness_out, long long n,
                                 void* stream) {
  if (n <= 0) return OK;
  if (!al16(w_ps) || !al16(g) || !step_word || !seen) {
    dtg_set_last_error("dtg_push_axpy_dyn", "null counter or unaligned buffer");
    return ERR_BAD_ARG;
  }
  push_axpy_dyn_kernel<<<el_grid(n, 8), EL_THREADS, 0, (cudaStream_t)stream>>>(
      w_ps, g, alpha, step_word, seen, staleness_out, n);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}
