// Variable initialisers on the device: the reference's
// tf.truncated_normal([2]) (Non-Distributed_Setup.py:13-14) and the N(0, std) / uniform /
// constant initialisers of the model families, written straight into the flat fp32 parameter
// buffer instead of being drawn on the host and copied (355 M parameters for GPT-2 medium).
//
// Counter-based: element e of stream s under seed k is Philox4x32-10(counter = {e_lo, e_hi,
// attempt, s}, key = k), so the value of a parameter depends only on (seed, tensor, element) --
// every rank that runs the kernel with the same seed holds bit-identical parameters (what
// "same seed on every rank" gives the host path), whatever the grid, the padding of the stored
// layout or the order the tensors are initialised in.
#include "common.cuh"

namespace dtg {
namespace {
constexpr int IN_THREADS = 256;

__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}
// (0, 1]: never 0, so the logarithm below is finite
__device__ __forceinline__ float u01(uint32_t x) { return (static_cast<float>(x >> 8) + 1.0f) * (1.0f / 16777216.0f); }
__device__ __forceinline__ float normal_from(uint32_t a, uint32_t b) {
  const float r = sqrtf(-2.0f * __logf(u01(a)));
  return r * __cosf(6.283185307179586f * u01(b));
}
}  // namespace

enum InitKind { INIT_NORMAL = 0, INIT_TRUNC_NORMAL = 1, INIT_UNIFORM = 2 };

// out[r * ld + c] for r < rows, c < cols; logical element index = first + r * cols + c.
//   INIT_NORMAL        a + b * z
//   INIT_TRUNC_NORMAL  a + b * z with z redrawn until |z| <= 2 (TF's truncated_normal)
//   INIT_UNIFORM       uniform in [a, b)
__global__ void __launch_bounds__(IN_THREADS)
init_random_kernel(float* __restrict__ out, long long rows, long long cols, long long ld,
                   long long first, int kind, float a, float b, uint2 key, uint32_t stream_id) {
  const long long n = rows * cols;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const unsigned long long e = static_cast<unsigned long long>(first + i);
    float v;
    if (kind == INIT_UNIFORM) {
      const uint4 r = philox4x32_10(make_uint4((uint32_t)e, (uint32_t)(e >> 32), 0u, stream_id), key);
      v = a + (b - a) * (static_cast<float>(r.x >> 8) * (1.0f / 16777216.0f));
    } else {
      float z = 0.f;
      // acceptance 95.45 % per draw: 16 attempts leave 2^-71 of the mass, clamped below
      for (uint32_t attempt = 0; attempt < 16u; ++attempt) {
        const uint4 r = philox4x32_10(make_uint4((uint32_t)e, (uint32_t)(e >> 32), attempt, stream_id), key);
        z = normal_from(r.x, r.y);
        if (kind != INIT_TRUNC_NORMAL || fabsf(z) <= 2.0f) break;
        z = normal_from(r.z, r.w);
        if (fabsf(z) <= 2.0f) break;
      }
      if (kind == INIT_TRUNC_NORMAL) z = fminf(fmaxf(z, -2.0f), 2.0f);
      v = a + b * z;
    }
    out[(i / cols) * ld + (i % cols)] = v;
  }
}

}  // namespace dtg

using namespace dtg;

extern "C" int dtg_init_random(float* out, long long rows, long long cols, long long ld,
                               long long first, int kind, float a, float b,
                               unsigned long long seed, unsigned int stream_id, void* stream) {
  if (rows <= 0 || cols <= 0) return OK;
  if (out == nullptr || ld < cols || kind < 0 || kind > 2 || first < 0) {
    dtg_set_last_error("dtg_init_random", "null output, ld < cols, or unknown kind");
    return ERR_BAD_ARG;
  }
  const long long n = rows * cols;
  long long blocks = (n + IN_THREADS - 1) / IN_THREADS;
  if (blocks > 148 * 8) blocks = 148 * 8;
  init_random_kernel<<<(int)blocks, IN_THREADS, 0, (cudaStream_t)stream>>>(
      out, rows, cols, ld, first, kind, a, b,
      make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)), stream_id);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}
