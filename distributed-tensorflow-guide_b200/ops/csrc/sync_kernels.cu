// W9 + W10: synchronous aggregation as ONE kernel per rank --
//   reduce-scatter of the workers' gradients (peer loads over NVLink, or NVLS
//   multimem.ld_reduce when a multicast mapping exists)
//   -> mean over the replicas that arrived (N-of-M, late ones are dropped)
//   -> SGD / Adagrad on this rank's fp32 master shard (the "ps shard")
//   -> all-gather of the updated parameters into every rank's bf16 (and fp32)
//      parameter buffers (peer stores, or multimem.st)
//   -> token release: a done-flag carrying the new global_step.
//
// It replaces, in the reference's synchronous schemes (Synchronous-SGD/ssgd.py:51-62):
// ConditionalAccumulator.apply_grad (K9), take_grad + ApplyGradientDescent on the ps
// (K10), the FIFO token queue (K11) and the parameter re-read (K12).  The per-worker
// learning-rate trick of Synchronous-SGD-different-learning-rates/ssgd.py:59-64 is a
// scale the producer folds into its gradient buffer (K13), so nothing changes here.
//
// Every rank runs this kernel at the same point of its stream; ranks meet through
// flag words in each other's signal pads (st.release.sys / ld.acquire.sys).  All
// waits are bounded: a dead peer yields status != 0, not a hang.
#include "common.cuh"

namespace dtg {

constexpr int MAX_WORLD = 8;
// 128 threads and at most 88 registers: a block of this kernel then fits on an SM NEXT to a
// persistent GEMM CTA (which leaves 11 776 registers and one block slot), so a bucket's
// aggregation really runs while the rest of backward computes; a 512-thread block could only
// start in the gaps between GEMM kernels.
constexpr int SYNC_THREADS = 128;

struct SyncParams {
  const float* grad[MAX_WORLD];        // rank p's gradient buffer (bucket base)
  float* param_f32[MAX_WORLD];         // rank p's fp32 parameter buffer or null
  __nv_bfloat16* param_bf16[MAX_WORLD];  // rank p's bf16 parameter buffer or null
  uint32_t* flags[MAX_WORLD];          // rank p's signal pad
  const float* mc_grad;                // multicast view of grad buffers (NVLS) or null
  float* mc_param_f32;                 // multicast view of fp32 params or null
  __nv_bfloat16* mc_param_bf16;        // multicast view of bf16 params or null
  float* master;                       // local fp32 master shard, element 0 == shard_begin
  float* slot;                         // local Adagrad accumulator for the shard or null
  unsigned long long* decision;        // local: (step << 32) | contributor mask
  unsigned int* cta_counter;           // local
  int* status;                         // local: 0 ok, 1 ready-timeout, 2 done-timeout
  long long shard_begin, shard_end;    // this rank's slice of the bucket, in elements
  long long timeout_cycles;
  int world, rank;
  int ready_base, done_base;           // flag slots: ready_base + p, done_base + p
  int replicas_to_aggregate;
  uint32_t step;                       // value the flags must reach
  float lr, gscale;                    // update uses gscale * mean(grads)
  int adagrad;
  int contributes;                     // 0: pure shard owner (dedicated ps rank), pushes no gradient
  unsigned int owner_mask;             // ranks that own a non-empty shard (their done flag is awaited)
};

__device__ __forceinline__ float4 mm_ld_reduce_v4(const float* mc) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];\n"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}
__device__ __forceinline__ void mm_st_v4(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};\n" ::"l"(mc),
               "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void mm_st_v2_b32(void* mc, uint32_t a, uint32_t b) {
  asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};\n" ::"l"(mc),
               "f"(__uint_as_float(a)), "f"(__uint_as_float(b))
               : "memory");
}

__global__ void __maxnreg__(88)
sync_rs_update_ag_kernel(const SyncParams p) {
  __shared__ unsigned int s_mask;
  __shared__ int s_last;
  const long long t0 = clock64();

  // ---- phase 0: announce "my gradients for `step` are in my buffer" ---------
  if (blockIdx.x == 0 && threadIdx.x < p.world && p.contributes) {
    __threadfence_system();
    st_release_sys_u32(p.flags[threadIdx.x] + p.ready_base + p.rank, p.step);
  }
  // ---- phase 1: block 0 decides who contributes (first N to arrive) --------
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    unsigned int mask = 0;
    const uint32_t* mine = p.flags[p.rank];
    while (true) {
      mask = 0;
      int cnt = 0;
      for (int r = 0; r < p.world; ++r)
        if ((int32_t)(ld_acquire_sys_u32(mine + p.ready_base + r) - p.step) >= 0) {
          mask |= 1u << r;
          ++cnt;
        }
      if (cnt >= p.replicas_to_aggregate) break;
      if (clock64() - t0 > p.timeout_cycles) { atomicExch(p.status, 1); break; }
    }
    const unsigned long long word = ((unsigned long long)p.step << 32) | mask;
    asm volatile("st.release.gpu.global.u64 [%0], %1;\n" ::"l"(p.decision), "l"(word) : "memory");
  }
  if (threadIdx.x == 0) {
    unsigned long long word;
    while (true) {
      asm volatile("ld.acquire.gpu.global.u64 %0, [%1];\n" : "=l"(word) : "l"(p.decision)
                   : "memory");
      if ((uint32_t)(word >> 32) == p.step) break;
      if (clock64() - t0 > 2 * p.timeout_cycles) { word = 0; break; }
    }
    s_mask = (unsigned int)(word & 0xffffffffu);
  }
  __syncthreads();
  const unsigned int mask = s_mask;
  const int cnt = __popc(mask);

  // ---- phase 2: reduce my shard, update the master copy, broadcast ---------
  if (cnt > 0) {
    const float inv = p.gscale / (float)cnt;
    const long long n = p.shard_end - p.shard_begin;
    const long long n4 = n >> 2;
    const bool use_mc_ld = p.mc_grad != nullptr && cnt == p.world;
    // Two independent 16-byte columns per thread and iteration: every load of both (one
    // multimem.ld_reduce each, or one peer load per contributing rank each) is issued before the
    // first result is consumed, which is what keeps enough bytes in flight over NVLink / HBM.
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < n4; i0 += 2 * stride) {
      float4 sum[2];
      const bool have1 = i0 + stride < n4;
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        sum[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (u == 1 && !have1) break;
        const long long e = p.shard_begin + 4 * (i0 + u * stride);   // element index in the bucket
        if (use_mc_ld) {
          sum[u] = mm_ld_reduce_v4(p.mc_grad + e);
        } else {
          float4 v[MAX_WORLD];
#pragma unroll
          for (int r = 0; r < MAX_WORLD; ++r)
            if (r < p.world && (mask >> r & 1u))
              v[r] = ld_v4_relaxed_sys(reinterpret_cast<const float4*>(p.grad[r] + e));
#pragma unroll
          for (int r = 0; r < MAX_WORLD; ++r)
            if (r < p.world && (mask >> r & 1u)) {
              sum[u].x += v[r].x; sum[u].y += v[r].y; sum[u].z += v[r].z; sum[u].w += v[r].w;
            }
        }
      }
      float4 w2[2], a2[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        if (u == 1 && !have1) break;
        w2[u] = reinterpret_cast<float4*>(p.master)[i0 + u * stride];
        if (p.adagrad) a2[u] = reinterpret_cast<float4*>(p.slot)[i0 + u * stride];
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        if (u == 1 && !have1) break;
        const long long i = i0 + u * stride;
        const long long e = p.shard_begin + 4 * i;
        float4 s = sum[u];
        s.x *= inv; s.y *= inv; s.z *= inv; s.w *= inv;
        float4 w = w2[u];
        if (p.adagrad) {
          float4 a = a2[u];
          a.x += s.x * s.x; a.y += s.y * s.y; a.z += s.z * s.z; a.w += s.w * s.w;
          reinterpret_cast<float4*>(p.slot)[i] = a;
          w.x -= p.lr * s.x * rsqrtf(a.x); w.y -= p.lr * s.y * rsqrtf(a.y);
          w.z -= p.lr * s.z * rsqrtf(a.z); w.w -= p.lr * s.w * rsqrtf(a.w);
        } else {
          w.x -= p.lr * s.x; w.y -= p.lr * s.y; w.z -= p.lr * s.z; w.w -= p.lr * s.w;
        }
        reinterpret_cast<float4*>(p.master)[i] = w;
        const uint32_t b0 = pack_bf16x2(w.x, w.y), b1 = pack_bf16x2(w.z, w.w);
        if (p.mc_param_bf16 != nullptr) {
          mm_st_v2_b32(p.mc_param_bf16 + e, b0, b1);
          if (p.mc_param_f32 != nullptr) mm_st_v4(p.mc_param_f32 + e, w);
        } else {
#pragma unroll
          for (int r = 0; r < MAX_WORLD; ++r) {
            if (r >= p.world) break;
            if (p.param_bf16[r] != nullptr)
              asm volatile("st.relaxed.sys.global.v2.b32 [%0], {%1, %2};\n" ::"l"(p.param_bf16[r] + e),
                           "r"(b0), "r"(b1) : "memory");
            if (p.param_f32[r] != nullptr)
              st_v4_relaxed_sys(reinterpret_cast<float4*>(p.param_f32[r] + e), w);
          }
        }
      }
    }
    // tail (< 4 elements), scalar, by one thread
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      for (long long t = n4 << 2; t < n; ++t) {
        const long long e = p.shard_begin + t;
        float s = 0.f;
        for (int r = 0; r < p.world; ++r)
          if (mask >> r & 1u) s += p.grad[r][e];
        s *= inv;
        float w = p.master[t];
        if (p.adagrad) { float a = p.slot[t] + s * s; p.slot[t] = a; w -= p.lr * s * rsqrtf(a); }
        else w -= p.lr * s;
        p.master[t] = w;
        for (int r = 0; r < p.world; ++r) {
          if (p.param_bf16[r]) p.param_bf16[r][e] = __float2bfloat16(w);
          if (p.param_f32[r]) p.param_f32[r][e] = w;
        }
      }
    }
  }

  // ---- phase 3: last CTA releases the token and waits for every owner ------
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(p.cta_counter, 1u);
    s_last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last) {
    if (threadIdx.x < p.world) {
      __threadfence_system();
      st_release_sys_u32(p.flags[threadIdx.x] + p.done_base + p.rank, p.step);
    }
    if (threadIdx.x == 0) {
      *p.cta_counter = 0;
      const uint32_t* mine = p.flags[p.rank];
      for (int r = 0; r < p.world; ++r) {
        if (!(p.owner_mask >> r & 1u)) continue;
        // a dead shard owner ends in a time-out status, never a hang
        while ((int32_t)(ld_acquire_sys_u32(mine + p.done_base + r) - p.step) < 0) {
          if (clock64() - t0 > 3 * p.timeout_cycles) { atomicExch(p.status, 2); break; }
        }
      }
    }
  }
}

// Plain cross-rank barrier on the signal pads (init / shutdown / tests).
__global__ void fabric_barrier_kernel(uint32_t* const* flags, int world, int rank, int slot_base,
                                      uint32_t value, long long timeout_cycles, int* status) {
  const long long t0 = clock64();
  if (threadIdx.x < world) {
    __threadfence_system();
    st_release_sys_u32(flags[threadIdx.x] + slot_base + rank, value);
    const uint32_t* mine = flags[rank];
    while ((int32_t)(ld_acquire_sys_u32(mine + slot_base + threadIdx.x) - value) < 0) {
      if (clock64() - t0 > timeout_cycles) { atomicExch(status, 3); break; }
    }
  }
}

}  // namespace dtg

using namespace dtg;

// Flat argument block shared with Python (ctypes.Structure mirrors this layout).
extern "C" int dtg_sync_rs_update_ag(const SyncParams* params, int num_blocks, void* stream) {
  if (params == nullptr || params->world < 1 || params->world > MAX_WORLD) {
    dtg_set_last_error("dtg_sync_rs_update_ag", "world must be 1..8");
    return ERR_BAD_ARG;
  }
  if ((params->shard_begin & 3) != 0) {
    dtg_set_last_error("dtg_sync_rs_update_ag", "shard_begin must be a multiple of 4 elements");
    return ERR_BAD_ARG;
  }
  if (params->adagrad && params->slot == nullptr) {
    dtg_set_last_error("dtg_sync_rs_update_ag", "adagrad needs a slot buffer");
    return ERR_BAD_ARG;
  }
  if (params->step == 0) {
    // a freshly zeroed decision word reads as "step 0, nobody contributes": blocks that look
    // before block 0 has published would skip their part of the shard without any error
    dtg_set_last_error("dtg_sync_rs_update_ag", "steps are numbered from 1 (step 0 is the cleared state)");
    return ERR_BAD_ARG;
  }
  if (num_blocks <= 0) num_blocks = 1184;
  sync_rs_update_ag_kernel<<<num_blocks, SYNC_THREADS, 0, (cudaStream_t)stream>>>(*params);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

extern "C" int dtg_sync_params_size() { return (int)sizeof(SyncParams); }

extern "C" int dtg_fabric_barrier(void* const* flags_dev_table, int world, int rank, int slot_base,
                                  unsigned int value, long long timeout_cycles, int* status,
                                  void* stream) {
  fabric_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<uint32_t* const*>(flags_dev_table), world, rank, slot_base, value,
      timeout_cycles, status);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}

// ---------------------------------------------------------------------------
// dtg_sync_rs_update_ag_noalloc: the same aggregation kernel for callers that must not
// allocate.
//
// dtg_sync_rs_update_ag needs two small pieces of scratch per bucket besides its inputs
// and outputs: the 64-bit word through which block 0 publishes which replicas
// contribute, and the counter that finds the last block.  The Python engine gets them
// from torch's allocator.  A trainer that embeds this library through the C ABI may not
// be able to do that: one that hands all of HBM to its own arena at start-up and
// treats any later cudaMalloc as an error; an asynchronous (Hogwild-style) worker loop
// in which cudaMalloc's implicit device-wide synchronisation would stall the other
// streams; or a host with a fixed, audited memory budget where a library must not grow
// it.  For them the scratch lives in the array below, which is part of the library
// image: every GPU that loads the library holds these bytes until the process exits.
//
// Limit: DTG_NOALLOC_MAX_BUCKETS buckets (16 bytes each).  A bucket index beyond that
// is rejected with a message; nothing is launched.
//
// Sharing: every call on a device uses the same array and finds what the previous call
// left there.  The entry point therefore clears its bucket's slot on the caller's
// stream before the launch (the block counter must start at zero, and a stale decision
// word carrying the same step number would otherwise be taken for the new one).  Two
// calls for the same bucket that overlap on the GPU -- different streams, no ordering
// between them -- race on that slot; give overlapping calls different buckets or order
// them.
//
// Current device: the array exists once per GPU, and its address is looked up for the device
// that is current when the call is made.  Call with the device that owns `stream` current
// (cudaSetDevice first in a multi-GPU process); a mismatch is refused where the driver can
// tell which device a stream belongs to.
//
// Step numbers start at 1: a cleared slot is indistinguishable from "step 0 decided, nobody
// contributes", so step 0 is refused by both entry points.
//
// Set-up: the first call on each device looks up the device address of the array
// (cudaGetSymbolAddress, with that device current) and remembers it.  Make that first
// call once, up front, before starting any stream capture.  No call here reaches
// cudaMalloc, cudaFree or any allocating library routine.
// ---------------------------------------------------------------------------
#define DTG_NOALLOC_MAX_BUCKETS 32
#define DTG_NOALLOC_MAX_DEVICES 64

struct __align__(16) NoAllocSlot {
  unsigned long long decision;
  unsigned int cta_counter;
  unsigned int pad;
};

// In host code this name is only a handle for cudaGetSymbolAddress, never an address.
static __device__ __align__(256) NoAllocSlot g_noalloc_scratch[DTG_NOALLOC_MAX_BUCKETS];

static NoAllocSlot* g_noalloc_dev_ptr[DTG_NOALLOC_MAX_DEVICES] = {nullptr};

extern "C" int dtg_sync_noalloc_max_buckets() { return DTG_NOALLOC_MAX_BUCKETS; }

extern "C" int dtg_sync_rs_update_ag_noalloc(const SyncParams* params, int bucket, int num_blocks,
                                             void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if (params == nullptr || params->world < 1 || params->world > MAX_WORLD) {
    dtg_set_last_error("dtg_sync_rs_update_ag_noalloc", "world must be 1..8");
    return ERR_BAD_ARG;
  }
  if (bucket < 0 || bucket >= DTG_NOALLOC_MAX_BUCKETS) {
    char msg[160];
    snprintf(msg, sizeof(msg),
             "bucket %d is outside the built-in scratch (0..%d); use dtg_sync_rs_update_ag "
             "with caller-provided scratch for more buckets",
             bucket, DTG_NOALLOC_MAX_BUCKETS - 1);
    dtg_set_last_error("dtg_sync_rs_update_ag_noalloc", msg);
    return ERR_TOO_LARGE;
  }
  if ((params->shard_begin & 3) != 0) {
    dtg_set_last_error("dtg_sync_rs_update_ag_noalloc",
                       "shard_begin must be a multiple of 4 elements");
    return ERR_BAD_ARG;
  }
  if (params->adagrad && params->slot == nullptr) {
    dtg_set_last_error("dtg_sync_rs_update_ag_noalloc", "adagrad needs a slot buffer");
    return ERR_BAD_ARG;
  }
  if (params->status == nullptr) {
    dtg_set_last_error("dtg_sync_rs_update_ag_noalloc", "status (an output) must be provided");
    return ERR_BAD_ARG;
  }
  if (params->step == 0) {
    // the slot is cleared to zero below, and zero is what "step 0, no contributors" looks like
    dtg_set_last_error("dtg_sync_rs_update_ag_noalloc",
                       "steps are numbered from 1 (step 0 is the cleared state of the scratch)");
    return ERR_BAD_ARG;
  }
  int dev = -1;
  DTG_CUDA_OK(cudaGetDevice(&dev));
  // The scratch address is looked up for the CURRENT device; a stream that belongs to another
  // device would run the kernel against the wrong GPU's copy of the array.  Refuse that.
  if (stream != nullptr && stream != cudaStreamLegacy && stream != cudaStreamPerThread) {
    int sdev = -1;
    if (cudaStreamGetDevice(stream, &sdev) == cudaSuccess) {
      if (sdev != dev) {
        dtg_set_last_error("dtg_sync_rs_update_ag_noalloc",
                           "the current device must be the device that owns `stream`");
        return ERR_BAD_ARG;
      }
    } else {
      (void)cudaGetLastError();   // older drivers: fall back to the documented precondition
    }
  }
  if (dev < 0 || dev >= DTG_NOALLOC_MAX_DEVICES) {
    dtg_set_last_error("dtg_sync_rs_update_ag_noalloc", "device index out of range");
    return ERR_UNSUPPORTED;
  }
  if (g_noalloc_dev_ptr[dev] == nullptr) {
    // once per device, with that device current
    NoAllocSlot* p = nullptr;
    DTG_CUDA_OK(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch));
    if (p == nullptr) {
      dtg_set_last_error("dtg_sync_rs_update_ag_noalloc", "cudaGetSymbolAddress returned null");
      return ERR_CUDA;
    }
    g_noalloc_dev_ptr[dev] = p;
  }
  NoAllocSlot* slot = g_noalloc_dev_ptr[dev] + bucket;
  // the slot keeps the previous call's contents: start from zero
  DTG_CUDA_OK(cudaMemsetAsync(slot, 0, sizeof(NoAllocSlot), stream));
  SyncParams p = *params;
  p.decision = &slot->decision;
  p.cta_counter = &slot->cta_counter;
  if (num_blocks <= 0) num_blocks = 1184;
  sync_rs_update_ag_kernel<<<num_blocks, SYNC_THREADS, 0, stream>>>(p);
  DTG_CUDA_OK(cudaGetLastError());
  return OK;
}
