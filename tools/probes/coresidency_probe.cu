// Can a small streaming kernel run on an SM that a persistent 320-thread CTA already holds?
// Kernel A: one CTA per SM, REGS registers per thread (forced by a live array), SMEM bytes of
// dynamic shared memory, spins for a fixed time.  Kernel B: 128-thread blocks streaming a buffer.
// B alone, A alone, then A + B on two streams.  If B cannot co-reside it finishes only after A.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o coresidency_probe coresidency_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int NREG>
__global__ void __launch_bounds__(320, 1) hog(long long cycles, float* sink) {
  extern __shared__ float sm[];
  float r[NREG];
#pragma unroll
  for (int i = 0; i < NREG; ++i) r[i] = threadIdx.x * 0.001f + i;
  const long long t0 = clock64();
  while (clock64() - t0 < cycles) {
#pragma unroll
    for (int i = 0; i < NREG; ++i) r[i] = r[i] * 1.0001f + r[(i + 1) % NREG];
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NREG; ++i) s += r[i];
  if (s == 12345.678f) sink[0] = s + sm[0];
}

__global__ void __launch_bounds__(128) stream_copy(const float4* __restrict__ a, float4* __restrict__ b, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    b[i] = a[i];
}

template <int NREG>
void run(int smem, const float4* a, float4* b, long long n, float* sink) {
  cudaStream_t s1, s2;
  cudaStreamCreate(&s1); cudaStreamCreate(&s2);
  cudaFuncSetAttribute(hog<NREG>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, hog<NREG>);
  cudaEvent_t e0, e1, e2; cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&e2);
  const long long cyc = 4000000;   // ~2 ms
  float tA, tB, tBoth, tBin;
  for (int rep = 0; rep < 2; ++rep) {
    cudaDeviceSynchronize();
    cudaEventRecord(e0, s1); hog<NREG><<<148, 320, smem, s1>>>(cyc, sink); cudaEventRecord(e1, s1);
    cudaDeviceSynchronize(); cudaEventElapsedTime(&tA, e0, e1);
    cudaEventRecord(e0, s2); for (int i = 0; i < 8; ++i) stream_copy<<<148 * 16, 128, 0, s2>>>(a, b, n); cudaEventRecord(e1, s2);
    cudaDeviceSynchronize(); cudaEventElapsedTime(&tB, e0, e1);
    cudaEventRecord(e0, s1);
    hog<NREG><<<148, 320, smem, s1>>>(cyc, sink);
    cudaEventRecord(e1, s1);
    cudaStreamWaitEvent(s2, e0, 0);
    for (int i = 0; i < 8; ++i) stream_copy<<<148 * 16, 128, 0, s2>>>(a, b, n);
    cudaEventRecord(e2, s2);
    cudaDeviceSynchronize();
    cudaEventElapsedTime(&tBoth, e0, e1); cudaEventElapsedTime(&tBin, e0, e2);
  }
  printf("hog regs=%d smem=%6d B: A alone %.2f ms, B alone %.2f ms | together: A %.2f ms, B done at %.2f ms -> %s\n",
         fa.numRegs, smem, tA, tB, tBoth, tBin, tBin < tA + 0.5f * tB ? "B ran beside A" : "B waited for A");
}

int main() {
  const long long n = 64ll << 20;   // 1 GiB of float4
  float4 *a, *b; float* sink;
  cudaMalloc(&a, n * 16); cudaMalloc(&b, n * 16); cudaMalloc(&sink, 16);
  cudaMemset(a, 0, n * 16);
  const int sm[] = {227328, 228352, 228864, 229376, 229632, 230144, 230400};
  for (int i = 0; i < 7; ++i) run<116>(sm[i], a, b, n, sink);
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
