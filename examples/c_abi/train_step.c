/* A two-layer MLP (784 -> 512 -> 10) trained on one fixed synthetic batch, written in C99 against
 * the C ABI of libdtg_b200.so only -- no Python, no torch.  The calls are the ones
 * models/mlp_native.py makes: tcgen05 GEMMs with fused bias / ReLU / ReLU' epilogues, the
 * softmax-cross-entropy kernel, MN-major weight-gradient GEMMs, the fused local SGD step that also
 * refreshes the bf16 mirror, and the device-side initialisers.
 *
 *   make            (needs nvcc's headers and libcudart; produces build/train_step)
 *   ./build/train_step [steps]
 *
 * Memory: ONE flat fp32 parameter buffer (+ bf16 mirror + fp32 gradient buffer) with every tensor
 * on a 64-element boundary, like parallel/flat.py; the last layer's 10 outputs are padded to 16 rows
 * so every row pitch is a multiple of 16 bytes (TMA).
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <cuda_runtime.h>

#include "dtg_b200.h"

#define CK(call)                                                                       \
  do {                                                                                 \
    cudaError_t e_ = (call);                                                           \
    if (e_ != cudaSuccess) {                                                           \
      fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_));       \
      return 2;                                                                        \
    }                                                                                  \
  } while (0)
#define DK(call)                                                                       \
  do {                                                                                 \
    int r_ = (call);                                                                   \
    if (r_ != 0) {                                                                     \
      fprintf(stderr, "%s:%d status %d: %s\n", __FILE__, __LINE__, r_, dtg_last_error()); \
      return 3;                                                                        \
    }                                                                                  \
  } while (0)

enum { B = 1024, IN = 784, HID = 512, OUT = 10, OUTP = 16 };
enum { OFF_W1 = 0, OFF_B1 = OFF_W1 + HID * IN, OFF_W2 = OFF_B1 + HID, OFF_B2 = OFF_W2 + OUTP * HID,
       TOTAL = ((OFF_B2 + OUTP + 63) / 64) * 64 };

int main(int argc, char** argv) {
  const int steps = argc > 1 ? atoi(argv[1]) : 40;
  const float lr = 0.05f;
  cudaStream_t st;
  float *p32, *g32, *x32, *logits, *loss;
  void *p16, *x16, *h16, *dz0, *dz1; /* bf16 */
  long long* labels;
  long long host_labels[B];
  float first = 0.f, last = 0.f;
  unsigned lcg = 12345u;
  int i, s;

  CK(cudaSetDevice(0));
  CK(cudaStreamCreate(&st));
  CK(cudaMalloc((void**)&p32, TOTAL * 4));
  CK(cudaMalloc((void**)&g32, TOTAL * 4));
  CK(cudaMalloc(&p16, TOTAL * 2));
  CK(cudaMalloc((void**)&x32, B * IN * 4));
  CK(cudaMalloc(&x16, B * IN * 2));
  CK(cudaMalloc(&h16, B * HID * 2));
  CK(cudaMalloc(&dz0, B * HID * 2));
  CK(cudaMalloc(&dz1, B * HID * 2));
  CK(cudaMalloc((void**)&logits, B * OUTP * 4));
  CK(cudaMalloc((void**)&loss, 4));
  CK(cudaMalloc((void**)&labels, B * 8));
  CK(cudaMemsetAsync(p32, 0, TOTAL * 4, st));
  CK(cudaMemsetAsync(g32, 0, TOTAL * 4, st));

  /* Glorot-uniform weights and a normal input batch, drawn on the device (kind 2 / kind 0);
   * biases and the 6 padding rows of the last layer stay zero */
  DK(dtg_init_random(p32 + OFF_W1, 1, (long long)HID * IN, (long long)HID * IN, 0, 2,
                     -sqrtf(6.f / (IN + HID)), sqrtf(6.f / (IN + HID)), 7ull, 1u, st));
  DK(dtg_init_random(p32 + OFF_W2, OUT, HID, HID, 0, 2, -sqrtf(6.f / (HID + OUT)),
                     sqrtf(6.f / (HID + OUT)), 7ull, 2u, st));
  DK(dtg_init_random(x32, 1, (long long)B * IN, (long long)B * IN, 0, 0, 0.f, 1.f, 7ull, 3u, st));
  DK(dtg_cast_f32_bf16(p32, p16, TOTAL, st));
  DK(dtg_cast_f32_bf16(x32, x16, (long long)B * IN, st));
  for (i = 0; i < B; ++i) {
    lcg = lcg * 1664525u + 1013904223u;
    host_labels[i] = (long long)((lcg >> 16) % OUT);
  }
  CK(cudaMemcpyAsync(labels, host_labels, sizeof(host_labels), cudaMemcpyHostToDevice, st));

  for (s = 0; s < steps; ++s) {
    const char* w1 = (const char*)p16 + 2 * (size_t)OFF_W1;
    const char* w2 = (const char*)p16 + 2 * (size_t)OFF_W2;
    /* forward: h = relu(x W1^T + b1); logits = h W2^T + b2 (fp32) */
    DK(dtg_gemm_bf16(x16, w1, B, HID, IN, IN, IN, 0, 0, h16, NULL, p32 + OFF_B1, NULL, NULL, HID, 0,
                     1.f, 0, 1, 0, 0, st));
    DK(dtg_gemm_bf16(h16, w2, B, OUTP, HID, HID, HID, 0, 0, logits, NULL, p32 + OFF_B2, NULL, NULL,
                     OUTP, 0, 1.f, 1, 0, 0, 0, st));
    /* loss (mean) and dlogits = (softmax - onehot) / B in bf16, padding columns zero */
    CK(cudaMemsetAsync(loss, 0, 4, st));
    DK(dtg_softmax_xent(logits, labels, dz0, loss, B, OUT, OUTP, HID, 1.f / B, st));
    /* backward, last layer: dW2 = dz^T h (both operands MN-major), db2, dh = (dz W2) * relu'(h) */
    DK(dtg_gemm_bf16(dz0, h16, OUTP, HID, B, HID, HID, 1, 1, g32 + OFF_W2, NULL, NULL, NULL, NULL,
                     HID, 0, 1.f, 1, 0, 0, 0, st));
    CK(cudaMemsetAsync(g32 + OFF_B2, 0, OUTP * 4, st));
    DK(dtg_bias_grad(dz0, g32 + OFF_B2, B, OUTP, HID, 1.f, st));
    DK(dtg_gemm_bf16(dz0, w2, B, HID, OUTP, HID, HID, 0, 1, dz1, NULL, NULL, h16, NULL, HID, HID,
                     1.f, 0, 0, 1, 0, st));
    /* first layer: dW1 = dh^T x, db1 */
    DK(dtg_gemm_bf16(dz1, x16, HID, IN, B, HID, IN, 1, 1, g32 + OFF_W1, NULL, NULL, NULL, NULL, IN,
                     0, 1.f, 1, 0, 0, 0, st));
    CK(cudaMemsetAsync(g32 + OFF_B1, 0, HID * 4, st));
    DK(dtg_bias_grad(dz1, g32 + OFF_B1, B, HID, HID, 1.f, st));
    /* p -= lr * g over the whole flat buffer, bf16 mirror refreshed in the same pass */
    DK(dtg_local_step(p32, p16, NULL, NULL, g32, lr, 0, TOTAL, st));
    if (s == 0 || s == steps - 1 || s % 10 == 0) {
      float l;
      CK(cudaMemcpyAsync(&l, loss, 4, cudaMemcpyDeviceToHost, st));
      CK(cudaStreamSynchronize(st));
      printf("step %3d  loss %.4f\n", s, l);
      if (s == 0) first = l;
      last = l;
    }
  }
  CK(cudaStreamSynchronize(st));
  printf("C_ABI_TRAIN %s: loss %.4f -> %.4f in %d steps\n", last < 0.9f * first ? "ok" : "FAILED",
         first, last, steps);
  return last < 0.9f * first ? 0 : 1;
}
