/* C ABI of libdtg_b200.so -- the sm_100a kernels of the dtg_b200 framework.
 *
 * Every function returns 0 on success and a dtg status code otherwise
 * (1 bad argument, 2 CUDA error, 3 TMA descriptor encode failed, 4 too large,
 * 5 unsupported); dtg_last_error() gives the message.  `stream` is a cudaStream_t
 * passed as void*.  All pointers are device pointers unless stated; "peer" means the
 * pointer may be another GPU's memory mapped over NVLink (CUDA IPC / peer access).
 *
 * Citations name the place in the reference (/root/reference) whose behaviour the
 * entry point reproduces.
 */
#ifndef DTG_B200_H_
#define DTG_B200_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ runtime */
const char* dtg_last_error(void);
int dtg_version(void);

/* ------------------------------------------------------------------ GEMM
 * D[M,N] = epilogue(alpha * op(A) . op(B)), bf16 operands, fp32 accumulation in TMEM
 * (tcgen05.mma, TMA-fed).  a_mn = 0: A is [M,K] row-major; 1: A is [K,M] row-major.
 * b_mn = 0: B is [N,K] row-major; 1: B is [K,N] row-major.  lda/ldb/ldo/ldmask are
 * row pitches in elements; operands need 16-byte aligned bases and pitches.
 *   mode      0 bf16 store | 1 fp32 store | 2 fp32 += | 3 fp32 red.global.add (push;
 *             `out` may be peer memory) | 4 same, staged for 128 B NVLink payloads
 *   act       0 none | 1 ReLU | 2 GELU(tanh)        bias: fp32 [N] or NULL
 *   mask_mode 0 none | 1 relu' of saved activation | 2 gelu' of saved pre-activation
 *   out2      optional bf16 copy of the pre-activation; residual: optional bf16 addend
 *   force_bn  0 auto | 64/128/256 one-CTA tile width | 512 two-CTA (cta_group::2) 256x256 pair tile | 768 tall pair kernel (512x256 tiles)
 * Replaces the reference's matmul-free toy forward/backward (Non-Distributed_Setup.py:15-20)
 * for the BASELINE.json models, and fuses ApplyGradientDescent on the ps
 * (Hogwild/Hogwild.py:44) into the weight-gradient epilogue (mode 3, alpha = -lr).
 */
int dtg_gemm_bf16(const void* A, const void* B, int M, int N, int K, long long lda,
                  long long ldb, int a_mn, int b_mn, void* out, void* out2, const float* bias,
                  const void* mask, const void* residual, long long ldo, long long ldmask,
                  float alpha, int mode, int act, int mask_mode, int force_bn, void* stream);
void dtg_gemm_clear_tensor_map_cache(void);

/* ------------------------------------------------------- ps data plane
 * PULL: dst_bf16[i] = bf16(src[i]), dst_f32[i] = src[i]; src may be peer memory
 * (assign_global_to_local, DOWNPOUR/DOWNPOUR.py:149-159; variable reads of
 * Hogwild/Hogwild.py:53). */
int dtg_pull_convert(const float* src, void* dst_bf16, float* dst_f32, long long n, void* stream);
/* Zero `count` fp32 tensors with one launch; table_dev (device memory) holds {address, element
 * count} pairs as 64-bit words, `longest` = the largest count (sizes the grid). */
int dtg_zero_many(const void* table_dev, int count, long long longest, void* stream);
/* fp32 copy of up to 16 small element ranges [lo, hi) of a (peer) shard into dst_f32 at the same
 * offsets, one launch: what a Hogwild pull still needs in fp32 (biases) next to the bf16 mirror. */
int dtg_pull_ranges(const float* src, float* dst_f32, const long long* lo, const long long* hi, int n,
                    void* stream);
/* PUSH, SGD on the ps: w_ps += alpha * g, unlocked fp32 reductions (Hogwild/Hogwild.py:44;
 * ADAG/ADAG.py:84-86 with alpha = -lr/T). */
int dtg_push_axpy(float* w_ps, const float* g, float alpha, long long n, void* stream);
/* PUSH, Adagrad on the ps: acc += (gscale*g)^2 atomically, w -= lr*g*rsqrt(acc)
 * (DOWNPOUR/DOWNPOUR.py:92,96-98). */
int dtg_push_adagrad(float* w_ps, float* acc_ps, const float* g, float lr, float gscale,
                     long long n, void* stream);
/* Worker-local SGD/Adagrad step + window accumulate + bf16 refresh in one pass
 * (DOWNPOUR/DOWNPOUR.py:63-78, ADAG/ADAG.py:69-83). */
int dtg_local_step(float* p, void* p_bf16, float* slot, float* window, const float* g, float lr,
                   int adagrad, long long n, void* stream);
/* global_step += delta on a (peer) 64-bit word (DOWNPOUR/DOWNPOUR.py:88,96-98). */
int dtg_counter_add(void* counter, unsigned long long delta, void* stream);

/* ------------------------------------------------------- model kernels */
int dtg_softmax_xent(const float* logits, const long long* labels, void* dlogits_bf16,
                     float* loss_sum, int B, int C, int ldl, int ldd, float scale, void* stream);
/* Same loss for bf16 logits, in place: the row is read once (into shared memory) and overwritten with
 * scale * (softmax - onehot); columns [C, ld) are padding (ignored, zeroed).  ld % 8 == 0,
 * ld <= 65536.  The LM-head loss of the GPT-2 / BERT configs. */
int dtg_softmax_xent_bf16(void* logits_inout, const long long* labels, float* loss_sum, int B, int C,
                          long long ld, float scale, void* stream);
int dtg_bias_grad(const void* dY_bf16, float* target, int M, int N, long long ldy, float alpha,
                  void* stream);
int dtg_cast_f32_bf16(const float* src, void* dst_bf16, long long n, void* stream);
int dtg_add_bf16(const void* a, const void* b, void* out, long long n, void* stream);
/* Batched GEMM (nb1 * nb2 problems, operand / output batch strides in elements): attention's
 * QK^T, PV and their backward products on per-head views of the packed QKV buffer. */
int dtg_gemm_bf16_batched(const void* A, const void* B, int M, int N, int K, long long lda,
                          long long ldb, int a_mn, int b_mn, void* out, long long ldo, float alpha,
                          int mode, int nb1, int nb2, long long a_bs1, long long a_bs2,
                          long long b_bs1, long long b_bs2, long long o_bs1, long long o_bs2,
                          int force_bn, void* stream);
/* Transformer glue (GPT-2 medium / BERT-large of BASELINE.json configs 4-5; no counterpart
 * in the reference, whose only model is the two-vector toy).  Parameter gradients are
 * reduction adds of alpha * grad (target may be a gradient buffer, a window or a ps shard). */
int dtg_layernorm_fwd(const void* x_bf16, const float* gamma, const float* beta, void* y_bf16,
                      float* mean, float* rstd, int rows, int cols, float eps, void* stream);
int dtg_layernorm_bwd(const void* dy_bf16, const void* x_bf16, const float* gamma,
                      const float* mean, const float* rstd, const void* dx_add_bf16,
                      void* dx_bf16, float* dgamma, float* dbeta, int rows, int cols, float alpha,
                      void* stream);
int dtg_embedding_fwd(const long long* ids, const void* wte_bf16, const void* wpe_bf16,
                      void* out_bf16, int rows, int seq_len, int dim, void* stream);
int dtg_embedding_bwd(const long long* ids, const void* dout_bf16, float* dwte, float* dwpe,
                      int rows, int seq_len, int dim, float alpha, void* stream);
int dtg_attn_softmax_fwd(const float* S, void* P_bf16, long long nrows, int T, float scale,
                         int causal, void* stream);
int dtg_attn_softmax_bwd(const void* P_bf16, const float* dP, void* dS_bf16, long long nrows,
                         int T, float scale, void* stream);
/* Batched GEMM with the attention epilogues (mode 5: per-tile softmax statistics, 6: P =
 * exp2(S*scale2 - rowv1) * rowv2 in bf16, 7: dS = alpha * P * (acc - rowv1)) and causal tile
 * skipping.  Used when head_dim != 64; head_dim 64 runs the fused kernels below. */
int dtg_gemm_bf16_batched_ex(const void* A, const void* B, int M, int N, int K, long long lda,
                             long long ldb, int a_mn, int b_mn, void* out, long long ldo,
                             float alpha, int mode, int nb1, int nb2, long long a_bs1,
                             long long a_bs2, long long b_bs1, long long b_bs2, long long o_bs1,
                             long long o_bs2, const float* rowv1, const float* rowv2, float* stats,
                             const void* pmat, float scale2, int causal, void* stream);
int dtg_attn_stats_merge(const float* stats, float* rowm, float* rowinvl, long long nrows,
                         int tiles_n, int T, int causal, void* stream);
/* D[(b*H+h)*T+t] = sum_c dO[b*T+t, h*hd+c] * O[b*T+t, h*hd+c] */
int dtg_attn_rowdot(const void* dO_bf16, const void* O_bf16, float* D, int B, int T, int H, int hd,
                    void* stream);
/* Fused tcgen05 attention, head_dim 64 (csrc/attn_sm100.cu).  q/k/v/dk/dv: [B, T, H, 64] views
 * of packed buffers with element strides (ld = row pitch, hs = head stride, bs = batch stride).
 * Forward writes ctx[b*T+t, h*64..] and the row statistics (max in log2 domain, 1/sum-exp);
 * P (optional, [B,H,T,T] bf16, zero-initialised when causal) is stored by TMA when non-null.
 * Backward recomputes P from the statistics, accumulates dK/dV in TMEM and stores
 * dS = scale * P * (dO V^T - D) (bf16 [B,H,T,T], zero-initialised when causal) for the dQ GEMM. */
int dtg_attn_fwd(const void* q, const void* k, const void* v, long long ld, long long hs,
                 long long bs, int B, int H, int T, int hd, void* P, void* ctx, long long ctx_ld,
                 float* rowm, float* rowinvl, float scale, int causal, void* stream);
int dtg_attn_bwd(const void* q, const void* k, const void* v, void* dk, void* dv, long long ld,
                 long long hs, long long bs, const void* dO, long long do_ld, long long do_hs,
                 long long do_bs, int B, int H, int T, int hd, void* dS, const float* rowm,
                 const float* rowinvl, const float* D, float scale, int causal, void* stream);
/* Data plane of the reference's TODO schemes (csrc/elastic_kernels.cu). */
int dtg_easgd_elastic(float* centre, float* x, void* x_bf16, const float* velocity, float delta,
                      float alpha, long long n, void* stream);
int dtg_momentum_step(float* x, float* v, const float* g, void* x_bf16, float lr, float delta,
                      long long n, void* stream);
int dtg_push_axpy_dyn(float* w_ps, const float* g, float alpha, const long long* step_word,
                      const long long* seen, long long* staleness_out, long long n, void* stream);
/* Host replay of the tall-pair GEMM's static tile schedule (no GPU needed): OK iff every 256x256
 * output block of an M x N problem on `clusters` CTA pairs is produced exactly once.
 * info[6] = big m-blocks, big tiles, small tiles, extra rounds, clusters used, 1000 x worst load. */
int dtg_tall_schedule_check(int M, int N, int clusters, float small_cost, int group, int* info);
/* Host check: the grouped tile order of the persistent GEMM kernels is a bijection onto the
 * tiles_m x tiles_n grid for this group size. */
int dtg_tile_order_check(int tiles_m, int tiles_n, int group);
/* Variable initialisers on the device (csrc/init_kernels.cu): counter-based (Philox4x32-10 over
 * {element, attempt, stream_id} keyed by seed), so equal seeds give bit-identical parameters on
 * every rank.  Writes out[r*ld + c], r < rows, c < cols; element index = first + r*cols + c.
 * kind 0: a + b*N(0,1); 1: the same truncated at 2 sigma by redrawing (tf.truncated_normal,
 * /root/reference/Non-Distributed_Setup.py:13-14); 2: uniform in [a, b). */
int dtg_init_random(float* out, long long rows, long long cols, long long ld, long long first,
                    int kind, float a, float b, unsigned long long seed, unsigned int stream_id,
                    void* stream);
/* Striped ps memory (CUDA VMM): one physical allocation per stripe, one virtual range. */
int dtg_vmm_granularity(size_t* out);
int dtg_vmm_create(size_t bytes, unsigned long long* handle, int* fd);
int dtg_vmm_import(int fd, unsigned long long* handle);
int dtg_vmm_map_striped(const unsigned long long* handles, size_t stripe_bytes, size_t stripes,
                        void** va_out);
int dtg_vmm_unmap(void* va, size_t total);
int dtg_vmm_release(unsigned long long handle);
/* NVLS multicast objects (in-switch reduction / broadcast for the synchronous step). */
int dtg_mc_supported(void);
int dtg_mc_granularity(int world, size_t size, size_t* out);
int dtg_mc_create(int world, size_t size, unsigned long long* handle, int* fd);
int dtg_mc_add_device(unsigned long long handle);
int dtg_mc_bind(unsigned long long mc_handle, unsigned long long mem_handle, size_t size);
/* teardown: detach this device's memory from the object (after its mappings are gone) */
int dtg_mc_unbind(unsigned long long mc_handle, size_t size);
int dtg_vmm_map_one(unsigned long long handle, size_t size, size_t align, void** va_out);
/* CRC-32C (Castagnoli), host side: TF checkpoint V2 checksums. */
unsigned int dtg_crc32c(const void* data, size_t n, unsigned int crc);
/* Convolution as an implicit GEMM (k x k, NHWC bf16, C % 64 == 0): the patch matrix is never
 * written; one GEMM operand is gathered from x by TMA im2col-mode loads.
 *   which = 1 forward:  out[N*Ho*Wo, Cout] = im2col(x) . other[Cout, k*k*C]^T   (other = weights)
 *   which = 2 wgrad:    out[Cout, k*k*C] (+)= alpha * other[N*Ho*Wo, Cout]^T . im2col(x)  (other = dY)
 * A stride-1 data gradient is which = 1 over dY with the flipped, channel-swapped filter and
 * pad' = k-1-pad (ops/conv.py::conv_gemm_dgrad).  mode / alpha / bias / force_bn as dtg_gemm_bf16.
 * No counterpart in the reference (its graphs have no convolution); model family of config 3. */
int dtg_conv_gemm_bf16(const void* x, int N, int H, int W, int C, int k, int stride, int pad,
                       int which, const void* other, long long ld_other, int Cout, void* out,
                       long long ldo, const float* bias, float alpha, int mode, int force_bn,
                       void* stream);
/* Conv-net glue (ResNet-50 of config 3), NHWC bf16, channels % 8 == 0. */
int dtg_im2col(const void* x, void* col, int N, int H, int W, int C, int kh, int kw, int stride,
               int pad, void* stream);
int dtg_col2im(const void* dcol, const void* dx_add, void* dx, int N, int H, int W, int C, int kh,
               int kw, int stride, int pad, void* stream);
int dtg_bn_stats(const void* x, float* sums, float* sumsq, long long rows, int C, void* stream);
int dtg_bn_apply(const void* x, const float* sums, const float* sumsq, const float* gamma,
                 const float* beta, const void* residual, void* y, long long rows, int C,
                 float eps, int relu, void* stream);
int dtg_bn_bwd(const void* dy, const void* x, const void* y, const float* sums, const float* sumsq,
               const float* gamma, float* dsum, float* dsumx, void* dx, void* dres, float* dgamma,
               float* dbeta, long long rows, int C, float eps, int relu, float alpha, void* stream);
int dtg_maxpool_fwd(const void* x, void* y, int N, int H, int W, int C, int k, int stride, int pad,
                    void* stream);
int dtg_maxpool_bwd(const void* x, const void* y, const void* dy, void* dx, int N, int H, int W,
                    int C, int k, int stride, int pad, void* stream);
int dtg_avgpool_fwd(const void* x, void* y, int N, int HW, int C, void* stream);
int dtg_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream);

/* ------------------------------------------------------- synchronous path
 * One kernel per rank: reduce-scatter of gradients over NVLink peers (first N of M
 * to arrive; late ones dropped) -> mean -> SGD/Adagrad on the local fp32 master
 * shard -> all-gather of bf16/fp32 parameters -> token (new global_step) release.
 * `params` points to a host-side dtg::SyncParams (see ops/csrc/sync_kernels.cu; the
 * Python mirror is parallel/peer_store.py::SyncParams).
 * params->step must be >= 1 (steps are numbered from 1; a cleared decision word means
 * "step 0", so step 0 is refused with status 1 = bad argument).
 * (SyncReplicasOptimizer, Synchronous-SGD/ssgd.py:51-62.) */
int dtg_sync_rs_update_ag(const void* params, int num_blocks, void* stream);
int dtg_sync_params_size(void);
/* Same kernel for callers that must not allocate: the per-bucket scratch (decision word +
 * last-block counter) comes from an array built into the library instead of from the
 * caller (params->decision / params->cta_counter are ignored; params->status is still the
 * caller's output).  For embedders whose HBM is fully owned by their own arena, or whose
 * worker loop cannot take the device-wide synchronisation of cudaMalloc.  `bucket` selects
 * one of dtg_sync_noalloc_max_buckets() (= 32) 16-byte slots; larger indices are refused
 * with status 4.  The slot is cleared on `stream` before each launch; calls for the same
 * bucket that overlap on the GPU race.  The first call on each device resolves the array's
 * address (cudaGetSymbolAddress) -- make it before any stream capture begins.  The array
 * exists once per GPU and is looked up for the CURRENT device: call with the device that
 * owns `stream` current (a mismatch is refused when the driver can report the stream's
 * device).  params->step must be >= 1, as above.  Never calls cudaMalloc/cudaFree.  This entry point has no counterpart in the reference. */
int dtg_sync_rs_update_ag_noalloc(const void* params, int bucket, int num_blocks, void* stream);
int dtg_sync_noalloc_max_buckets(void);
int dtg_fabric_barrier(void* const* flags_dev_table, int world, int rank, int slot_base,
                       unsigned int value, long long timeout_cycles, int* status, void* stream);

/* ------------------------------------------------------- peer-memory fabric */
int dtg_ipc_handle_size(void);
int dtg_ipc_alloc(size_t bytes, void** ptr, void* handle_out);
int dtg_ipc_open(const void* handle, void** ptr);
int dtg_ipc_close(void* ptr);
int dtg_free(void* ptr);
int dtg_enable_peer_access(int dev, int peer);
int dtg_memset_async(void* ptr, int value, size_t bytes, void* stream);
int dtg_memcpy_async(void* dst, const void* src, size_t bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DTG_B200_H_ */
