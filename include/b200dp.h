/*
 * b200dp -- C ABI of the native B200 (sm_100a) library.
 *
 * Every entry point is `extern "C"`, takes raw device pointers plus a CUDA
 * stream, launches asynchronously on that stream and returns 0 on success or a
 * non-zero code (a cudaError_t value, or a negative B200DP_E* code) with a
 * human-readable message available from b200dp_last_error().  Python reaches
 * these through ctypes (cs744_pytorch_distributed_tutorial_b200/ops/_lib.py);
 * C/C++ callers can link libb200dp.so directly.
 *
 * Layout conventions (internal to the framework; converted at the
 * state_dict boundary):
 *   activations   bf16, NHWC  [B][H][W][C]
 *   conv weights  fp32 master in KRSC [Cout][3][3][Cin]; bf16 shadows
 *                 Wf [Cout][9*Cin] (forward) and Wd [Cin][9*Cout] (dgrad, taps flipped)
 *   gradients     fp32, same layout as the master parameters, in one flat buffer
 *
 * Citations of the form (ref: path:line) point into the reference repository
 * kkyyhh96/CS744_PyTorch_Distributed_Tutorial and name the behaviour the entry
 * point reproduces; none of its code is reused.
 */
#ifndef B200DP_H_
#define B200DP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* b200dp_stream_t; /* cudaStream_t */

#define B200DP_EINVAL (-22)
#define B200DP_ENOSYS (-38)
#define B200DP_ETIMEDOUT (-110)

/* Last error message recorded on the calling thread ("" if none). */
const char* b200dp_last_error(void);
/* Library version / build string (contains "sm_100a"). */
const char* b200dp_version(void);

/* ------------------------------------------------------------------------
 * tcgen05 / TMEM / TMA implicit-GEMM convolution
 * (ref: master/part1/model.py:18-23 nn.Conv2d(k=3,s=1,p=1,bias=True) and its autograd)
 * ------------------------------------------------------------------------ */

/* out[B,H,W,Cout] (bf16) = conv3x3_pad1(act[B,H,W,Cin] bf16, wgt[Cout][taps*Cin] bf16) + bias.
 * taps = 9 (3x3, pad 1) or 1 (pointwise GEMM on a pre-gathered matrix, used for layer 0).
 * Cin % 64 == 0, Cout % block_n == 0, block_n in {32,64,128,256} or 0 = choose.
 * The same entry point computes dgrad when given dY as `act` and the
 * tap-flipped transposed weights Wd as `wgt` (bias = NULL).
 * If stats != NULL (2*Cout floats, zero on entry) the epilogue also adds the
 * per-channel sum and sum of squares of the bf16-rounded outputs into
 * stats[0..Cout) and stats[Cout..2Cout) (fused BatchNorm statistics). */
int b200dp_conv_igemm(const void* act, const void* wgt, const float* bias, void* out, float* stats, int B,
                      int H, int W, int Cin, int Cout, int taps, int block_n, b200dp_stream_t stream);

/* The same with the K loop (taps x Cin/64 blocks) split over split_k CTAs per output tile -- for the
 * small-spatial layers whose pixel count yields too few tiles to fill 148 SMs.  Partial tiles are
 * combined with red.global.add in `workspace` (B*H*W*Cout floats, ZERO on entry; left zero on exit),
 * then a finalize kernel adds the bias, rounds to bf16 and takes the BatchNorm statistics. */
int b200dp_conv_igemm_splitk(const void* act, const void* wgt, const float* bias, void* out, float* stats,
                             float* workspace, int B, int H, int W, int Cin, int Cout, int taps, int block_n,
                             int split_k, b200dp_stream_t stream);

/* CTA-pair variant (tcgen05.mma.cta_group::2): one 256-pixel x block_n tile per pair of SMs; each SM
 * stages its own 128 pixel rows and only half of the weight tile, the pair's single MMA thread issues
 * for both and commits by cluster multicast.  Needs an even number of 128-pixel tiles and block_n in
 * {64,128,256}; split_k / workspace as above (split_k = 1: workspace may be NULL). */
int b200dp_conv_igemm_pair(const void* act, const void* wgt, const float* bias, void* out, float* stats,
                           float* workspace, int B, int H, int W, int Cin, int Cout, int taps, int block_n,
                           int split_k, b200dp_stream_t stream);

/* Data gradient that also produces the BatchNorm-backward sums of the block below it.  act = dY of
 * this conv, wgt = its tap-flipped transposed weights, out = dL/da for the previous block's output
 * a = relu(bn(z)) (max-pooled or not).  bn_out is that stored output (same [B,H,W,Cout] shape as out);
 * on return  sums[0..Cout) += sum_{a>0} g  and  sums[Cout..2Cout) += sum_{a>0} g*(a-beta)/gamma  with g
 * the gradient just written -- exactly what b200dp_bn_relu_pool_bwd_reduce_from_out computes, fused into
 * the GEMM epilogue (split_k > 1: into the finalize kernel).  No bias, no forward statistics.
 * block_n in {32,64,128,256} ({64,128,256} with cta_pair); workspace as for b200dp_conv_igemm_splitk.
 * The executor's data gradient for every block but the first. */
int b200dp_conv_dgrad_bn_reduce(const void* act, const void* wgt, void* out, float* workspace, const void* bn_out,
                                const float* gamma, const float* beta, float* sums, int B, int H, int W, int Cin,
                                int Cout, int taps, int block_n, int split_k, int cta_pair, b200dp_stream_t stream);

/* dW[Cout][taps][cin_valid] (fp32) += sum over pixels dY[p][Cout] * X[p + tap][Cin].
 * dy: bf16 [B,H,W,Cout]; x: bf16 [B,H,W,Cin].  Both operands are consumed
 * MN-major straight from their NHWC layout.  split_k = 0 chooses a split that
 * fills the 148 SMs; with split_k > 1 partial tiles are combined with
 * red.global.add, so dw must be zero (or hold the value to accumulate onto).
 * cin_valid <= Cin restricts which input-channel columns are written (layer 0:
 * 27 of the 64 padded im2col columns). */
int b200dp_conv_wgrad(const void* dy, const void* x, float* dw, int B, int H, int W, int Cin, int Cout,
                      int taps, int cin_valid, int split_k, int accumulate, b200dp_stream_t stream);

/* Weight gradient fused with the transfer half of the data-parallel reduce-scatter (optional path of part3,
 * B200DP_WGRAD_PUSH=1; measured slower than the bucket reducer at 2 and 8 GPUs).  As b200dp_conv_wgrad_tiled (dw zero on entry, all Cin columns), plus: the tensor sits
 * at element flat_off of the flat gradient buffer, inside the bucket [own_start, own_start + own_len) that
 * the ranks own in `world` equal float4 slices (the slicing of the two-shot all-reduce).  When an output
 * tile is complete on this rank its float4s owned by rank r are stored over NVLink into peer_stage[r] at
 * stage_base + (rank * per4 + j) * 4, j = index inside r's slice, per4 = ceil(own_len / 4 / world).
 * tile_counters: >= taps * ceil(Cout/128) * Cin/block_n zeroed uints (left zero); peer_stage: host array
 * of `world` device pointers.  The owner finishes the reduction in b200dp_reduce_staged_sgd. */
int b200dp_conv_wgrad_push(const void* dy, const void* x, float* dw, int B, int H, int W, int Cin, int Cout,
                           int taps, int split_k, int block_n, unsigned int* tile_counters, int n_counters,
                           int world, int rank, void* const* peer_stage, long long flat_off, long long own_start,
                           long long own_len, long long stage_base, b200dp_stream_t stream);

/* As b200dp_conv_wgrad with an explicit input-channel (GEMM N) tile: 64, 128 or 256; 0 = default. */
int b200dp_conv_wgrad_tiled(const void* dy, const void* x, float* dw, int B, int H, int W, int Cin, int Cout,
                            int taps, int cin_valid, int split_k, int accumulate, int block_n,
                            b200dp_stream_t stream);

/* Precision-generic forms of the three GEMM entry points above.  dtype 0 = bf16 (identical to the entry points
 * without _ex), dtype 1 = tf32: activations (act / out / bn_out / dy / x) and the operand copy of the weights
 * are fp32 in memory, the tensor cores read them as tf32 (tcgen05.mma kind::tf32, the TMA unit rounds to
 * nearest on the way into shared memory) and accumulate in fp32 -- the precision of the reference's fp32
 * model under torch's TF32 convolutions (ref: master/part1/model.py:18-25, no autocast anywhere).
 * With dtype 1, Cin must be a multiple of 32 (bf16: 64); the wgrad N tile may also be 32.
 * cta_pair != 0 selects the cta_group::2 kernel; split_k / workspace as for b200dp_conv_igemm_splitk. */
int b200dp_conv_igemm_ex(const void* act, const void* wgt, const float* bias, void* out, float* stats,
                         float* workspace, int B, int H, int W, int Cin, int Cout, int taps, int block_n, int split_k,
                         int cta_pair, int dtype, b200dp_stream_t stream);
int b200dp_conv_dgrad_bn_reduce_ex(const void* act, const void* wgt, void* out, float* workspace, const void* bn_out,
                                   const float* gamma, const float* beta, float* sums, int B, int H, int W, int Cin,
                                   int Cout, int taps, int block_n, int split_k, int cta_pair, int dtype,
                                   b200dp_stream_t stream);
int b200dp_conv_wgrad_ex(const void* dy, const void* x, float* dw, int B, int H, int W, int Cin, int Cout, int taps,
                         int cin_valid, int split_k, int accumulate, int block_n, int dtype, b200dp_stream_t stream);

/* ------------------------------------------------------------------------
 * Bandwidth-bound layer kernels (bf16 NHWC activations, 8 channels per thread)
 * ------------------------------------------------------------------------ */

/* Layer 0 patch gather: x fp32 NCHW [B,3,H,W] -> a0 bf16 [B*H*W][64]; column
 * k = (r*3+s)*3 + c for k < 27, zero above.  Turns the 3-channel conv into a
 * K=64 GEMM for b200dp_conv_igemm(taps=1).  (ref: master/part1/model.py:13,18) */
int b200dp_im2col3x3_c3(const float* x_nchw, void* a0, int B, int H, int W, b200dp_stream_t stream);

/* Precision-generic forms of the bandwidth-bound layer kernels below: the same contracts with a trailing
 * `dtype` before the stream -- 0 = bf16 activations (identical to the entry point without _ex), 1 = fp32
 * activations (the tf32 arm; im2col then writes 32-column patch rows).  Parameters, statistics and gradients of
 * parameters are fp32 either way.  (ref: the fp32 model of master/part1/model.py:16-25,40-46) */
int b200dp_im2col3x3_c3_ex(const float* x_nchw, void* a0, int B, int H, int W, int dtype, b200dp_stream_t stream);
/* flags bit 0 (bf16 only): columns 32..63 of a0 are already zero (they are zero in every step) and are left alone,
 * which halves the bytes written -- for a caller that owns a persistent zero-initialised patch matrix. */
int b200dp_im2col3x3_c3_ex2(const float* x_nchw, void* a0, int B, int H, int W, int dtype, int flags,
                            b200dp_stream_t stream);
int b200dp_bn_stats_ex(const void* y, float* stats, long long P, int C, int dtype, b200dp_stream_t stream);
int b200dp_bn_relu_pool_fwd_ex(const void* y, const float* stats, const float* gamma, const float* beta,
                               float* running_mean, float* running_var, long long* num_batches_tracked,
                               float* save_mean, float* save_invstd, void* out, int B, int H, int W, int C, int pool,
                               int training, float momentum, float eps, int dtype, b200dp_stream_t stream);
int b200dp_bn_relu_pool_bwd_reduce_ex(const void* y, const void* dout, const float* mean, const float* invstd,
                                      const float* gamma, const float* beta, float* sums, int B, int H, int W, int C,
                                      int pool, int dtype, b200dp_stream_t stream);
int b200dp_bn_relu_pool_bwd_reduce_from_out_ex(const void* out, const void* dout, const float* gamma,
                                               const float* beta, float* sums, long long Q, int C, int dtype,
                                               b200dp_stream_t stream);
int b200dp_bn_relu_pool_bwd_apply_ex(const void* y, const void* dout, const float* mean, const float* invstd,
                                     const float* gamma, const float* beta, const float* sums, void* dy,
                                     float* dgamma, float* dbeta, int B, int H, int W, int C, int pool, int dtype,
                                     b200dp_stream_t stream);
int b200dp_bn_relu_pool_fwd_tma_ex(const void* y, const float* stats, const float* gamma, const float* beta,
                                   float* running_mean, float* running_var, long long* num_batches_tracked,
                                   float* save_mean, float* save_invstd, void* out, int B, int H, int W, int C,
                                   int pool, int training, float momentum, float eps, int dtype,
                                   b200dp_stream_t stream);
int b200dp_bn_bwd_apply_nopool_tma_ex(const void* y, const void* dout, const float* mean, const float* invstd,
                                      const float* gamma, const float* beta, const float* sums, void* dy,
                                      float* dgamma, float* dbeta, int B, int H, int W, int C, int dtype,
                                      b200dp_stream_t stream);
int b200dp_bn_bwd_reduce_from_out_tma_ex(const void* out, const void* dout, const float* gamma, const float* beta,
                                         float* sums, long long Q, int C, int dtype, b200dp_stream_t stream);
int b200dp_linear_fwd_ex(const void* x, const float* w, const float* b, float* logits, int B, int F, int O, int dtype,
                         b200dp_stream_t stream);
int b200dp_linear_bwd_ex(const void* x, const float* w, const float* dlogits, float* dw, float* db, void* dx, int B,
                         int F, int O, int dtype, b200dp_stream_t stream);
int b200dp_head_fwd_bwd_ex(const void* x, const float* w, const float* bias, const long long* target, float* logits,
                           float* loss, int* correct, float* dlogits, void* dx, float* dw, float* db, int B, int F,
                           int O, float grad_scale, int dtype, b200dp_stream_t stream);

/* stats[0..C) += sum_p y[p][c], stats[C..2C) += sum_p y[p][c]^2 (stats zero on entry). */
int b200dp_bn_stats(const void* y, float* stats, long long P, int C, b200dp_stream_t stream);

/* out = maxpool2x2?(relu(batchnorm(y))).  training=1: batch statistics from `stats`
 * (count B*H*W), saves mean / invstd, updates running_mean / running_var (unbiased)
 * / num_batches_tracked exactly like nn.BatchNorm2d.  training=0: running stats.
 * (ref: master/part1/model.py:16,24-25) */
int b200dp_bn_relu_pool_fwd(const void* y, const float* stats, const float* gamma, const float* beta,
                            float* running_mean, float* running_var, long long* num_batches_tracked,
                            float* save_mean, float* save_invstd, void* out, int B, int H, int W, int C,
                            int pool, int training, float momentum, float eps, b200dp_stream_t stream);

/* Backward of the fused block, pass 1: sums[0..C) += sum dz, sums[C..2C) += sum dz*xhat. */
int b200dp_bn_relu_pool_bwd_reduce(const void* y, const void* dout, const float* mean, const float* invstd,
                                   const float* gamma, const float* beta, float* sums, int B, int H, int W,
                                   int C, int pool, b200dp_stream_t stream);
/* pass 1 from the block OUTPUT `out` (= maxpool?(relu(bn(y))), same shape as dout; Q = its pixel count):
 * S1 = sum_{out>0} dout, S2 = sum_{out>0} dout*(out-beta)/gamma.  Reads two quarter-size tensors on
 * pooled layers instead of the full-resolution y. */
int b200dp_bn_relu_pool_bwd_reduce_from_out(const void* out, const void* dout, const float* gamma,
                                            const float* beta, float* sums, long long Q, int C,
                                            b200dp_stream_t stream);
/* pass 2: dy = gamma*invstd*(dz - S1/P - xhat*S2/P); dgamma = S2, dbeta = S1 (either may be NULL). */
int b200dp_bn_relu_pool_bwd_apply(const void* y, const void* dout, const float* mean, const float* invstd,
                                  const float* gamma, const float* beta, const float* sums, void* dy,
                                  float* dgamma, float* dbeta, int B, int H, int W, int C, int pool,
                                  b200dp_stream_t stream);

/* BatchNorm + ReLU (+ pool) backward that allocates nothing and needs no scratch from the caller.
 * This entry point has no counterpart in the reference.  It is for hosts that cannot call
 * cudaMalloc where the backward runs -- all device memory already belongs to the host
 * application's own arena or to a fixed memory plan, or the implicit device synchronisation of an
 * allocation is not tolerable there -- and therefore cannot supply the [2][C] `sums` buffer that
 * b200dp_bn_relu_pool_bwd_reduce/_apply require.  Both passes run back to back on `stream`; the
 * per-channel sums live in a 16 KiB __device__ array inside the library (present on every GPU that
 * loads it, until exit).  Results equal reduce + apply.
 *   - C must be <= 2048; larger C is rejected with B200DP_EINVAL and a message.
 *   - The first call on each device looks up the array's device address (cudaGetSymbolAddress,
 *     with that device current).  Make that first call before any stream capture begins, and keep
 *     the device that owns `stream` current when calling.
 *   - All calls share the one array; it is cleared at the start of each call.  Two calls that
 *     overlap on the GPU (different streams, or different host threads) race on it -- serialise them.
 * Inputs and outputs are still the caller's; no cudaMalloc / cudaFree / torch allocation happens. */
int b200dp_bn_relu_pool_bwd_noalloc(const void* y, const void* dout, const float* mean, const float* invstd,
                                    const float* gamma, const float* beta, void* dy, float* dgamma,
                                    float* dbeta, int B, int H, int W, int C, int pool, b200dp_stream_t stream);

/* Pooled apply pass of the BatchNorm backward as a bulk-copy (cp.async.bulk) pipeline: one thread streams
 * 20 KiB stages of y / dout into shared memory, 256 threads compute, one thread streams dy back.  Same
 * contract and bit-identical results as b200dp_bn_relu_pool_bwd_apply with pool = 1; (W/2)*(C/8) must divide
 * 256.  b200dp_bn_relu_pool_bwd_apply routes here whenever the shape fits. */
int b200dp_bn_bwd_apply_pool_tma(const void* y, const void* dout, const float* mean, const float* invstd,
                                 const float* gamma, const float* beta, const float* sums, void* dy, float* dgamma,
                                 float* dbeta, int B, int H, int W, int C, b200dp_stream_t stream);

/* The rest of the BatchNorm family as bulk-copy pipelines (csrc/bn_stream.cu; the plain entry points route to
 * them whenever the shape fits).  Same contracts as the kernels they
 * replace: b200dp_bn_relu_pool_bwd_apply (pool = 0), b200dp_bn_relu_pool_fwd and
 * b200dp_bn_relu_pool_bwd_reduce_from_out; the first two are bit-identical, the reduction agrees up to
 * summation order.  C/8 must divide 256 (and (W/2)*(C/8) with pooling). */
int b200dp_bn_bwd_apply_nopool_tma(const void* y, const void* dout, const float* mean, const float* invstd,
                                   const float* gamma, const float* beta, const float* sums, void* dy,
                                   float* dgamma, float* dbeta, int B, int H, int W, int C, b200dp_stream_t stream);
int b200dp_bn_relu_pool_fwd_tma(const void* y, const float* stats, const float* gamma, const float* beta,
                                float* running_mean, float* running_var, long long* num_batches_tracked,
                                float* save_mean, float* save_invstd, void* out, int B, int H, int W, int C, int pool,
                                int training, float momentum, float eps, b200dp_stream_t stream);
int b200dp_bn_bwd_reduce_from_out_tma(const void* out, const void* dout, const float* gamma, const float* beta,
                                      float* sums, long long Q, int C, b200dp_stream_t stream);

/* Classifier and loss (ref: master/part1/model.py:40,45; master/part1/part1.py:36,94). */
int b200dp_linear_fwd(const void* x_bf16, const float* w, const float* b, float* logits, int B, int F, int O,
                      b200dp_stream_t stream);
/* loss = mean CE; dlogits = (softmax - onehot) * grad_scale / B; correct = #argmax==target.
 * loss, dlogits and correct may each be NULL. */
int b200dp_cross_entropy(const float* logits, const long long* target, float* loss, float* dlogits,
                         int* correct, int B, int O, float grad_scale, b200dp_stream_t stream);
int b200dp_linear_bwd(const void* x_bf16, const float* w, const float* dlogits, float* dw, float* db,
                      void* dx_bf16, int B, int F, int O, b200dp_stream_t stream);
/* The three calls above as ONE launch (what a training step runs): logits, mean cross-entropy, top-1 count, dlogits (may be NULL), dx, and
 * dw += dlogits^T x, db += sum dlogits.  dw / db are accumulated onto (zero them first), loss / correct
 * are overwritten.  F % 8 == 0, F <= 2048, O <= 16, B <= 4096; one stream at a time. */
int b200dp_head_fwd_bwd(const void* x_bf16, const float* w, const float* bias, const long long* target,
                        float* logits, float* loss, int* correct, float* dlogits, void* dx_bf16, float* dw,
                        float* db, int B, int F, int O, float grad_scale, b200dp_stream_t stream);

/* SGD with momentum and weight decay on flat fp32 buffers:
 * d = g*grad_scale + wd*p; buf = momentum*buf + d; p -= lr*buf.  buf starts at zero,
 * which reproduces torch.optim.SGD's first step.  (ref: master/part1/part1.py:38,98-99) */
int b200dp_sgd_momentum(float* p, const float* g, float* buf, long long n, float lr, float momentum, float wd,
                        float grad_scale, b200dp_stream_t stream);

/* fp32 master [Cout][taps][Cin] -> bf16 Wf [Cout][kpad] (zero padded) and, if wd != NULL,
 * bf16 Wd [Cin][taps*Cout] with the taps flipped (the dgrad operand). */
/* Clear nbytes (multiple of 16) at dst (16-byte aligned): the per-step clear of the flat gradient buffer and
 * of the statistics region, as a kernel of this library so that a step launches no framework fill
 * (ref: optimizer.zero_grad() at master/part1/part1.py:34). */
int b200dp_fill_zero(void* dst, long long nbytes, b200dp_stream_t stream);
/* Copy two device ranges (multiples of 16 bytes; either may be empty) in one launch: a step's staged images
 * and labels into the static buffers of the captured step (ref: data.to(device) at master/part1/part1.py:32). */
int b200dp_copy_pair(void* dst0, const void* src0, long long nbytes0, void* dst1, const void* src1,
                     long long nbytes1, b200dp_stream_t stream);

int b200dp_prep_conv_weights(const float* master, void* wf, void* wd, int Cout, int Cin, int taps, int kpad,
                             b200dp_stream_t stream);
/* The same for n layers in ONE launch (tiled, coalesced, transposed through shared memory);
 * every layer needs Cout % 64 == 0, Cin % 64 == 0 and kpad == taps*Cin. */
/* Precision-generic forms: dtype 0 = bf16 operand copies, 1 = fp32 copies for the tf32 kernels (wf may then be NULL
 * where the master tensor itself is the forward operand, i.e. whenever kpad == taps * Cin). */
int b200dp_prep_conv_weights_ex(const float* master, void* wf, void* wd, int Cout, int Cin, int taps, int kpad,
                                int dtype, b200dp_stream_t stream);
int b200dp_prep_conv_weights_batched_ex(int n, const float* const* master, void* const* wf, void* const* wd,
                                        const int* cout, const int* cin, const int* taps, const int* kpad, int dtype,
                                        b200dp_stream_t stream);
int b200dp_prep_conv_weights_batched(int n, const float* const* master, void* const* wf, void* const* wd,
                                     const int* cout, const int* cin, const int* taps, const int* kpad,
                                     b200dp_stream_t stream);

/* ------------------------------------------------------------------------
 * Rendezvous store (TCP key-value, rank 0 hosts it on master_ip:port)
 * (ref: master/part2a/part2a.py:80-85 init_process_group via MASTER_ADDR/MASTER_PORT)
 * ------------------------------------------------------------------------ */
int b200dp_store_create(const char* master_ip, int port, int rank, int world, int timeout_ms, void** out_store);
int b200dp_store_set(void* store, const char* key, const void* data, int len);
/* Blocks until the key exists (or the store's timeout expires -> B200DP_ETIMEDOUT). */
int b200dp_store_get(void* store, const char* key, void* buf, int cap, int* out_len);
int b200dp_store_add(void* store, const char* key, long long delta, long long* result);
int b200dp_store_barrier(void* store, const char* name);
/* Erase every key that starts with `prefix` (finished barrier rounds, host-side reductions). */
int b200dp_store_delete_prefix(void* store, const char* prefix);
/* The timeout given to b200dp_store_create, in milliseconds. */
int b200dp_store_timeout_ms(void* store);
int b200dp_store_destroy(void* store);

/* ------------------------------------------------------------------------
 * Communicator: symmetric heap mapped into every rank over CUDA IPC + signal pad.
 * Offsets passed to the collectives are BYTE offsets from the heap base and are
 * identical on every rank.  `channel` (0..3) selects an independent set of flags
 * and epoch counter; launches on one channel must be stream-ordered on each rank
 * and issued in the same order by all ranks.
 * ------------------------------------------------------------------------ */
int b200dp_comm_create(void* store, int rank, int world, int device, long long heap_bytes, void** out_comm);
/* flags bit 0: allocate the heap with the CUDA virtual-memory API (cuMemCreate, POSIX-fd exchange over
 * Unix sockets) and map an NVSwitch MULTICAST view of it, which enables the NVLS all-reduce (algo 3). */
int b200dp_comm_create_ex(void* store, int rank, int world, int device, long long heap_bytes, int flags,
                          void** out_comm);
int b200dp_comm_has_multicast(void* comm);
/* Let algo 0 of b200dp_allreduce (and the bucket reducer) pick the NVLS kernel for ranges > 256 KiB. */
int b200dp_comm_prefer_nvls(void* comm, int on);
int b200dp_comm_destroy(void* comm);
int b200dp_comm_heap(void* comm, void** base, long long* usable_bytes, long long* data_offset);
int b200dp_comm_peer_base(void* comm, int rank, void** base);
int b200dp_comm_set_timeout(void* comm, double seconds);
int b200dp_comm_barrier(void* comm, int channel, b200dp_stream_t stream);

/* Gradient all-reduce (mean) over elements [start, end) of the flat fp32 gradient buffer at
 * byte offset grad_off (ref: master/part2b/part2b.py:43-45; DDP buckets master/part3/part3.py:116).
 * algo 0 = choose by size, 1 = one-shot (every rank reduces everything), 2 = two-shot
 * (reduce-scatter + all-gather), 3 = NVLS: multimem.ld_reduce sums the owned slice inside the
 * NVSwitch and multimem.st writes the result to every rank (needs a multicast communicator).  prescale=1 divides before summing (part2b / DDP order).
 * fuse_sgd=1 applies SGD(momentum, weight decay) in the epilogue to the parameters at
 * param_off / momentum at mom_off; the two-shot variant then all-gathers updated PARAMETERS. */
int b200dp_allreduce(void* comm, long long grad_off, long long param_off, long long mom_off, long long start,
                     long long end, int fuse_sgd, int prescale, float lr, float momentum, float weight_decay,
                     int algo, int blocks, int channel, b200dp_stream_t stream);

/* Owner-side half of the fused wgrad -> reduce-scatter path (optional): bucket [start, end); inside
 * [pushed_lo, pushed_hi) the mean gradient of my slice is (own + W-1 staged copies) / W, elsewhere it is
 * pulled from the peers; then SGD and the updated parameters go to every rank (nvls = 1: multimem.st). */
int b200dp_reduce_staged_sgd(void* comm, long long grad_off, long long param_off, long long mom_off,
                             long long stage_buf_off, long long start, long long end, long long pushed_lo,
                             long long pushed_hi, long long stage_base, float lr, float momentum,
                             float weight_decay, int nvls, int blocks, int channel, b200dp_stream_t stream);

/* Parameter server: `root` sums every rank's gradient in rank order starting from zero, divides
 * by the world size and stores the mean into every rank's gradient buffer
 * (ref: master/part2a/part2a.py:42-52).  push=1: workers first store their gradients into the
 * root's staging area at stage_off ((world-1) * range bytes), the isend/irecv data flow of
 * master/part2a/part2a_extra.py:41-58. */
int b200dp_ps_reduce_bcast(void* comm, long long grad_off, long long stage_off, long long start, long long end,
                           int root, int push, int blocks, int channel, b200dp_stream_t stream);

/* Copy nbytes (multiple of 16) at heap byte offset byte_off from `root` to every rank
 * (DDP module-state and per-forward BatchNorm buffer sync). */
int b200dp_broadcast(void* comm, long long byte_off, long long nbytes, int root, int blocks, int channel,
                     b200dp_stream_t stream);

/* All-gather of owner slices: elements [start, end) of the flat fp32 buffer at heap byte offset byte_off are
 * owned in `world` equal float4 slices (ceil(n4 / world) each, the slicing of the two-shot / NVLS collectives);
 * every rank stores its slice into every peer's copy.  Assembles the SGD momentum -- which the fused
 * collectives keep only on a slice's owner -- before a checkpoint, so that the saved optimizer state equals
 * torch.optim.SGD's (ref: optimizer at master/part3/part3.py:117-118). */
int b200dp_allgather_slices(void* comm, long long byte_off, long long start, long long end, int blocks, int channel,
                            b200dp_stream_t stream);

/* Bucket reducer (the role of DDP's C++ Reducer, ref: master/part3/part3.py:116): per-bucket
 * all-reduce(+SGD) on a high-priority side stream with a capped CTA count, forked from and
 * joined back into the compute stream with events (CUDA-graph capturable). */
int b200dp_reducer_create(void* comm, int n_buckets, const long long* starts, const long long* ends,
                          long long grad_off, long long param_off, long long mom_off, int fuse_sgd,
                          int prescale, float lr, float momentum, float weight_decay, int blocks, int channel,
                          void** out_reducer);
int b200dp_reducer_set_lr(void* reducer, float lr);
int b200dp_reducer_bucket_ready(void* reducer, int bucket, b200dp_stream_t compute_stream);
int b200dp_reducer_finalize(void* reducer, b200dp_stream_t compute_stream);
int b200dp_reducer_destroy(void* reducer);
/* The reducer's side stream (cudaStream_t): work queued on it after bucket_ready(k) runs behind bucket
 * k's collective and is joined by finalize(). */
int b200dp_reducer_stream(void* reducer, void** stream_out);
/* CTAs per bucket collective for subsequent bucket_ready() calls (re-capture the step afterwards). */
int b200dp_reducer_set_blocks(void* reducer, int blocks);
/* on = 1 (the default): all bucket kernels of a step but the last drop their exit barrier (implied by the entry
 * barrier of the next bucket kernel); requires bucket_ready() in index order. */
int b200dp_reducer_set_lazy_exit(void* reducer, int on);
/* Optional path: per bucket, the element range delivered by b200dp_conv_wgrad_push and its staging area;
 * such buckets run b200dp_reduce_staged_sgd instead of the all-reduce (pushed_lo == pushed_hi: unchanged). */
int b200dp_reducer_set_staged(void* reducer, long long stage_buf_off, const long long* pushed_lo,
                              const long long* pushed_hi, const long long* stage_base);

/* Native batch loader (csrc/loader.cpp; the role of torch's DataLoader worker processes for the reference):
 * a pool of threads turns uint8 HWC CIFAR images into normalised fp32 NCHW batches -- RandomCrop(32,
 * padding 4) + RandomHorizontalFlip + ToTensor + Normalize in one table-lookup pass -- inside caller-owned
 * (pinned) buffers, in order, `ring` batches ahead.  Host-only.
 *   images: uint8 [n][32][32][3], labels: int64 [n] (both must outlive the loader); xs / ys: `ring` output
 *   buffers of batch*3*32*32 floats / batch int64; mean / std: 3 floats each.
 *   start_epoch: `order` = the epoch's sample indices in consumption order (sharding / shuffling is the
 *   caller's).  next: blocks for the next batch, returns 1 when the epoch is exhausted.  release: the
 *   slot's contents are no longer needed (its host->device copy has completed). */
int b200dp_loader_create(const void* images, const long long* labels, long long n, int batch, int ring,
                         void* const* xs, void* const* ys, const float* mean, const float* std_, int augment,
                         unsigned long long seed, int threads, void** out_loader);
int b200dp_loader_start_epoch(void* loader, const long long* order, long long count, long long epoch,
                              int drop_last);
int b200dp_loader_next(void* loader, int* slot, int* count);
int b200dp_loader_release(void* loader, int slot);
int b200dp_loader_destroy(void* loader);

/* Descriptor hand-off between the ranks of one host (SCM_RIGHTS over abstract Unix sockets); how the
 * virtual-memory heap and the NVSwitch multicast object are shared.  Host-only, testable without a GPU.
 * b200dp_fd_recv returns the descriptor, -2 on timeout (timeout_ms < 0: wait for ever), -1 on error. */
int b200dp_fd_listen(const char* name);
int b200dp_fd_send(const char* name, int fd, int tag, int timeout_ms);
int b200dp_fd_recv(int listen_fd, int* tag, int timeout_ms);

#ifdef __cplusplus
}
#endif
#endif /* B200DP_H_ */
