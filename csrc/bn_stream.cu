// BatchNorm / ReLU / pool kernels as bulk-copy (cp.async.bulk, the TMA engine without a tensor map)
// pipelines.  These are what the plain BatchNorm entry points run whenever the shape fits the stage geometry
// (measured in round 2: 0.700 -> 0.668 ms per step against the register-file kernels they replaced).
//
// Why: the register-file kernels in elementwise.cu sit at 45-56 % of the L2 / HBM rate with the issue
// slots half idle (ncu, profiles/RESULTS_round1.md section 5): every warp alternates between waiting for its own
// loads and doing arithmetic.  NHWC activations are contiguous, so a stage of work is ONE contiguous chunk
// per tensor; one thread keeps a ring of such chunks in flight per CTA, all 256 threads compute from
// shared memory, and bulk stores drain the results -- loads, arithmetic and stores of different stages
// overlap without relying on occupancy.  The pooled backward apply pass lives in elementwise.cu
// (bn_bwd_apply_pool_tma_kernel, written first); this file holds the generic skeleton and the remaining
// members of the family.  Every op keeps the arithmetic of the kernel it replaces, statement for
// statement; tests/test_gpu_kernels.py::test_bn_relu_pool_fwd_bwd checks both families against torch.
//
// No counterpart in the reference (its BatchNorm is ATen's; ref: master/part1/model.py:20-22).
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "../include/b200dp.h"
#include "common.cuh"
#include "host_util.h"

namespace b200dp {
namespace {

struct V8 {
  float v[8];
};

template <class T>
B200DP_DEVINL V8 ld_v8(const T* p) {
  const F8 f = ld8(p);
  V8 r;
#pragma unroll
  for (int j = 0; j < 8; ++j) r.v[j] = f.v[j];
  return r;
}
template <class T>
B200DP_DEVINL void st_v8(T* p, const V8& r) {
  F8 f;
#pragma unroll
  for (int j = 0; j < 8; ++j) f.v[j] = r.v[j];
  st8(p, f);
}
// work items per thread and stage, chosen so that a stage holds the same BYTES for bf16 and fp32 activations
template <class T>
constexpr int items_2in() { return sizeof(T) == 4 ? 1 : 2; }   // two input streams (apply, reduce)
template <class T>
constexpr int items_1in() { return sizeof(T) == 4 ? 2 : 4; }   // one input stream (forward, no pool)

B200DP_DEVINL V8 ld_f32x8(const float* p) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  const float4 b = *reinterpret_cast<const float4*>(p + 4);
  V8 r;
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
  r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}

B200DP_DEVINL V8 ld_cg_f32x8(const float* p) {   // two 16-byte L2 loads (values written by atomics)
  const float4 a = __ldcg(reinterpret_cast<const float4*>(p));
  const float4 b = __ldcg(reinterpret_cast<const float4*>(p + 4));
  V8 r;
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
  r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}

constexpr int kThreads = 256;
constexpr int kStages = 3;

// ---------------------------------------------------------------------------
// Skeleton.  The tensor is cut into `units` (an op-defined number of contiguous bytes per stream:
// a pixel, or a pooled row); a stage holds `units_per_stage` consecutive units.  Streams: A and B are
// inputs (B optional), OUT is optional.  Op supplies
//   Op(const Params&, int tid)                      per-thread constants (after griddepcontrol.wait)
//   void compute(a, b, out, units_valid)            this thread's share of one stage, shared memory only
//   void finish(const Params&, float* scratch)      after the last stage (reductions); scratch = stage ring
// and the geometry below.
// ---------------------------------------------------------------------------
struct StreamGeom {
  long units;               // total units in the tensor
  int units_per_stage;      // G
  int a_bytes, b_bytes, out_bytes;   // per unit (b / out may be 0)
};

template <class Op>
__global__ void __launch_bounds__(kThreads, 2) bn_stream_kernel(const typename Op::Params p, const StreamGeom g,
                                                                const typename Op::elem* __restrict__ a_ptr,
                                                                const typename Op::elem* __restrict__ b_ptr,
                                                                typename Op::elem* __restrict__ out_ptr) {
  using elem = typename Op::elem;
  extern __shared__ uint8_t stream_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(stream_raw) + 127) & ~uintptr_t(127));
  const int a_stage = g.units_per_stage * g.a_bytes, b_stage = g.units_per_stage * g.b_bytes;
  const int o_stage = g.units_per_stage * g.out_bytes;
  const int in_stage = a_stage + b_stage;
  uint8_t* out_buf = smem + kStages * in_stage;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(out_buf + 2 * o_stage);
  pdl_launch_dependents();
  const long num_chunks = (g.units + g.units_per_stage - 1) / g.units_per_stage;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) mbar_init(&full_bar[s], 1);
    mbar_fence_init();
  }
  __syncthreads();
  pdl_wait();

  auto issue = [&](long j, int stage) {
    const long chunk_id = blockIdx.x + j * static_cast<long>(gridDim.x);
    if (chunk_id >= num_chunks) return;
    const long u0 = chunk_id * g.units_per_stage;
    const int n = static_cast<int>(min(static_cast<long>(g.units_per_stage), g.units - u0));
    uint8_t* dst = smem + stage * in_stage;
    const uint32_t ab = static_cast<uint32_t>(n) * g.a_bytes, bb = static_cast<uint32_t>(n) * g.b_bytes;
    mbar_arrive_expect_tx(&full_bar[stage], ab + bb);
    bulk_load_1d(dst, reinterpret_cast<const uint8_t*>(a_ptr) + u0 * g.a_bytes, ab, &full_bar[stage]);
    if (bb != 0) bulk_load_1d(dst + a_stage, reinterpret_cast<const uint8_t*>(b_ptr) + u0 * g.b_bytes, bb, &full_bar[stage]);
  };
  if (threadIdx.x == 0)
    for (int s = 0; s < kStages; ++s) issue(s, s);

  Op op(p, threadIdx.x);

  int stage = 0;
  uint32_t phase = 0;
  for (long j = 0;; ++j) {
    const long chunk_id = blockIdx.x + j * static_cast<long>(gridDim.x);
    if (chunk_id >= num_chunks) break;
    const long u0 = chunk_id * g.units_per_stage;
    const int n = static_cast<int>(min(static_cast<long>(g.units_per_stage), g.units - u0));
    uint8_t* ob = out_buf + (j & 1) * o_stage;
    if (o_stage != 0 && threadIdx.x == 0) bulk_wait_group_read<1>();   // store of two chunks ago has drained `ob`
    mbar_wait(&full_bar[stage], phase);
    __syncthreads();
    const uint8_t* in = smem + stage * in_stage;
    op.compute(reinterpret_cast<const elem*>(in), reinterpret_cast<const elem*>(in + a_stage),
               reinterpret_cast<elem*>(ob), n);
    if (o_stage != 0) fence_proxy_async_smem();
    __syncthreads();   // outputs written, input stage no longer read
    if (threadIdx.x == 0) {
      if (o_stage != 0) {
        bulk_store_1d(reinterpret_cast<uint8_t*>(out_ptr) + u0 * g.out_bytes, ob, static_cast<uint32_t>(n) * g.out_bytes);
        bulk_commit_group();
      }
      issue(j + kStages, stage);
    }
    if (++stage == kStages) {
      stage = 0;
      phase ^= 1;
    }
  }
  if (o_stage != 0 && threadIdx.x == 0) bulk_wait_group<0>();
  __syncthreads();
  op.finish(p, reinterpret_cast<float*>(smem));
}

template <class Op>
int launch_stream(const typename Op::Params& p, const StreamGeom& g, const void* a, const void* b, void* out,
                  cudaStream_t s) {
  const size_t smem = static_cast<size_t>(kStages) * g.units_per_stage * (g.a_bytes + g.b_bytes) +
                      2u * g.units_per_stage * g.out_bytes + kStages * 8 + 128 + 64;
  static size_t configured = 0;
  if (smem > configured) {
    B200DP_CUDA_TRY(cudaFuncSetAttribute(bn_stream_kernel<Op>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem)));
    configured = smem;
  }
  const long num_chunks = (g.units + g.units_per_stage - 1) / g.units_per_stage;
  const int grid = static_cast<int>(num_chunks < 148 * 2 ? num_chunks : 148 * 2);   // 3 per SM: no gain (A/B)
  using elem = typename Op::elem;
  B200DP_CUDA_TRY(launch_pdl(bn_stream_kernel<Op>, dim3(grid), dim3(kThreads), smem, s, p, g,
                             static_cast<const elem*>(a), static_cast<const elem*>(b), static_cast<elem*>(out)));
  return 0;
}

// ---------------------------------------------------------------------------
// op: BatchNorm + ReLU backward, apply pass, no pooling  (bn_bwd_apply_nopool_kernel)
//   unit = one pixel (C channels); 4 items per thread per stage
// ---------------------------------------------------------------------------
struct ApplyParams {
  const float *mean, *invstd, *gamma, *beta, *sums;
  float *dgamma, *dbeta;
  int C;
  long count;   // B*H*W of the conv output
};

template <class T>
struct ApplyNoPoolOp {
  using Params = ApplyParams;
  using elem = T;
  static constexpr int kItemsPerThread = items_2in<T>();   // 2 x 8 KiB per stage, 64 KiB per CTA, 3 CTAs per SM
  V8 scale, shift, c0, c1;
  int C, off0, pix0, pix_step;
  __device__ ApplyNoPoolOp(const Params& p, int tid) {
    const int chunks = p.C / 8;
    const int chunk = tid % chunks, c0i = chunk * 8;
    C = p.C;
    pix0 = tid / chunks;
    pix_step = kThreads / chunks;
    off0 = c0i;
    const float inv_count = 1.f / static_cast<float>(p.count);
    const V8 mean = ld_f32x8(p.mean + c0i), invstd = ld_f32x8(p.invstd + c0i);
    const V8 g = ld_f32x8(p.gamma + c0i), b = ld_f32x8(p.beta + c0i);
    const V8 s1 = ld_cg_f32x8(p.sums + c0i), s2 = ld_cg_f32x8(p.sums + p.C + c0i);   // L2: atomics wrote them
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      scale.v[j] = g.v[j] * invstd.v[j];
      shift.v[j] = b.v[j] - mean.v[j] * scale.v[j];
      c1.v[j] = -scale.v[j] * invstd.v[j] * s2.v[j] * inv_count;
      c0.v[j] = -scale.v[j] * s1.v[j] * inv_count - c1.v[j] * mean.v[j];
    }
    if (blockIdx.x == 0 && pix0 == 0 && p.dgamma != nullptr) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        p.dgamma[c0i + j] = s2.v[j];
        p.dbeta[c0i + j] = s1.v[j];
      }
    }
  }
  __device__ void compute(const T* y, const T* dout, T* dy, int n) const {
#pragma unroll
    for (int u = 0; u < kItemsPerThread; ++u) {
      const int pix = pix0 + u * pix_step;
      if (pix >= n) break;
      const int off = pix * C + off0;
      const V8 yv = ld_v8(y + off);
      const V8 go = ld_v8(dout + off);
      V8 o;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float z = fmaf(yv.v[j], scale.v[j], shift.v[j]);
        const float dz = z > 0.f ? go.v[j] : 0.f;
        o.v[j] = fmaf(scale.v[j], dz, fmaf(c1.v[j], yv.v[j], c0.v[j]));
      }
      st_v8(dy + off, o);
    }
  }
  __device__ void finish(const Params&, float*) const {}
};

// ---------------------------------------------------------------------------
// op: BatchNorm(train or eval) + ReLU (+ 2x2 max-pool) forward  (bn_relu_pool_fwd_kernel)
//   POOL: unit = one pooled row of one image (two image rows of y in, Wo*C out); one item per thread
//   else: unit = one pixel; kFwdItems items per thread
// ---------------------------------------------------------------------------
struct FwdParams {
  const float *stats, *gamma, *beta;
  float *running_mean, *running_var;
  long long* num_batches_tracked;
  float *save_mean, *save_invstd;
  int W, C, training;
  long count;   // B*H*W
  float momentum, eps;
};

template <class T, bool POOL>
struct FwdOp {
  using Params = FwdParams;
  using elem = T;
  static constexpr int kFwdItems = items_1in<T>();   // one input stream: 16 KiB in + 16 KiB out per stage
  V8 scale, shift;
  int C, c0i, pix0, pix_step, y_off, o_off, rowb, gt;
  __device__ FwdOp(const Params& p, int tid) {
    const int chunks = p.C / 8;
    const int chunk = tid % chunks;
    C = p.C;
    c0i = chunk * 8;
    pix0 = tid / chunks;
    pix_step = kThreads / chunks;
    rowb = p.W * p.C;
    if (POOL) {
      const int Wo = p.W / 2;
      gt = pix0 / Wo;
      const int wo = pix0 - gt * Wo;
      y_off = gt * 2 * rowb + (2 * wo) * p.C + c0i;
      o_off = gt * Wo * p.C + wo * p.C + c0i;
    } else {
      gt = 0;
      y_off = o_off = c0i;
    }
    const float count = static_cast<float>(p.count);
    const V8 g = ld_f32x8(p.gamma + c0i), b = ld_f32x8(p.beta + c0i);
    if (p.training) {
      const V8 s1 = ld_f32x8(p.stats + c0i), s2 = ld_f32x8(p.stats + p.C + c0i);
      V8 mean, var, invstd;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        mean.v[j] = s1.v[j] / count;
        var.v[j] = fmaxf(s2.v[j] / count - mean.v[j] * mean.v[j], 0.f);
        invstd.v[j] = rsqrtf(var.v[j] + p.eps);
        scale.v[j] = g.v[j] * invstd.v[j];
        shift.v[j] = b.v[j] - mean.v[j] * scale.v[j];
      }
      if (blockIdx.x == 0 && pix0 == 0) {
        const float unbias = count > 1.f ? count / (count - 1.f) : 1.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          p.save_mean[c0i + j] = mean.v[j];
          p.save_invstd[c0i + j] = invstd.v[j];
          p.running_mean[c0i + j] = (1.f - p.momentum) * p.running_mean[c0i + j] + p.momentum * mean.v[j];
          p.running_var[c0i + j] = (1.f - p.momentum) * p.running_var[c0i + j] + p.momentum * var.v[j] * unbias;
        }
        if (chunk == 0 && p.num_batches_tracked != nullptr) *p.num_batches_tracked += 1;
      }
    } else {
      const V8 rm = ld_f32x8(p.running_mean + c0i), rv = ld_f32x8(p.running_var + c0i);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        scale.v[j] = g.v[j] * rsqrtf(rv.v[j] + p.eps);
        shift.v[j] = b.v[j] - rm.v[j] * scale.v[j];
      }
    }
  }
  __device__ void compute(const T* y, const T*, T* out, int n) const {
    if (POOL) {
      if (gt >= n) return;
      const V8 a = ld_v8(y + y_off);
      const V8 bq = ld_v8(y + y_off + C);
      const V8 c = ld_v8(y + y_off + rowb);
      const V8 d = ld_v8(y + y_off + rowb + C);
      V8 o;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float za = fmaf(a.v[j], scale.v[j], shift.v[j]), zb = fmaf(bq.v[j], scale.v[j], shift.v[j]);
        const float zc = fmaf(c.v[j], scale.v[j], shift.v[j]), zd = fmaf(d.v[j], scale.v[j], shift.v[j]);
        o.v[j] = fmaxf(fmaxf(fmaxf(za, zb), fmaxf(zc, zd)), 0.f);
      }
      st_v8(out + o_off, o);
    } else {
#pragma unroll
      for (int u = 0; u < kFwdItems; ++u) {
        const int pix = pix0 + u * pix_step;
        if (pix >= n) break;
        const int off = pix * C + c0i;
        const V8 a = ld_v8(y + off);
        V8 o;
#pragma unroll
        for (int j = 0; j < 8; ++j) o.v[j] = fmaxf(fmaf(a.v[j], scale.v[j], shift.v[j]), 0.f);
        st_v8(out + off, o);
      }
    }
  }
  __device__ void finish(const Params&, float*) const {}
};

// ---------------------------------------------------------------------------
// op: BatchNorm-backward reduction from the block output  (bn_bwd_reduce_from_out_kernel)
//   S1 = sum_{a>0} g,  S2 = sum_{a>0} g * (a - beta) / gamma   per channel, added to sums[0..C) / [C..2C)
//   unit = one pixel of the (pooled) output; two input streams, no output stream.  The per-thread partial
//   sums live in registers across all stages; finish() combines the threads of a block in shared memory
//   and issues one atomic per channel per block.  (Summation order differs from the register-file
//   kernel, so the sums agree to rounding, not bit for bit.)
// ---------------------------------------------------------------------------
struct ReduceParams {
  const float *gamma, *beta;
  float* sums;
  int C;
};

template <class T>
struct ReduceFromOutOp {
  using Params = ReduceParams;
  using elem = T;
  static constexpr int kItemsPerThread = items_2in<T>();
  V8 bt, ig, a1, a2;
  int C, c0i, pix0, pix_step;
  __device__ ReduceFromOutOp(const Params& p, int tid) {
    const int chunks = p.C / 8;
    C = p.C;
    c0i = (tid % chunks) * 8;
    pix0 = tid / chunks;
    pix_step = kThreads / chunks;
    const V8 g = ld_f32x8(p.gamma + c0i);
    bt = ld_f32x8(p.beta + c0i);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      ig.v[j] = fabsf(g.v[j]) > 1e-20f ? 1.f / g.v[j] : 0.f;
      a1.v[j] = a2.v[j] = 0.f;
    }
  }
  __device__ void compute(const T* out, const T* dout, T*, int n) {
#pragma unroll
    for (int u = 0; u < kItemsPerThread; ++u) {
      const int pix = pix0 + u * pix_step;
      if (pix >= n) break;
      const int off = pix * C + c0i;
      const V8 o = ld_v8(out + off);
      const V8 g = ld_v8(dout + off);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float d = o.v[j] > 0.f ? g.v[j] : 0.f;
        a1.v[j] += d;
        a2.v[j] = fmaf(d, (o.v[j] - bt.v[j]) * ig.v[j], a2.v[j]);
      }
    }
  }
  __device__ void finish(const Params& p, float* scratch) const {
    // scratch: [rows][C] x 2 with rows = kThreads / chunks (<= 16 KiB for C <= 2048; the stage ring is idle)
    const int rows = pix_step;
    float* r1 = scratch;
    float* r2 = scratch + rows * C;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      r1[pix0 * C + c0i + j] = a1.v[j];
      r2[pix0 * C + c0i + j] = a2.v[j];
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += kThreads) {
      float sa = 0.f, sb = 0.f;
      for (int r = 0; r < rows; ++r) {
        sa += r1[r * C + c];
        sb += r2[r * C + c];
      }
      atomicAdd(p.sums + c, sa);
      atomicAdd(p.sums + C + c, sb);
    }
  }
};

}  // namespace
}  // namespace b200dp

using namespace b200dp;

static bool chunks_ok(int C) { return C >= 8 && C % 8 == 0 && C <= 2048 && kThreads % (C / 8) == 0; }

template <class T>
static int apply_nopool_impl(const void* y, const void* dout, const float* mean, const float* invstd,
                             const float* gamma, const float* beta, const float* sums, void* dy, float* dgamma,
                             float* dbeta, int B, int H, int W, int C, cudaStream_t s) {
  constexpr int kBytes = static_cast<int>(sizeof(T));
  const long P = static_cast<long>(B) * H * W;
  const ApplyParams p{mean, invstd, gamma, beta, sums, dgamma, dbeta, C, P};
  const StreamGeom g{P, kThreads * ApplyNoPoolOp<T>::kItemsPerThread / (C / 8), kBytes * C, kBytes * C, kBytes * C};
  return launch_stream<ApplyNoPoolOp<T>>(p, g, y, dout, dy, s);
}

/* BatchNorm(+ReLU) backward apply pass without pooling as a bulk-copy pipeline: same contract as
 * b200dp_bn_relu_pool_bwd_apply_ex with pool = 0 (which routes here).  C/8 must divide 256.
 * dtype 0 = bf16 activations, 1 = fp32. */
extern "C" int b200dp_bn_bwd_apply_nopool_tma_ex(const void* y, const void* dout, const float* mean,
                                                 const float* invstd, const float* gamma, const float* beta,
                                                 const float* sums, void* dy, float* dgamma, float* dbeta, int B,
                                                 int H, int W, int C, int dtype, b200dp_stream_t s) {
  if (B <= 0 || H <= 0 || W <= 0 || !chunks_ok(C)) {
    set_error("bn_bwd_apply_nopool_tma: unsupported shape B=%d H=%d W=%d C=%d (C/8 must divide 256)", B, H, W, C);
    return B200DP_EINVAL;
  }
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (dtype == 1)
    return apply_nopool_impl<float>(y, dout, mean, invstd, gamma, beta, sums, dy, dgamma, dbeta, B, H, W, C, st);
  return apply_nopool_impl<__nv_bfloat16>(y, dout, mean, invstd, gamma, beta, sums, dy, dgamma, dbeta, B, H, W, C, st);
}
extern "C" int b200dp_bn_bwd_apply_nopool_tma(const void* y, const void* dout, const float* mean, const float* invstd,
                                              const float* gamma, const float* beta, const float* sums, void* dy,
                                              float* dgamma, float* dbeta, int B, int H, int W, int C,
                                              b200dp_stream_t s) {
  return b200dp_bn_bwd_apply_nopool_tma_ex(y, dout, mean, invstd, gamma, beta, sums, dy, dgamma, dbeta, B, H, W, C, 0, s);
}

template <class T>
static int fwd_impl(const FwdParams& p, const void* y, void* out, int B, int H, int W, int C, int pool,
                    cudaStream_t s) {
  constexpr int kBytes = static_cast<int>(sizeof(T));
  if (pool) {
    const StreamGeom g{static_cast<long>(B) * (H / 2), kThreads / ((W / 2) * (C / 8)), 2 * kBytes * W * C, 0,
                       kBytes * (W / 2) * C};
    return launch_stream<FwdOp<T, true>>(p, g, y, nullptr, out, s);
  }
  const StreamGeom g{static_cast<long>(B) * H * W, kThreads * FwdOp<T, false>::kFwdItems / (C / 8), kBytes * C, 0,
                     kBytes * C};
  return launch_stream<FwdOp<T, false>>(p, g, y, nullptr, out, s);
}

/* BatchNorm (training or eval) + ReLU (+ 2x2 max-pool) forward as a bulk-copy pipeline: same contract as
 * b200dp_bn_relu_pool_fwd_ex (which routes here).  C/8 must divide 256; with pool, (W/2)*(C/8) must. */
extern "C" int b200dp_bn_relu_pool_fwd_tma_ex(const void* y, const float* stats, const float* gamma, const float* beta,
                                              float* running_mean, float* running_var, long long* num_batches_tracked,
                                              float* save_mean, float* save_invstd, void* out, int B, int H, int W,
                                              int C, int pool, int training, float momentum, float eps, int dtype,
                                              b200dp_stream_t s) {
  const bool ok = B > 0 && H > 0 && W > 0 && chunks_ok(C) &&
                  (!pool || (((H | W) & 1) == 0 && (W / 2) * (C / 8) <= kThreads && kThreads % ((W / 2) * (C / 8)) == 0));
  if (!ok) {
    set_error("bn_relu_pool_fwd_tma: unsupported shape B=%d H=%d W=%d C=%d pool=%d", B, H, W, C, pool);
    return B200DP_EINVAL;
  }
  const FwdParams p{stats, gamma, beta, running_mean, running_var, num_batches_tracked, save_mean, save_invstd,
                    W, C, training, static_cast<long>(B) * H * W, momentum, eps};
  cudaStream_t st = static_cast<cudaStream_t>(s);
  return dtype == 1 ? fwd_impl<float>(p, y, out, B, H, W, C, pool, st)
                    : fwd_impl<__nv_bfloat16>(p, y, out, B, H, W, C, pool, st);
}
extern "C" int b200dp_bn_relu_pool_fwd_tma(const void* y, const float* stats, const float* gamma, const float* beta,
                                           float* running_mean, float* running_var, long long* num_batches_tracked,
                                           float* save_mean, float* save_invstd, void* out, int B, int H, int W,
                                           int C, int pool, int training, float momentum, float eps,
                                           b200dp_stream_t s) {
  return b200dp_bn_relu_pool_fwd_tma_ex(y, stats, gamma, beta, running_mean, running_var, num_batches_tracked, save_mean,
                                        save_invstd, out, B, H, W, C, pool, training, momentum, eps, 0, s);
}

/* Reduction pass of the BatchNorm backward from the block output as a bulk-copy pipeline: the contract of
 * b200dp_bn_relu_pool_bwd_reduce_from_out_ex (sums accumulated onto; equal up to summation order). */
extern "C" int b200dp_bn_bwd_reduce_from_out_tma_ex(const void* out, const void* dout, const float* gamma,
                                                    const float* beta, float* sums, long long Q, int C, int dtype,
                                                    b200dp_stream_t s) {
  if (Q <= 0 || !chunks_ok(C)) {
    set_error("bn_bwd_reduce_from_out_tma: unsupported shape Q=%lld C=%d (C/8 must divide 256)", Q, C);
    return B200DP_EINVAL;
  }
  const ReduceParams p{gamma, beta, sums, C};
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (dtype == 1) {
    const StreamGeom g{static_cast<long>(Q), kThreads * ReduceFromOutOp<float>::kItemsPerThread / (C / 8), 4 * C, 4 * C, 0};
    return launch_stream<ReduceFromOutOp<float>>(p, g, out, dout, nullptr, st);
  }
  const StreamGeom g{static_cast<long>(Q), kThreads * ReduceFromOutOp<__nv_bfloat16>::kItemsPerThread / (C / 8), 2 * C,
                     2 * C, 0};
  return launch_stream<ReduceFromOutOp<__nv_bfloat16>>(p, g, out, dout, nullptr, st);
}
extern "C" int b200dp_bn_bwd_reduce_from_out_tma(const void* out, const void* dout, const float* gamma,
                                                 const float* beta, float* sums, long long Q, int C,
                                                 b200dp_stream_t s) {
  return b200dp_bn_bwd_reduce_from_out_tma_ex(out, dout, gamma, beta, sums, Q, C, 0, s);
}
