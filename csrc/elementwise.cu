// Bandwidth-bound kernels of the VGG step: layer-0 patch gather, BatchNorm(train)
// statistics, fused BN + ReLU (+ 2x2 max-pool) forward and backward, the
// 512->10 classifier with softmax cross-entropy, weight re-layout and the SGD
// update.  All activations are bf16 NHWC; every thread works on 8 channels
// (one 16-byte vector) so that loads and stores are fully coalesced.
//
// Semantics follow the torch modules the reference instantiates
// (ref: master/part1/model.py:16,24-25,40; master/part1/part1.py:94,98-99):
// BatchNorm2d eps 1e-5 / momentum 0.1 / biased batch variance for normalisation,
// unbiased for running_var; MaxPool2d(2,2) routes the gradient to the first
// maximum in scan order; CrossEntropyLoss(mean); SGD lr/momentum/weight_decay
// with dampening 0 and weight decay on every parameter.
#include <stdio.h>
#include <stdlib.h>

#include "../include/b200dp.h"
#include "common.cuh"
#include "host_util.h"

namespace b200dp {

struct Vec8 {
  float v[8];
};

B200DP_DEVINL Vec8 load_bf16x8(const __nv_bfloat16* p) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  Vec8 r;
  float2 f;
  f = unpack_bf16x2(u.x); r.v[0] = f.x; r.v[1] = f.y;
  f = unpack_bf16x2(u.y); r.v[2] = f.x; r.v[3] = f.y;
  f = unpack_bf16x2(u.z); r.v[4] = f.x; r.v[5] = f.y;
  f = unpack_bf16x2(u.w); r.v[6] = f.x; r.v[7] = f.y;
  return r;
}

B200DP_DEVINL void store_bf16x8(__nv_bfloat16* p, const Vec8& r) {
  uint4 u;
  u.x = pack_bf16x2(r.v[0], r.v[1]);
  u.y = pack_bf16x2(r.v[2], r.v[3]);
  u.z = pack_bf16x2(r.v[4], r.v[5]);
  u.w = pack_bf16x2(r.v[6], r.v[7]);
  *reinterpret_cast<uint4*>(p) = u;
}

// raw (still packed) eight activations: one 16-byte vector of bf16 or two of fp32
template <class T>
struct Raw8;
template <>
struct Raw8<__nv_bfloat16> {
  uint4 a;
};
template <>
struct Raw8<float> {
  float4 a, b;
};
B200DP_DEVINL Raw8<__nv_bfloat16> ldraw8(const __nv_bfloat16* p) {
  Raw8<__nv_bfloat16> r;
  r.a = *reinterpret_cast<const uint4*>(p);
  return r;
}
B200DP_DEVINL Raw8<float> ldraw8(const float* p) {
  Raw8<float> r;
  r.a = *reinterpret_cast<const float4*>(p);
  r.b = *(reinterpret_cast<const float4*>(p) + 1);
  return r;
}
B200DP_DEVINL Vec8 load8(const __nv_bfloat16* p) { return load_bf16x8(p); }
B200DP_DEVINL void store8(__nv_bfloat16* p, const Vec8& r) { store_bf16x8(p, r); }
B200DP_DEVINL Vec8 load8(const float* p) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  const float4 b = *(reinterpret_cast<const float4*>(p) + 1);
  Vec8 r;
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
  r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}
B200DP_DEVINL void store8(float* p, const Vec8& r) {
  *reinterpret_cast<float4*>(p) = make_float4(r.v[0], r.v[1], r.v[2], r.v[3]);
  *(reinterpret_cast<float4*>(p) + 1) = make_float4(r.v[4], r.v[5], r.v[6], r.v[7]);
}

template <class T>
B200DP_DEVINL T operand_cast(float v) {
  if constexpr (sizeof(T) == 4) return v;
  else return __float2bfloat16(v);
}
template <class T>
B200DP_DEVINL void operand_store4(T* dst, float a, float b, float c, float d) {
  if constexpr (sizeof(T) == 4) {
    *reinterpret_cast<float4*>(dst) = make_float4(a, b, c, d);
  } else {
    uint2 o;
    o.x = pack_bf16x2(a, b);
    o.y = pack_bf16x2(c, d);
    *reinterpret_cast<uint2*>(dst) = o;
  }
}

B200DP_DEVINL Vec8 load_f32x8(const float* p) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  const float4 b = *reinterpret_cast<const float4*>(p + 4);
  Vec8 r;
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
  r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}

B200DP_DEVINL Vec8 load_cg_f32x8(const float* p) {   // two 16-byte L2 loads (values written by atomics)
  const float4 a = __ldcg(reinterpret_cast<const float4*>(p));
  const float4 b = __ldcg(reinterpret_cast<const float4*>(p + 4));
  Vec8 r;
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
  r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}

// ---------------------------------------------------------------------------
// layer 0: gather 3x3x3 patches of the fp32 NCHW input into a bf16 [pixels][64]
// matrix (27 real columns, k = (r*3+s)*3 + c, the rest zero) so that the first
// conv is a plain K=64 GEMM for the tcgen05 kernel.
// ---------------------------------------------------------------------------
// One block per group of kIm2colRows image rows: the (rows+2) x 3 x (W+2) input window is staged in
// shared memory with coalesced loads, then every thread emits 16-byte chunks of the 64-wide patch rows.
constexpr int kIm2colRows = 4;

// blockDim is a multiple of 8, so a thread emits the same 16-byte chunk of every patch row it touches: the
// (tap, channel) -> window offset decode of its 8 columns is done once, outside the pixel loop; the upper
// half of the row (columns 32..63) is a constant zero store.
// KW: columns written per row.  KW < KP leaves columns KW..KP-1 alone -- they are zero in every step, so a caller
// that owns a persistent, zero-initialised matrix only pays for the 32 columns that carry data.
template <class T, int KP, int KW>
__global__ void __launch_bounds__(256) im2col_c3_v2_kernel(const float* __restrict__ x, T* __restrict__ a0, int B,
                                                           int H, int W) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ float win[];  // [rows + 2][3 channels][W + 2]
  constexpr int kChunks = KW / 8;   // 16-byte-aligned groups of 8 columns per patch row (KP = 64 bf16 / 32 fp32)
  const int Wp = W + 2;
  const int groups = (H + kIm2colRows - 1) / kIm2colRows;
  const int n = blockIdx.x / groups, h0 = (blockIdx.x - n * groups) * kIm2colRows;
  const int rows = min(kIm2colRows, H - h0);
  for (int i = threadIdx.x; i < (rows + 2) * 3 * Wp; i += blockDim.x) {
    const int r = i / (3 * Wp), rem = i - r * 3 * Wp;
    const int c = rem / Wp, ww = rem - c * Wp - 1;
    const int hh = h0 + r - 1;
    win[i] = (hh >= 0 && hh < H && ww >= 0 && ww < W) ? __ldg(x + ((static_cast<long>(n) * 3 + c) * H + hh) * W + ww) : 0.f;
  }
  const int chunk = threadIdx.x % kChunks;
  int off[8];
  unsigned live = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int k = chunk * 8 + j;  // k = (r*3 + s)*3 + c
    const bool real = k < 27;
    off[j] = real ? (k / 9) * 3 * Wp + (k % 3) * Wp + (k / 3) % 3 : 0;
    live |= static_cast<unsigned>(real) << j;
  }
  __syncthreads();
  T* dst_base = a0 + (static_cast<long>(n) * H + h0) * W * KP + chunk * 8;
  for (int pix = threadIdx.x / kChunks; pix < rows * W; pix += blockDim.x / kChunks) {
    Vec8 o;
#pragma unroll
    for (int j = 0; j < 8; ++j) o.v[j] = 0.f;
    if (live != 0) {
      const int pr = pix / W, w = pix - pr * W;
      const float* base = win + pr * 3 * Wp + w;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float t = base[off[j]];
        o.v[j] = ((live >> j) & 1u) ? t : 0.f;
      }
    }
    store8(dst_base + static_cast<long>(pix) * KP, o);
  }
}

// ---------------------------------------------------------------------------
// per-channel sum / sum of squares of a bf16 [P][C] matrix (stand-alone variant;
// the conv epilogue can produce the same numbers fused).
// ---------------------------------------------------------------------------
template <class T>
__global__ void bn_stats_kernel(const T* __restrict__ y, float* __restrict__ stats, long P, int C) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ float red[];  // [rows_per_block][C] x 2
  const int chunks = C / 8;
  const int rows_per_block = blockDim.x / chunks;
  const int chunk = threadIdx.x % chunks;
  const int rl = threadIdx.x / chunks;
  Vec8 s1, s2;
#pragma unroll
  for (int j = 0; j < 8; ++j) s1.v[j] = s2.v[j] = 0.f;
  if (rl < rows_per_block) {
    for (long r = static_cast<long>(blockIdx.x) * rows_per_block + rl; r < P;
         r += static_cast<long>(gridDim.x) * rows_per_block) {
      const Vec8 v = load8(y + r * C + chunk * 8);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        s1.v[j] += v.v[j];
        s2.v[j] += v.v[j] * v.v[j];
      }
    }
  }
  float* r1 = red;
  float* r2 = red + rows_per_block * C;
  if (rl < rows_per_block) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      r1[rl * C + chunk * 8 + j] = s1.v[j];
      r2[rl * C + chunk * 8 + j] = s2.v[j];
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float a = 0.f, b = 0.f;
    for (int r = 0; r < rows_per_block; ++r) {
      a += r1[r * C + c];
      b += r2[r * C + c];
    }
    atomicAdd(stats + c, a);
    atomicAdd(stats + C + c, b);
  }
}

// ---------------------------------------------------------------------------
// BN(train or eval) + ReLU (+ 2x2/2 max-pool) forward
// ---------------------------------------------------------------------------
struct BnFwdParams {   // y / out: activations of type T (bf16 or fp32), selected by the kernel's template argument
  const void* y;
  const float* stats;  // [2][C] sum, sumsq over P = B*H*W (training only)
  const float* gamma;
  const float* beta;
  float* running_mean;
  float* running_var;
  long long* num_batches_tracked;
  float* save_mean;
  float* save_invstd;
  void* out;
  int B, H, W, C, pool, training;
  float momentum, eps;
};

B200DP_DEVINL Vec8 unpack8(const Raw8<float>& u) {
  Vec8 r;
  r.v[0] = u.a.x; r.v[1] = u.a.y; r.v[2] = u.a.z; r.v[3] = u.a.w;
  r.v[4] = u.b.x; r.v[5] = u.b.y; r.v[6] = u.b.z; r.v[7] = u.b.w;
  return r;
}
B200DP_DEVINL Vec8 unpack8(const uint4& u);
B200DP_DEVINL Vec8 unpack8(const Raw8<__nv_bfloat16>& u) { return unpack8(u.a); }

B200DP_DEVINL Vec8 unpack8(const uint4& u) {
  Vec8 r;
  float2 f;
  f = unpack_bf16x2(u.x); r.v[0] = f.x; r.v[1] = f.y;
  f = unpack_bf16x2(u.y); r.v[2] = f.x; r.v[3] = f.y;
  f = unpack_bf16x2(u.z); r.v[4] = f.x; r.v[5] = f.y;
  f = unpack_bf16x2(u.w); r.v[6] = f.x; r.v[7] = f.y;
  return r;
}

// Each thread owns 8 channels (scale/shift stay in registers) and walks output pixels two at a
// time so that up to eight 16-byte loads are in flight per thread.
template <class T>
__global__ void __launch_bounds__(256, 3) bn_relu_pool_fwd_kernel(const BnFwdParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const T* __restrict__ py = static_cast<const T*>(p.y);
  T* __restrict__ pout = static_cast<T*>(p.out);
  const int chunks = p.C / 8;
  const int rows_per_block = blockDim.x / chunks;
  const int chunk = threadIdx.x % chunks;
  const int rl = threadIdx.x / chunks;
  const int c0 = chunk * 8;
  const int Ho = p.pool ? p.H / 2 : p.H, Wo = p.pool ? p.W / 2 : p.W;
  const long Q = static_cast<long>(p.B) * Ho * Wo;
  const float count = static_cast<float>(static_cast<long>(p.B) * p.H * p.W);
  Vec8 scale, shift;
  {
    const Vec8 g = load_f32x8(p.gamma + c0), b = load_f32x8(p.beta + c0);
    if (p.training) {
      const Vec8 s1 = load_f32x8(p.stats + c0), s2 = load_f32x8(p.stats + p.C + c0);
      Vec8 mean, var, invstd;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        mean.v[j] = s1.v[j] / count;
        var.v[j] = fmaxf(s2.v[j] / count - mean.v[j] * mean.v[j], 0.f);
        invstd.v[j] = rsqrtf(var.v[j] + p.eps);
        scale.v[j] = g.v[j] * invstd.v[j];
        shift.v[j] = b.v[j] - mean.v[j] * scale.v[j];
      }
      if (blockIdx.x == 0 && rl == 0) {
        // one thread per channel chunk owns the side outputs and the running statistics
        const float unbias = count > 1.f ? count / (count - 1.f) : 1.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          p.save_mean[c0 + j] = mean.v[j];
          p.save_invstd[c0 + j] = invstd.v[j];
          p.running_mean[c0 + j] = (1.f - p.momentum) * p.running_mean[c0 + j] + p.momentum * mean.v[j];
          p.running_var[c0 + j] = (1.f - p.momentum) * p.running_var[c0 + j] + p.momentum * var.v[j] * unbias;
        }
        if (chunk == 0 && p.num_batches_tracked != nullptr) *p.num_batches_tracked += 1;
      }
    } else {
      const Vec8 rm = load_f32x8(p.running_mean + c0), rv = load_f32x8(p.running_var + c0);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        scale.v[j] = g.v[j] * rsqrtf(rv.v[j] + p.eps);
        shift.v[j] = b.v[j] - rm.v[j] * scale.v[j];
      }
    }
  }
  if (rl >= rows_per_block) return;
  const long stride = static_cast<long>(gridDim.x) * rows_per_block;
  const long rowb = static_cast<long>(p.W) * p.C;
  for (long qa = static_cast<long>(blockIdx.x) * rows_per_block + rl; qa < Q; qa += 2 * stride) {
    const long qb = qa + stride;
    const bool hasb = qb < Q;
    if (p.pool) {
      long off[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const long q = (u == 0 || hasb) ? (u == 0 ? qa : qb) : qa;
        const int wo = static_cast<int>(q % Wo);
        const int ho = static_cast<int>((q / Wo) % Ho);
        const long n = q / (static_cast<long>(Wo) * Ho);
        off[u] = ((n * p.H + 2 * ho) * p.W + 2 * wo) * p.C + c0;
      }
      Raw8<T> raw[2][4];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        raw[u][0] = ldraw8(py + off[u]);
        raw[u][1] = ldraw8(py + off[u] + p.C);
        raw[u][2] = ldraw8(py + off[u] + rowb);
        raw[u][3] = ldraw8(py + off[u] + rowb + p.C);
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        if (u == 1 && !hasb) break;
        const Vec8 a = unpack8(raw[u][0]), bq = unpack8(raw[u][1]), c = unpack8(raw[u][2]), d = unpack8(raw[u][3]);
        Vec8 o;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float za = fmaf(a.v[j], scale.v[j], shift.v[j]), zb = fmaf(bq.v[j], scale.v[j], shift.v[j]);
          const float zc = fmaf(c.v[j], scale.v[j], shift.v[j]), zd = fmaf(d.v[j], scale.v[j], shift.v[j]);
          o.v[j] = fmaxf(fmaxf(fmaxf(za, zb), fmaxf(zc, zd)), 0.f);
        }
        store8(pout + (u == 0 ? qa : qb) * p.C + c0, o);
      }
    } else {
      const Raw8<T> ra = ldraw8(py + qa * p.C + c0);
      const Raw8<T> rb = ldraw8(py + (hasb ? qb : qa) * p.C + c0);
      const Vec8 a = unpack8(ra), b2 = unpack8(rb);
      Vec8 oa, ob;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        oa.v[j] = fmaxf(fmaf(a.v[j], scale.v[j], shift.v[j]), 0.f);
        ob.v[j] = fmaxf(fmaf(b2.v[j], scale.v[j], shift.v[j]), 0.f);
      }
      store8(pout + qa * p.C + c0, oa);
      if (hasb) store8(pout + qb * p.C + c0, ob);
    }
  }
}

// ---------------------------------------------------------------------------
// BN + ReLU (+ pool) backward.  Pass 1 reduces sum(dz) and sum(dz * xhat) per
// channel; pass 2 recomputes dz and writes dY = gamma*invstd*(dz - S1/P - xhat*S2/P).
// dz is the gradient reaching the BN output: the incoming gradient routed through
// the pool arg-max (first maximum in scan order) and masked by ReLU.
// ---------------------------------------------------------------------------
struct BnBwdParams {   // y / dout / dy: activations of type T, selected by the kernel's template argument
  const void* y;
  const void* dout;
  const float* mean;
  const float* invstd;
  const float* gamma;
  const float* beta;
  float* sums;  // [2][C]: S1 = sum dz, S2 = sum dz*xhat
  void* dy;
  float* dgamma;
  float* dbeta;
  int B, H, W, C, pool;
};

// Register-lean formulation (64 regs -> 4 blocks / SM so that enough 16-byte loads are in flight):
//   z  = y*scale + shift                       scale = gamma*invstd, shift = beta - mean*scale
//   pass 1 accumulates  a1 = sum dz,  a2 = sum dz*y ;  S2 = invstd*(a2 - mean*a1) at the end
//   pass 2 writes       dy = scale*dz + c0 + c1*y   c1 = -scale*invstd*S2/P, c0 = -scale*S1/P - c1*mean
template <class T, bool APPLY>
B200DP_DEVINL void bn_relu_pool_bwd_body(const BnBwdParams& p, float* red) {
  const T* __restrict__ py = static_cast<const T*>(p.y);
  const T* __restrict__ pdout = static_cast<const T*>(p.dout);
  T* __restrict__ pdy = static_cast<T*>(p.dy);
  const int chunks = p.C / 8;
  const int rows_per_block = blockDim.x / chunks;
  const int chunk = threadIdx.x % chunks;
  const int rl = threadIdx.x / chunks;
  const int c0i = chunk * 8;
  const int Ho = p.pool ? p.H / 2 : p.H, Wo = p.pool ? p.W / 2 : p.W;
  const long Q = static_cast<long>(p.B) * Ho * Wo;
  const float inv_count = 1.f / static_cast<float>(static_cast<long>(p.B) * p.H * p.W);

  Vec8 scale, shift, c0, c1;
  {
    const Vec8 mean = load_f32x8(p.mean + c0i), invstd = load_f32x8(p.invstd + c0i);
    const Vec8 g = load_f32x8(p.gamma + c0i), b = load_f32x8(p.beta + c0i);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      scale.v[j] = g.v[j] * invstd.v[j];
      shift.v[j] = b.v[j] - mean.v[j] * scale.v[j];
    }
    if (APPLY) {
      // L2 loads: in the fused kernel these sums were produced by atomics earlier in the same launch
      const Vec8 s1 = load_cg_f32x8(p.sums + c0i), s2 = load_cg_f32x8(p.sums + p.C + c0i);   // L2: atomics wrote them
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        c1.v[j] = -scale.v[j] * invstd.v[j] * s2.v[j] * inv_count;
        c0.v[j] = -scale.v[j] * s1.v[j] * inv_count - c1.v[j] * mean.v[j];
      }
      if (blockIdx.x == 0 && rl == 0 && p.dgamma != nullptr) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          p.dgamma[c0i + j] = s2.v[j];
          p.dbeta[c0i + j] = s1.v[j];
        }
      }
    }
  }
  Vec8 a1, a2;
#pragma unroll
  for (int j = 0; j < 8; ++j) a1.v[j] = a2.v[j] = 0.f;

  if (rl < rows_per_block) {
    for (long q = static_cast<long>(blockIdx.x) * rows_per_block + rl; q < Q;
         q += static_cast<long>(gridDim.x) * rows_per_block) {
      const Vec8 go = load8(pdout + q * p.C + c0i);
      if (p.pool) {
        const int wo = static_cast<int>(q % Wo);
        const int ho = static_cast<int>((q / Wo) % Ho);
        const long n = q / (static_cast<long>(Wo) * Ho);
        const long off = ((n * p.H + 2 * ho) * p.W + 2 * wo) * p.C + c0i;
        const long rowb = static_cast<long>(p.W) * p.C;
        const Vec8 v0 = load8(py + off), v1 = load8(py + off + p.C);
        const Vec8 v2 = load8(py + off + rowb), v3 = load8(py + off + rowb + p.C);
        Vec8 o0, o1, o2, o3;
        {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float ya = v0.v[j], yb = v1.v[j], yc = v2.v[j], yd = v3.v[j];
            // same arithmetic as the forward pass so that the winner is identical
            const float za = fmaxf(fmaf(ya, scale.v[j], shift.v[j]), 0.f);
            const float zb = fmaxf(fmaf(yb, scale.v[j], shift.v[j]), 0.f);
            const float zc = fmaxf(fmaf(yc, scale.v[j], shift.v[j]), 0.f);
            const float zd = fmaxf(fmaf(yd, scale.v[j], shift.v[j]), 0.f);
            int win = 0;
            float zmax = za, yw = ya;
            if (zb > zmax) { zmax = zb; win = 1; yw = yb; }
            if (zc > zmax) { zmax = zc; win = 2; yw = yc; }
            if (zd > zmax) { zmax = zd; win = 3; yw = yd; }
            const float dz = zmax > 0.f ? go.v[j] : 0.f;
            if (APPLY) {
              const float sdz = scale.v[j] * dz;
              o0.v[j] = fmaf(c1.v[j], ya, c0.v[j]) + (win == 0 ? sdz : 0.f);
              o1.v[j] = fmaf(c1.v[j], yb, c0.v[j]) + (win == 1 ? sdz : 0.f);
              o2.v[j] = fmaf(c1.v[j], yc, c0.v[j]) + (win == 2 ? sdz : 0.f);
              o3.v[j] = fmaf(c1.v[j], yd, c0.v[j]) + (win == 3 ? sdz : 0.f);
            } else {
              a1.v[j] += dz;
              a2.v[j] = fmaf(dz, yw, a2.v[j]);
            }
          }
        }
        if (APPLY) {
          store8(pdy + off, o0);
          store8(pdy + off + p.C, o1);
          store8(pdy + off + rowb, o2);
          store8(pdy + off + rowb + p.C, o3);
        }
      } else {
        const Vec8 yv = load8(py + q * p.C + c0i);
        Vec8 o;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float z = fmaf(yv.v[j], scale.v[j], shift.v[j]);
          const float dz = z > 0.f ? go.v[j] : 0.f;
          if (APPLY) {
            o.v[j] = fmaf(scale.v[j], dz, fmaf(c1.v[j], yv.v[j], c0.v[j]));
          } else {
            a1.v[j] += dz;
            a2.v[j] = fmaf(dz, yv.v[j], a2.v[j]);
          }
        }
        if (APPLY) store8(pdy + q * p.C + c0i, o);
      }
    }
  }
  if (!APPLY) {
    float* r1 = red;
    float* r2 = red + rows_per_block * p.C;
    if (rl < rows_per_block) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        r1[rl * p.C + c0i + j] = a1.v[j];
        r2[rl * p.C + c0i + j] = a2.v[j];
      }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < p.C; c += blockDim.x) {
      float s = 0.f, t = 0.f;
      for (int r = 0; r < rows_per_block; ++r) {
        s += r1[r * p.C + c];
        t += r2[r * p.C + c];
      }
      // S2 = sum dz*xhat = invstd * (sum dz*y - mean * sum dz)
      const float m = p.mean[c], is = p.invstd[c];
      atomicAdd(p.sums + c, s);
      atomicAdd(p.sums + p.C + c, is * (t - m * s));
    }
  }
}

// Apply pass for blocks WITHOUT a pool: purely element-wise, so four rows per thread per iteration
// (eight 16-byte loads in flight) instead of one -- the generic body keeps too few bytes in flight for
// these layers (ncu: 42 us for 50 MB at 256 channels x 8x8).
template <class T>
__global__ void __launch_bounds__(256, 3) bn_bwd_apply_nopool_kernel(const BnBwdParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const T* __restrict__ py = static_cast<const T*>(p.y);
  const T* __restrict__ pdout = static_cast<const T*>(p.dout);
  T* __restrict__ pdy = static_cast<T*>(p.dy);
  const int chunks = p.C / 8;
  const int rows_per_block = blockDim.x / chunks;
  const int chunk = threadIdx.x % chunks, rl = threadIdx.x / chunks, c0i = chunk * 8;
  const long Q = static_cast<long>(p.B) * p.H * p.W;
  const float inv_count = 1.f / static_cast<float>(Q);
  Vec8 scale, shift, c0, c1;
  {
    const Vec8 mean = load_f32x8(p.mean + c0i), invstd = load_f32x8(p.invstd + c0i);
    const Vec8 g = load_f32x8(p.gamma + c0i), b = load_f32x8(p.beta + c0i);
    const Vec8 s1 = load_cg_f32x8(p.sums + c0i), s2 = load_cg_f32x8(p.sums + p.C + c0i);   // L2: atomics wrote them
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      scale.v[j] = g.v[j] * invstd.v[j];
      shift.v[j] = b.v[j] - mean.v[j] * scale.v[j];
      c1.v[j] = -scale.v[j] * invstd.v[j] * s2.v[j] * inv_count;
      c0.v[j] = -scale.v[j] * s1.v[j] * inv_count - c1.v[j] * mean.v[j];
    }
    if (blockIdx.x == 0 && rl == 0 && p.dgamma != nullptr) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        p.dgamma[c0i + j] = s2.v[j];
        p.dbeta[c0i + j] = s1.v[j];
      }
    }
  }
  if (rl >= rows_per_block) return;
  constexpr int U = 4;
  const long stride = static_cast<long>(gridDim.x) * rows_per_block;
  for (long q0 = static_cast<long>(blockIdx.x) * rows_per_block + rl; q0 < Q; q0 += U * stride) {
    Raw8<T> ry[U], rg[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long q = q0 + u * stride;
      const long qq = q < Q ? q : q0;
      ry[u] = ldraw8(py + qq * p.C + c0i);
      rg[u] = ldraw8(pdout + qq * p.C + c0i);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long q = q0 + u * stride;
      if (q >= Q) break;
      const Vec8 yv = unpack8(ry[u]), go = unpack8(rg[u]);
      Vec8 o;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float z = fmaf(yv.v[j], scale.v[j], shift.v[j]);
        const float dz = z > 0.f ? go.v[j] : 0.f;
        o.v[j] = fmaf(scale.v[j], dz, fmaf(c1.v[j], yv.v[j], c0.v[j]));
      }
      store8(pdy + q * p.C + c0i, o);
    }
  }
}

// Apply pass for POOLED blocks, two pooled pixels per thread per iteration (ten 16-byte loads in
// flight instead of five).  Same arithmetic as the generic body, so winners match the forward pass.
template <class T>
__global__ void __launch_bounds__(256, 2) bn_bwd_apply_pool_kernel(const BnBwdParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const T* __restrict__ py = static_cast<const T*>(p.y);
  const T* __restrict__ pdout = static_cast<const T*>(p.dout);
  T* __restrict__ pdy = static_cast<T*>(p.dy);
  const int chunks = p.C / 8;
  const int rows_per_block = blockDim.x / chunks;
  const int chunk = threadIdx.x % chunks, rl = threadIdx.x / chunks, c0i = chunk * 8;
  const int Ho = p.H / 2, Wo = p.W / 2;
  const long Q = static_cast<long>(p.B) * Ho * Wo;
  const float inv_count = 1.f / static_cast<float>(static_cast<long>(p.B) * p.H * p.W);
  Vec8 scale, shift, c0, c1;
  {
    const Vec8 mean = load_f32x8(p.mean + c0i), invstd = load_f32x8(p.invstd + c0i);
    const Vec8 g = load_f32x8(p.gamma + c0i), b = load_f32x8(p.beta + c0i);
    const Vec8 s1 = load_cg_f32x8(p.sums + c0i), s2 = load_cg_f32x8(p.sums + p.C + c0i);   // L2: atomics wrote them
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      scale.v[j] = g.v[j] * invstd.v[j];
      shift.v[j] = b.v[j] - mean.v[j] * scale.v[j];
      c1.v[j] = -scale.v[j] * invstd.v[j] * s2.v[j] * inv_count;
      c0.v[j] = -scale.v[j] * s1.v[j] * inv_count - c1.v[j] * mean.v[j];
    }
    if (blockIdx.x == 0 && rl == 0 && p.dgamma != nullptr) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        p.dgamma[c0i + j] = s2.v[j];
        p.dbeta[c0i + j] = s1.v[j];
      }
    }
  }
  if (rl >= rows_per_block) return;
  constexpr int U = 2;
  const long stride = static_cast<long>(gridDim.x) * rows_per_block;
  const long rowb = static_cast<long>(p.W) * p.C;
  for (long q0 = static_cast<long>(blockIdx.x) * rows_per_block + rl; q0 < Q; q0 += U * stride) {
    long off[U];
    Raw8<T> raw[U][4], rg[U];
    bool live[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long q = q0 + u * stride;
      live[u] = q < Q;
      const long qq = live[u] ? q : q0;
      const int wo = static_cast<int>(qq % Wo);
      const int ho = static_cast<int>((qq / Wo) % Ho);
      const long n = qq / (static_cast<long>(Wo) * Ho);
      off[u] = ((n * p.H + 2 * ho) * p.W + 2 * wo) * p.C + c0i;
      rg[u] = ldraw8(pdout + qq * p.C + c0i);
      raw[u][0] = ldraw8(py + off[u]);
      raw[u][1] = ldraw8(py + off[u] + p.C);
      raw[u][2] = ldraw8(py + off[u] + rowb);
      raw[u][3] = ldraw8(py + off[u] + rowb + p.C);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (!live[u]) break;
      const Vec8 go = unpack8(rg[u]);
      const Vec8 ya = unpack8(raw[u][0]), yb = unpack8(raw[u][1]), yc = unpack8(raw[u][2]), yd = unpack8(raw[u][3]);
      Vec8 o0, o1, o2, o3;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float za = fmaxf(fmaf(ya.v[j], scale.v[j], shift.v[j]), 0.f);
        const float zb = fmaxf(fmaf(yb.v[j], scale.v[j], shift.v[j]), 0.f);
        const float zc = fmaxf(fmaf(yc.v[j], scale.v[j], shift.v[j]), 0.f);
        const float zd = fmaxf(fmaf(yd.v[j], scale.v[j], shift.v[j]), 0.f);
        int win = 0;
        float zmax = za;
        if (zb > zmax) { zmax = zb; win = 1; }
        if (zc > zmax) { zmax = zc; win = 2; }
        if (zd > zmax) { zmax = zd; win = 3; }
        const float sdz = zmax > 0.f ? scale.v[j] * go.v[j] : 0.f;
        o0.v[j] = fmaf(c1.v[j], ya.v[j], c0.v[j]) + (win == 0 ? sdz : 0.f);
        o1.v[j] = fmaf(c1.v[j], yb.v[j], c0.v[j]) + (win == 1 ? sdz : 0.f);
        o2.v[j] = fmaf(c1.v[j], yc.v[j], c0.v[j]) + (win == 2 ? sdz : 0.f);
        o3.v[j] = fmaf(c1.v[j], yd.v[j], c0.v[j]) + (win == 3 ? sdz : 0.f);
      }
      store8(pdy + off[u], o0);
      store8(pdy + off[u] + p.C, o1);
      store8(pdy + off[u] + rowb, o2);
      store8(pdy + off[u] + rowb + p.C, o3);
    }
  }
}

template <class T, bool APPLY>
__global__ void __launch_bounds__(256, 3) bn_relu_pool_bwd_kernel(const BnBwdParams p) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ float red[];
  bn_relu_pool_bwd_body<T, APPLY>(p, red);
}

// ---------------------------------------------------------------------------
// Bulk-copy pipelined variant of the pooled apply pass (what b200dp_bn_relu_pool_bwd_apply runs with pool = 1
// whenever the shape fits).
// The register-file versions above keep at most ~10 KiB in flight per SM (37 % occupancy, a few 16-byte
// loads per thread) and measure 45-56 % of the HBM / L2 rate.  Here the memory-level parallelism no longer
// depends on the compute threads: one thread streams 20 KiB stages (256 work items: the two image rows of
// y that a row of pooled pixels needs are contiguous in NHWC, and so is the matching row of dout) into a
// ring of shared-memory buffers with cp.async.bulk, the 256 threads compute from shared memory into one of
// two output buffers, and one thread streams those back with bulk stores.  Arithmetic per element is the
// same as in bn_bwd_apply_pool_kernel, so the results are bit-identical.
// Work item = (pooled pixel, 8-channel chunk); a "tile" = one pooled row of one image = Wo * C/8 items;
// a stage = kBnTmaItems / (Wo * C/8) consecutive tiles (all tiles are contiguous in memory, image
// boundaries included).  Requires (Wo * C/8) to divide kBnTmaItems.
// ---------------------------------------------------------------------------
// Work items per stage = threads per block: 256 for bf16, 128 for fp32 activations (the same bytes per stage).
// Three stages, two CTAs per SM (184 KB).  Two stages with three CTAs per SM (24 warps instead of 16) measured
// slower (0.597 vs 0.578 ms per step).
constexpr int kBnTmaStages = 3;
constexpr int kBnTmaYBytes = 256 * 4 * 16;              // 16 KiB of y per stage (4 vectors of 8 channels per item)
constexpr int kBnTmaGBytes = 256 * 16;                  // 4 KiB of dout per stage
constexpr int kBnTmaStageBytes = kBnTmaYBytes + kBnTmaGBytes;
constexpr int kBnTmaSmem = kBnTmaStages * kBnTmaStageBytes + 2 * kBnTmaYBytes + 64 + 128;
template <class T>
constexpr int bn_tma_items() { return 512 / static_cast<int>(sizeof(T)); }

template <class T>
__global__ void __launch_bounds__(bn_tma_items<T>(), 2) bn_bwd_apply_pool_tma_kernel(const BnBwdParams p) {
  constexpr int kBnTmaItems = bn_tma_items<T>();
  const T* __restrict__ py = static_cast<const T*>(p.y);
  const T* __restrict__ pdout = static_cast<const T*>(p.dout);
  T* __restrict__ pdy = static_cast<T*>(p.dy);
  extern __shared__ uint8_t bn_tma_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(bn_tma_raw) + 127) & ~uintptr_t(127));
  uint8_t* out_buf = smem + kBnTmaStages * kBnTmaStageBytes;                     // 2 x 16 KiB
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(out_buf + 2 * kBnTmaYBytes);  // [kBnTmaStages]
  pdl_launch_dependents();
  const int chunks = p.C / 8;
  const int Ho = p.H / 2, Wo = p.W / 2;
  const int items_per_tile = Wo * chunks;
  const int G = kBnTmaItems / items_per_tile;                       // tiles per stage
  const long NT = static_cast<long>(p.B) * Ho;                      // tiles in the tensor
  const long num_chunks = (NT + G - 1) / G;
  const long tile_y_elems = 2L * p.W * p.C, tile_g_elems = static_cast<long>(Wo) * p.C;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kBnTmaStages; ++s) mbar_init(&full_bar[s], 1);
    mbar_fence_init();
  }
  __syncthreads();
  pdl_wait();

  // producer side (thread 0): chunk j of this block is global chunk blockIdx.x + j * gridDim.x
  auto issue = [&](long j, int stage) {
    const long chunk_id = blockIdx.x + j * static_cast<long>(gridDim.x);
    if (chunk_id >= num_chunks) return;
    const long t0 = chunk_id * G;
    const int tiles = static_cast<int>(min(static_cast<long>(G), NT - t0));
    uint8_t* dst = smem + stage * kBnTmaStageBytes;
    const uint32_t yb = static_cast<uint32_t>(tiles * tile_y_elems * sizeof(T)), gb = static_cast<uint32_t>(tiles * tile_g_elems * sizeof(T));
    mbar_arrive_expect_tx(&full_bar[stage], yb + gb);
    bulk_load_1d(dst, py + t0 * tile_y_elems, yb, &full_bar[stage]);
    bulk_load_1d(dst + kBnTmaYBytes, pdout + t0 * tile_g_elems, gb, &full_bar[stage]);
  };
  if (threadIdx.x == 0)
    for (int s = 0; s < kBnTmaStages; ++s) issue(s, s);

  // per-thread constants: the channel chunk of a thread never changes (kBnTmaItems % chunks == 0)
  const int chunk = threadIdx.x % chunks, pix = threadIdx.x / chunks, c0i = chunk * 8;
  const int gt = pix / Wo, wo = pix - gt * Wo;                        // tile inside the stage, pooled column
  const float inv_count = 1.f / static_cast<float>(static_cast<long>(p.B) * p.H * p.W);
  Vec8 scale, shift, c0, c1;
  {
    const Vec8 mean = load_f32x8(p.mean + c0i), invstd = load_f32x8(p.invstd + c0i);
    const Vec8 g = load_f32x8(p.gamma + c0i), b = load_f32x8(p.beta + c0i);
    const Vec8 s1 = load_cg_f32x8(p.sums + c0i), s2 = load_cg_f32x8(p.sums + p.C + c0i);   // L2: atomics wrote them
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      scale.v[j] = g.v[j] * invstd.v[j];
      shift.v[j] = b.v[j] - mean.v[j] * scale.v[j];
      c1.v[j] = -scale.v[j] * invstd.v[j] * s2.v[j] * inv_count;
      c0.v[j] = -scale.v[j] * s1.v[j] * inv_count - c1.v[j] * mean.v[j];
    }
    if (blockIdx.x == 0 && pix == 0 && p.dgamma != nullptr) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        p.dgamma[c0i + j] = s2.v[j];
        p.dbeta[c0i + j] = s1.v[j];
      }
    }
  }
  // element offsets of this thread's item inside a stage (same layout in the y, dy and dout buffers)
  const int y_off = gt * static_cast<int>(tile_y_elems) + (2 * wo) * p.C + c0i;
  const int g_off = gt * static_cast<int>(tile_g_elems) + wo * p.C + c0i;
  const int rowb = p.W * p.C;

  int stage = 0;
  uint32_t phase = 0;
  for (long j = 0;; ++j) {
    const long chunk_id = blockIdx.x + j * static_cast<long>(gridDim.x);
    if (chunk_id >= num_chunks) break;
    const long t0 = chunk_id * G;
    const int tiles = static_cast<int>(min(static_cast<long>(G), NT - t0));
    uint8_t* ob = out_buf + (j & 1) * kBnTmaYBytes;
    if (threadIdx.x == 0) bulk_wait_group_read<1>();   // the store issued two chunks ago has drained this buffer
    mbar_wait(&full_bar[stage], phase);                // this stage's y and dout have landed
    __syncthreads();                                   // ... and everybody knows the output buffer is free
    if (gt < tiles) {
      const T* ys = reinterpret_cast<const T*>(smem + stage * kBnTmaStageBytes);
      const T* gs = reinterpret_cast<const T*>(smem + stage * kBnTmaStageBytes + kBnTmaYBytes);
      T* os = reinterpret_cast<T*>(ob);
      const Vec8 go = load8(gs + g_off);
      const Vec8 ya = load8(ys + y_off);
      const Vec8 yb = load8(ys + y_off + p.C);
      const Vec8 yc = load8(ys + y_off + rowb);
      const Vec8 yd = load8(ys + y_off + rowb + p.C);
      Vec8 o0, o1, o2, o3;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float za = fmaxf(fmaf(ya.v[k], scale.v[k], shift.v[k]), 0.f);
        const float zb = fmaxf(fmaf(yb.v[k], scale.v[k], shift.v[k]), 0.f);
        const float zc = fmaxf(fmaf(yc.v[k], scale.v[k], shift.v[k]), 0.f);
        const float zd = fmaxf(fmaf(yd.v[k], scale.v[k], shift.v[k]), 0.f);
        int win = 0;
        float zmax = za;
        if (zb > zmax) { zmax = zb; win = 1; }
        if (zc > zmax) { zmax = zc; win = 2; }
        if (zd > zmax) { zmax = zd; win = 3; }
        const float sdz = zmax > 0.f ? scale.v[k] * go.v[k] : 0.f;
        o0.v[k] = fmaf(c1.v[k], ya.v[k], c0.v[k]) + (win == 0 ? sdz : 0.f);
        o1.v[k] = fmaf(c1.v[k], yb.v[k], c0.v[k]) + (win == 1 ? sdz : 0.f);
        o2.v[k] = fmaf(c1.v[k], yc.v[k], c0.v[k]) + (win == 2 ? sdz : 0.f);
        o3.v[k] = fmaf(c1.v[k], yd.v[k], c0.v[k]) + (win == 3 ? sdz : 0.f);
      }
      store8(os + y_off, o0);
      store8(os + y_off + p.C, o1);
      store8(os + y_off + rowb, o2);
      store8(os + y_off + rowb + p.C, o3);
    }
    fence_proxy_async_smem();   // generic-proxy writes above must be visible to the bulk store below
    __syncthreads();            // everyone has written its outputs and finished reading the input stage
    if (threadIdx.x == 0) {
      bulk_store_1d(pdy + t0 * tile_y_elems, ob, static_cast<uint32_t>(tiles * tile_y_elems * sizeof(T)));
      bulk_commit_group();
      issue(j + kBnTmaStages, stage);   // refill the stage just consumed
    }
    if (++stage == kBnTmaStages) {
      stage = 0;
      phase ^= 1;
    }
  }
  if (threadIdx.x == 0) bulk_wait_group<0>();   // all stores performed before the block retires
}

// Reduction pass from the block OUTPUT instead of the conv output.  With a = relu(z) (max-pooled or
// not) and g the gradient arriving at a: dz is g where a > 0 and lands on the element whose z equals
// a, so  S1 = sum_{a>0} g  and  S2 = sum dz*xhat = sum_{a>0} g * (a - beta) / gamma.  For pooled
// layers this reads two quarter-size tensors instead of the full-resolution y (2.5x fewer bytes) and
// needs no arg-max search; `a` is the stored bf16 value, i.e. z rounded to 8 bits of mantissa, whose
// error averages out over the B*H*W terms of the sums.
template <class T>
__global__ void __launch_bounds__(256, 4) bn_bwd_reduce_from_out_kernel(const T* __restrict__ out,
                                                                        const T* __restrict__ dout,
                                                                        const float* __restrict__ gamma,
                                                                        const float* __restrict__ beta,
                                                                        float* __restrict__ sums, long Q, int C) {
  extern __shared__ float red[];
  pdl_launch_dependents();
  pdl_wait();
  const int chunks = C / 8;
  const int rows_per_block = blockDim.x / chunks;
  const int chunk = threadIdx.x % chunks, rl = threadIdx.x / chunks, c0 = chunk * 8;
  Vec8 bt, ig;
  {
    const Vec8 g = load_f32x8(gamma + c0);
    bt = load_f32x8(beta + c0);
#pragma unroll
    for (int j = 0; j < 8; ++j) ig.v[j] = fabsf(g.v[j]) > 1e-20f ? 1.f / g.v[j] : 0.f;
  }
  Vec8 a1, a2;
#pragma unroll
  for (int j = 0; j < 8; ++j) a1.v[j] = a2.v[j] = 0.f;
  if (rl < rows_per_block) {
    const long stride = static_cast<long>(gridDim.x) * rows_per_block;
    for (long qa = static_cast<long>(blockIdx.x) * rows_per_block + rl; qa < Q; qa += 2 * stride) {
      const long qb = qa + stride;
      const bool hasb = qb < Q;
      const Raw8<T> oa = ldraw8(out + qa * C + c0);
      const Raw8<T> ga = ldraw8(dout + qa * C + c0);
      const Raw8<T> ob = ldraw8(out + (hasb ? qb : qa) * C + c0);
      const Raw8<T> gb = ldraw8(dout + (hasb ? qb : qa) * C + c0);
      const Vec8 o1 = unpack8(oa), g1 = unpack8(ga), o2 = unpack8(ob), g2 = unpack8(gb);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float d1 = o1.v[j] > 0.f ? g1.v[j] : 0.f;
        const float d2 = (hasb && o2.v[j] > 0.f) ? g2.v[j] : 0.f;
        a1.v[j] += d1 + d2;
        a2.v[j] = fmaf(d1, (o1.v[j] - bt.v[j]) * ig.v[j], a2.v[j]);
        a2.v[j] = fmaf(d2, (o2.v[j] - bt.v[j]) * ig.v[j], a2.v[j]);
      }
    }
  }
  float* r1 = red;
  float* r2 = red + rows_per_block * C;
  if (rl < rows_per_block) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      r1[rl * C + c0 + j] = a1.v[j];
      r2[rl * C + c0 + j] = a2.v[j];
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float sa = 0.f, sb = 0.f;
    for (int r = 0; r < rows_per_block; ++r) {
      sa += r1[r * C + c];
      sb += r2[r * C + c];
    }
    atomicAdd(sums + c, sa);
    atomicAdd(sums + C + c, sb);
  }
}

// ---------------------------------------------------------------------------
// classifier: logits = x W^T + b ; softmax cross-entropy (mean) ; backward
// ---------------------------------------------------------------------------
constexpr int kLinMaxO = 16;
constexpr int kLinTileB = 256;

// one warp per sample: the 16 activations of a lane are loaded once, all O dot products run as
// independent chains, then O warp reductions
template <class T>
__global__ void __launch_bounds__(128) linear_fwd_kernel(const T* __restrict__ x,
                                                         const float* __restrict__ w, const float* __restrict__ bias,
                                                         float* __restrict__ logits, int B, int F, int O) {
  pdl_launch_dependents();
  pdl_wait();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= B) return;
  const T* xr = x + static_cast<long>(warp) * F;
  float acc[kLinMaxO];
#pragma unroll
  for (int o = 0; o < kLinMaxO; ++o) acc[o] = 0.f;
  for (int f = lane * 8; f < F; f += 256) {
    const Vec8 xv = load8(xr + f);
#pragma unroll
    for (int o = 0; o < kLinMaxO; ++o) {
      if (o < O) {
        const Vec8 wv = load_f32x8(w + static_cast<long>(o) * F + f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[o] = fmaf(xv.v[j], wv.v[j], acc[o]);
      }
    }
  }
#pragma unroll
  for (int o = 0; o < kLinMaxO; ++o) {
    if (o < O) {
      const float t = warp_sum(acc[o]);
      if (lane == 0) logits[static_cast<long>(warp) * O + o] = t + bias[o];
    }
  }
}

// one block; loss is the batch mean; dlogits = (softmax - onehot) * grad_scale / B
__global__ void cross_entropy_kernel(const float* __restrict__ logits, const long long* __restrict__ target,
                                     float* __restrict__ loss, float* __restrict__ dlogits, int* __restrict__ correct,
                                     int B, int O, float grad_scale) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float s_loss[32];
  __shared__ int s_corr[32];
  float my_loss = 0.f;
  int my_corr = 0;
  for (int i = threadIdx.x; i < B; i += blockDim.x) {
    const float* l = logits + static_cast<long>(i) * O;
    float mx = l[0];
    int am = 0;
    for (int o = 1; o < O; ++o) {
      if (l[o] > mx) {
        mx = l[o];
        am = o;
      }
    }
    float se = 0.f;
    for (int o = 0; o < O; ++o) se += expf(l[o] - mx);
    const float lse = logf(se) + mx;
    const int t = static_cast<int>(target[i]);
    my_loss += lse - l[t];
    my_corr += (am == t);
    if (dlogits != nullptr) {
      const float sc = grad_scale / static_cast<float>(B);
      for (int o = 0; o < O; ++o)
        dlogits[static_cast<long>(i) * O + o] = (expf(l[o] - lse) - (o == t ? 1.f : 0.f)) * sc;
    }
  }
  my_loss = warp_sum(my_loss);
  for (int o = 16; o > 0; o >>= 1) my_corr += __shfl_xor_sync(0xffffffffu, my_corr, o);
  if ((threadIdx.x & 31) == 0) {
    s_loss[threadIdx.x >> 5] = my_loss;
    s_corr[threadIdx.x >> 5] = my_corr;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float tl = 0.f;
    int tc = 0;
    for (int w = 0; w < (blockDim.x + 31) / 32; ++w) {
      tl += s_loss[w];
      tc += s_corr[w];
    }
    if (loss != nullptr) *loss = tl / static_cast<float>(B);
    if (correct != nullptr) *correct = tc;
  }
}

// blocks [0, nW): dW (and db) for a tile of 64 features; blocks [nW, ...): dX.
// dW[o][f] = sum_b dl[b][o] * x[b][f]: 64 feature lanes x 4 batch groups per block, the dl tile
// staged in shared memory, O accumulators per thread, batch groups combined through smem.
template <class T>
__global__ void __launch_bounds__(256) linear_bwd_kernel(const T* __restrict__ x,
                                                         const float* __restrict__ w,
                                                         const float* __restrict__ dlogits, float* __restrict__ dw,
                                                         float* __restrict__ db, T* __restrict__ dx,
                                                         int B, int F, int O, int nW) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float s_dl[kLinTileB * kLinMaxO];
  __shared__ float s_red[4][kLinMaxO][64];
  if (static_cast<int>(blockIdx.x) < nW) {
    // nW = feature tiles x batch slices; partial sums are combined with atomics (dw, db zero on entry)
    const int ftiles = (F + 63) / 64;
    const int ft = blockIdx.x % ftiles, slice = blockIdx.x / ftiles, nslices = nW / ftiles;
    const int fl = threadIdx.x & 63, bg = threadIdx.x >> 6;
    const int f = ft * 64 + fl;
    const int per = (B + nslices - 1) / nslices;
    const int b_lo = slice * per, b_hi = min(B, b_lo + per);
    float acc[kLinMaxO];
#pragma unroll
    for (int o = 0; o < kLinMaxO; ++o) acc[o] = 0.f;
    const int nb = max(b_hi - b_lo, 0);
    for (int i = threadIdx.x; i < nb * O; i += blockDim.x) s_dl[i] = dlogits[static_cast<long>(b_lo) * O + i];
    __syncthreads();
    if (f < F) {
#pragma unroll 4
      for (int b = bg; b < nb; b += 4) {
        const float xv = static_cast<float>(x[static_cast<long>(b_lo + b) * F + f]);
#pragma unroll
        for (int o = 0; o < kLinMaxO; ++o)
          if (o < O) acc[o] = fmaf(s_dl[b * O + o], xv, acc[o]);
      }
    }
#pragma unroll
    for (int o = 0; o < kLinMaxO; ++o)
      if (o < O) s_red[bg][o][fl] = acc[o];
    __syncthreads();
    if (bg == 0 && f < F) {
#pragma unroll
      for (int o = 0; o < kLinMaxO; ++o)
        if (o < O)
          atomicAdd(dw + static_cast<long>(o) * F + f,
                    s_red[0][o][fl] + s_red[1][o][fl] + s_red[2][o][fl] + s_red[3][o][fl]);
    }
    if (ft == 0 && threadIdx.x < O) {
      float dbacc = 0.f;
      for (int b = 0; b < nb; ++b) dbacc += s_dl[b * O + threadIdx.x];
      atomicAdd(db + threadIdx.x, dbacc);
    }
  } else {
    const long idx = static_cast<long>(blockIdx.x - nW) * blockDim.x + threadIdx.x;
    if (idx < static_cast<long>(B) * F) {
      const int i = static_cast<int>(idx / F), f = static_cast<int>(idx - static_cast<long>(i) * F);
      float acc = 0.f;
      for (int o = 0; o < O; ++o) acc = fmaf(dlogits[static_cast<long>(i) * O + o], w[static_cast<long>(o) * F + f], acc);
      dx[idx] = operand_cast<T>(acc);
    }
  }
}

// ---------------------------------------------------------------------------
// classifier head of a training step in ONE launch: logits, mean softmax cross-entropy, top-1 count,
// dlogits, dX and this block's share of dW / db.  kHeadWarps samples per block, one warp each; the weight
// matrix is staged in shared memory once per block (all its loads in flight together, no per-warp L2 round
// trips), the block's activations and dlogits are parked there too so that dW needs one vector reduction per
// (class, 4 features) per block.  Loss and count are combined in block order by the last block to take a ticket,
// so neither needs clearing and the result does not depend on the order of atomics; the ticket is taken before
// the dW reductions are issued so that its round trip overlaps them.
// ---------------------------------------------------------------------------
constexpr int kHeadWarps = 4;
constexpr int kHeadMaxBlocks = 1024;
__device__ float g_head_loss_part[kHeadMaxBlocks];
__device__ int g_head_corr_part[kHeadMaxBlocks];
__device__ unsigned int g_head_blocks_done = 0;

template <class T>
B200DP_DEVINL float4 head_load4(const T* p) {
  if constexpr (sizeof(T) == 4) {
    return *reinterpret_cast<const float4*>(p);
  } else {
    const uint2 r = *reinterpret_cast<const uint2*>(p);
    return make_float4(__uint_as_float(r.x << 16), __uint_as_float(r.x & 0xffff0000u), __uint_as_float(r.y << 16),
                       __uint_as_float(r.y & 0xffff0000u));
  }
}

B200DP_DEVINL void head_red_add_v4(float* addr, const float4& v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

template <class T>
__global__ void __launch_bounds__(kHeadWarps * 32)
    head_fwd_bwd_kernel(const T* __restrict__ x, const float* __restrict__ w,
                        const float* __restrict__ bias, const long long* __restrict__ target,
                        float* __restrict__ logits, float* __restrict__ loss, int* __restrict__ correct,
                        float* __restrict__ dlogits, T* __restrict__ dx, float* __restrict__ dw,
                        float* __restrict__ db, int B, int F, int O, float grad_scale) {
  pdl_launch_dependents();
  extern __shared__ __align__(16) float head_sm[];
  float* ws = head_sm;                       // [O][F]
  float* xs = ws + O * F;                    // [kHeadWarps][F]
  float* dls = xs + kHeadWarps * F;          // [kHeadWarps][kLinMaxO]
  __shared__ float s_loss[kHeadWarps];
  __shared__ int s_corr[kHeadWarps];
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int i = blockIdx.x * kHeadWarps + wib;  // sample of this warp
  const bool live = i < B;
  pdl_wait();
  // weights -> shared memory (coalesced float4), this warp's activation row -> shared memory as fp32
  for (int k = threadIdx.x * 4; k < O * F; k += kHeadWarps * 32 * 4)
    *reinterpret_cast<float4*>(ws + k) = __ldg(reinterpret_cast<const float4*>(w + k));
  for (int f = lane * 4; f < F; f += 128) {
    const float4 v = live ? head_load4(x + static_cast<long>(i) * F + f) : make_float4(0.f, 0.f, 0.f, 0.f);
    *reinterpret_cast<float4*>(xs + wib * F + f) = v;
  }
  if (lane < kLinMaxO) dls[wib * kLinMaxO + lane] = 0.f;
  __syncthreads();
  float my_loss = 0.f;
  int my_corr = 0;
  if (live) {
    float acc[kLinMaxO];
#pragma unroll
    for (int o = 0; o < kLinMaxO; ++o) acc[o] = 0.f;
    for (int f = lane * 4; f < F; f += 128) {
      const float4 xv = *reinterpret_cast<const float4*>(xs + wib * F + f);
#pragma unroll
      for (int o = 0; o < kLinMaxO; ++o) {
        if (o < O) {
          const float4 wv = *reinterpret_cast<const float4*>(ws + o * F + f);
          acc[o] = fmaf(xv.x, wv.x, fmaf(xv.y, wv.y, fmaf(xv.z, wv.z, fmaf(xv.w, wv.w, acc[o]))));
        }
      }
    }
    // every lane ends up with every logit (butterfly), so the softmax below needs no broadcast
    float l[kLinMaxO];
#pragma unroll
    for (int o = 0; o < kLinMaxO; ++o) {
      float t = acc[o];
      if (o < O) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) t += __shfl_xor_sync(0xffffffffu, t, d);
      }
      l[o] = o < O ? t + __ldg(bias + o) : -INFINITY;
    }
    float mx = l[0];
    int am = 0;
#pragma unroll
    for (int o = 1; o < kLinMaxO; ++o) {
      if (o < O && l[o] > mx) {
        mx = l[o];
        am = o;
      }
    }
    float se = 0.f;
#pragma unroll
    for (int o = 0; o < kLinMaxO; ++o)
      if (o < O) se += expf(l[o] - mx);
    const float lse = logf(se) + mx;
    const int t = static_cast<int>(target[i]);
    const float sc = grad_scale / static_cast<float>(B);
    float dl[kLinMaxO], lt = 0.f, mine_l = 0.f, mine_d = 0.f;
#pragma unroll
    for (int o = 0; o < kLinMaxO; ++o) {
      dl[o] = o < O ? (expf(l[o] - lse) - (o == t ? 1.f : 0.f)) * sc : 0.f;
      lt = (o == t) ? l[o] : lt;
      mine_l = (o == lane) ? l[o] : mine_l;
      mine_d = (o == lane) ? dl[o] : mine_d;
    }
    my_loss = lse - lt;
    my_corr = (am == t) ? 1 : 0;
    if (lane < O) {
      logits[static_cast<long>(i) * O + lane] = mine_l;
      if (dlogits != nullptr) dlogits[static_cast<long>(i) * O + lane] = mine_d;
      dls[wib * kLinMaxO + lane] = mine_d;
    }
    for (int f = lane * 4; f < F; f += 128) {
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int o = 0; o < kLinMaxO; ++o) {
        if (o < O) {
          const float4 wv = *reinterpret_cast<const float4*>(ws + o * F + f);
          a.x = fmaf(dl[o], wv.x, a.x);
          a.y = fmaf(dl[o], wv.y, a.y);
          a.z = fmaf(dl[o], wv.z, a.z);
          a.w = fmaf(dl[o], wv.w, a.w);
        }
      }
      operand_store4(dx + static_cast<long>(i) * F + f, a.x, a.y, a.z, a.w);
    }
  }
  if (lane == 0) {
    s_loss[wib] = my_loss;
    s_corr[wib] = my_corr;
  }
  __syncthreads();
  // loss / count: per-block partials; warp 0 takes the block's ticket now and looks at it after its reductions
  unsigned ticket = 0;
  if (threadIdx.x == 0) {
    float tl = 0.f;
    int tc = 0;
#pragma unroll
    for (int k = 0; k < kHeadWarps; ++k) {
      tl += s_loss[k];
      tc += s_corr[k];
    }
    g_head_loss_part[blockIdx.x] = tl;
    g_head_corr_part[blockIdx.x] = tc;
    __threadfence();
    ticket = atomicAdd(&g_head_blocks_done, 1u);
  }
  // dW / db of the block's samples: one vector reduction per (class, 4 features) per block
  for (int f = threadIdx.x * 4; f < F; f += kHeadWarps * 32 * 4) {
    float4 xv[kHeadWarps];
#pragma unroll
    for (int sidx = 0; sidx < kHeadWarps; ++sidx) xv[sidx] = *reinterpret_cast<const float4*>(xs + sidx * F + f);
#pragma unroll
    for (int o = 0; o < kLinMaxO; ++o) {
      if (o < O) {
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int sidx = 0; sidx < kHeadWarps; ++sidx) {
          const float d = dls[sidx * kLinMaxO + o];
          a.x = fmaf(d, xv[sidx].x, a.x);
          a.y = fmaf(d, xv[sidx].y, a.y);
          a.z = fmaf(d, xv[sidx].z, a.z);
          a.w = fmaf(d, xv[sidx].w, a.w);
        }
        head_red_add_v4(dw + static_cast<long>(o) * F + f, a);
      }
    }
  }
  if (static_cast<int>(threadIdx.x) < O) {
    float a = 0.f;
#pragma unroll
    for (int sidx = 0; sidx < kHeadWarps; ++sidx) a += dls[sidx * kLinMaxO + threadIdx.x];
    atomicAdd(db + threadIdx.x, a);
  }
  if (wib == 0) {
    ticket = __shfl_sync(0xffffffffu, ticket, 0);
    if (ticket == gridDim.x - 1) {   // every other block has published its partials before taking its ticket
      __threadfence();
      float tl = 0.f;
      int tc = 0;
      for (unsigned k = lane; k < gridDim.x; k += 32) {   // lane-strided, then a fixed butterfly: deterministic
        tl += __ldcg(&g_head_loss_part[k]);
        tc += __ldcg(&g_head_corr_part[k]);
      }
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) {
        tl += __shfl_xor_sync(0xffffffffu, tl, d);
        tc += __shfl_xor_sync(0xffffffffu, tc, d);
      }
      if (lane == 0) {
        if (loss != nullptr) *loss = tl / static_cast<float>(B);
        if (correct != nullptr) *correct = tc;
        g_head_blocks_done = 0;  // ready for the next launch (stream order, graph replays)
      }
    }
  }
}

// ---------------------------------------------------------------------------
// optimizer and weight re-layout
// ---------------------------------------------------------------------------
// d = g*gscale + wd*p ; buf = mom*buf + d ; p -= lr*buf.  With buf zero-initialised the
// first step gives buf = d, which is torch.optim.SGD's first-step rule (dampening 0).
__global__ void sgd_momentum_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ buf,
                                    long n, float lr, float momentum, float wd, float gscale) {
  pdl_launch_dependents();
  pdl_wait();
  const long n4 = n >> 2;
  const long stride = static_cast<long>(gridDim.x) * blockDim.x;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4; i += stride) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    const float4 gv = reinterpret_cast<const float4*>(g)[i];
    float4 bv = reinterpret_cast<float4*>(buf)[i];
    bv.x = fmaf(momentum, bv.x, fmaf(wd, pv.x, gv.x * gscale));
    bv.y = fmaf(momentum, bv.y, fmaf(wd, pv.y, gv.y * gscale));
    bv.z = fmaf(momentum, bv.z, fmaf(wd, pv.z, gv.z * gscale));
    bv.w = fmaf(momentum, bv.w, fmaf(wd, pv.w, gv.w * gscale));
    pv.x = fmaf(-lr, bv.x, pv.x);
    pv.y = fmaf(-lr, bv.y, pv.y);
    pv.z = fmaf(-lr, bv.z, pv.z);
    pv.w = fmaf(-lr, bv.w, pv.w);
    reinterpret_cast<float4*>(buf)[i] = bv;
    reinterpret_cast<float4*>(p)[i] = pv;
  }
  for (long i = (n4 << 2) + blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n; i += stride) {
    const float b = fmaf(momentum, buf[i], fmaf(wd, p[i], g[i] * gscale));
    buf[i] = b;
    p[i] = fmaf(-lr, b, p[i]);
  }
}

// Clear n16 16-byte words (the flat gradient buffer before the split-K weight gradients are accumulated
// into it, the per-step statistics region).  The framework's own fill, so that a training step launches no
// library kernel and the fill takes part in the programmatic-dependent-launch chain.
__global__ void __launch_bounds__(256) fill_zero_kernel(uint4* __restrict__ dst, long n16) {
  pdl_launch_dependents();
  pdl_wait();
  const long stride = static_cast<long>(gridDim.x) * blockDim.x;
  const uint4 z = make_uint4(0u, 0u, 0u, 0u);
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n16; i += stride) dst[i] = z;
}

// Clear two fp32 ranges of any length (the classifier's dW / db before the atomics of linear_bwd).
__global__ void __launch_bounds__(256) fill_zero_f32_pair_kernel(float* __restrict__ a, long na, float* __restrict__ b,
                                                                 long nb) {
  pdl_launch_dependents();
  pdl_wait();
  const long stride = static_cast<long>(gridDim.x) * blockDim.x;
  const long first = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x;
  for (long i = first; i < na; i += stride) a[i] = 0.f;
  for (long i = first; i < nb; i += stride) b[i] = 0.f;
}

// Device-to-device copy of up to two ranges of 16-byte words in one launch (a step's staged images and
// labels into the buffers the captured step reads).
__global__ void __launch_bounds__(256) copy_pair_kernel(uint4* __restrict__ d0, const uint4* __restrict__ s0, long n0,
                                                        uint4* __restrict__ d1, const uint4* __restrict__ s1, long n1) {
  pdl_launch_dependents();
  pdl_wait();
  const long stride = static_cast<long>(gridDim.x) * blockDim.x;
  const long first = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x;
  for (long i = first; i < n0; i += stride) d0[i] = s0[i];
  for (long i = first; i < n1; i += stride) d1[i] = s1[i];
}

// master fp32 [Cout][taps][Cin] -> Wf bf16 [Cout][kpad] (zero padded past taps*Cin)
//                               -> Wd bf16 [Cin][taps*Cout] with taps flipped (dgrad operand), optional
// T = __nv_bfloat16: operand copies of the bf16 path.  T = float: the tf32 path, whose forward operand is the master
// tensor itself wherever no padding is needed (wf == nullptr), so usually only Wd is produced.
template <class T>
__global__ void prep_conv_weights_kernel(const float* __restrict__ master, T* __restrict__ wf, T* __restrict__ wd,
                                         int Cout, int Cin, int taps, int kpad) {
  pdl_launch_dependents();
  pdl_wait();
  const long nf = wf != nullptr ? static_cast<long>(Cout) * kpad : 0;
  const long nd = wd != nullptr ? static_cast<long>(Cin) * taps * Cout : 0;
  const int kvalid = taps * Cin;
  const long stride = static_cast<long>(gridDim.x) * blockDim.x;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < nf + nd; i += stride) {
    if (i < nf) {
      const int co = static_cast<int>(i / kpad), k = static_cast<int>(i - static_cast<long>(co) * kpad);
      wf[i] = operand_cast<T>(k < kvalid ? master[static_cast<long>(co) * kvalid + k] : 0.f);
    } else {
      const long j = i - nf;
      const int co = static_cast<int>(j % Cout);
      const int tap = static_cast<int>((j / Cout) % taps);
      const int ci = static_cast<int>(j / (static_cast<long>(Cout) * taps));
      wd[j] = operand_cast<T>(master[(static_cast<long>(co) * taps + (taps - 1 - tap)) * Cin + ci]);
    }
  }
}

// All conv layers in ONE launch: 64(co) x 64(ci) tiles of one tap are read with coalesced float4
// loads, written to Wf in the operand type and transposed through shared memory into Wd (tap flipped).
constexpr int kPrepMaxLayers = 20;
struct PrepLayer {
  const float* master;
  void* wf;   // may be null (tf32: the master tensor is the forward operand)
  void* wd;   // may be null
  int cout, cin, taps, kpad;
  int tile_begin;  // first global tile index of this layer
  int ci_tiles;
};
struct PrepTable {
  PrepLayer l[kPrepMaxLayers];
  int n;
};

template <class T>
__global__ void __launch_bounds__(256) prep_conv_weights_batched_kernel(const PrepTable t) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float tile[64][65];
  int li = 0;
  while (li + 1 < t.n && static_cast<int>(blockIdx.x) >= t.l[li + 1].tile_begin) ++li;
  const PrepLayer L = t.l[li];
  T* wf = static_cast<T*>(L.wf);
  T* wd = static_cast<T*>(L.wd);
  int tidx = blockIdx.x - L.tile_begin;
  const int ci_t = tidx % L.ci_tiles;
  tidx /= L.ci_tiles;
  const int tap = tidx % L.taps;
  const int co_t = tidx / L.taps;
  const int co0 = co_t * 64, ci0 = ci_t * 64;
  const int kvalid = L.taps * L.cin;
  const int c4 = (threadIdx.x & 15) * 4;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int r = (threadIdx.x >> 4) + 16 * k;
    const float4 v = *reinterpret_cast<const float4*>(L.master + static_cast<long>(co0 + r) * kvalid + tap * L.cin + ci0 + c4);
    tile[r][c4] = v.x; tile[r][c4 + 1] = v.y; tile[r][c4 + 2] = v.z; tile[r][c4 + 3] = v.w;
    if (wf != nullptr)
      operand_store4(wf + static_cast<long>(co0 + r) * L.kpad + tap * L.cin + ci0 + c4, v.x, v.y, v.z, v.w);
  }
  if (wd == nullptr) return;
  __syncthreads();
  const int ftap = L.taps - 1 - tap;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int ci = (threadIdx.x >> 4) + 16 * k;
    operand_store4(wd + (static_cast<long>(ci0 + ci) * L.taps + ftap) * L.cout + co0 + c4, tile[c4][ci],
                   tile[c4 + 1][ci], tile[c4 + 2][ci], tile[c4 + 3][ci]);
  }
}

// The BatchNorm kernels run as bulk-copy pipelines (csrc/bn_stream.cu, bn_bwd_apply_pool_tma_kernel) wherever the
// shape fits their stage geometry; the register-file kernels in this file take the remaining shapes.
static bool bn_tma_shape_ok(int H, int W, int C, int pool) {
  if (C < 8 || C % 8 != 0 || C > 2048 || 256 % (C / 8) != 0) return false;
  if (!pool) return true;
  const int items = (W / 2) * (C / 8);
  return ((H | W) & 1) == 0 && items >= 1 && items <= 256 && 256 % items == 0;
}

static inline int grid_for(long work_items, int threads, int max_blocks = 148 * 16) {
  long b = (work_items + threads - 1) / threads;
  if (b > max_blocks) b = max_blocks;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}

}  // namespace b200dp

using namespace b200dp;

/* dtype 0: bf16 patch rows of 64 columns; dtype 1: fp32 patch rows of 32 columns (27 real columns either way).
 * flags bit 0: the zero half of the bf16 rows (columns 32..63) is already zero in a0 and is not rewritten. */
extern "C" int b200dp_im2col3x3_c3_ex2(const float* x_nchw, void* a0, int B, int H, int W, int dtype, int flags,
                                       b200dp_stream_t s) {
  const size_t smem = (kIm2colRows + 2) * 3 * (W + 2) * sizeof(float);
  const int groups = (H + kIm2colRows - 1) / kIm2colRows;
  if (dtype == 1)
    B200DP_CUDA_TRY(launch_pdl(im2col_c3_v2_kernel<float, 32, 32>, dim3(B * groups), dim3(256), smem,
                               static_cast<cudaStream_t>(s), x_nchw, static_cast<float*>(a0), B, H, W));
  else if (flags & 1)
    B200DP_CUDA_TRY(launch_pdl(im2col_c3_v2_kernel<__nv_bfloat16, 64, 32>, dim3(B * groups), dim3(256), smem,
                               static_cast<cudaStream_t>(s), x_nchw, static_cast<__nv_bfloat16*>(a0), B, H, W));
  else
    B200DP_CUDA_TRY(launch_pdl(im2col_c3_v2_kernel<__nv_bfloat16, 64, 64>, dim3(B * groups), dim3(256), smem,
                               static_cast<cudaStream_t>(s), x_nchw, static_cast<__nv_bfloat16*>(a0), B, H, W));
  return 0;
}
extern "C" int b200dp_im2col3x3_c3_ex(const float* x_nchw, void* a0, int B, int H, int W, int dtype,
                                      b200dp_stream_t s) {
  return b200dp_im2col3x3_c3_ex2(x_nchw, a0, B, H, W, dtype, 0, s);
}
extern "C" int b200dp_im2col3x3_c3(const float* x_nchw, void* a0, int B, int H, int W, b200dp_stream_t s) {
  return b200dp_im2col3x3_c3_ex(x_nchw, a0, B, H, W, 0, s);
}

// Forward declarations of the bulk-copy pipeline entry points (csrc/bn_stream.cu); dtype 0 = bf16, 1 = fp32.
extern "C" int b200dp_bn_relu_pool_fwd_tma_ex(const void* y, const float* stats, const float* gamma, const float* beta,
                                              float* running_mean, float* running_var, long long* num_batches_tracked,
                                              float* save_mean, float* save_invstd, void* out, int B, int H, int W,
                                              int C, int pool, int training, float momentum, float eps, int dtype,
                                              b200dp_stream_t s);
extern "C" int b200dp_bn_bwd_apply_nopool_tma_ex(const void* y, const void* dout, const float* mean,
                                                 const float* invstd, const float* gamma, const float* beta,
                                                 const float* sums, void* dy, float* dgamma, float* dbeta, int B,
                                                 int H, int W, int C, int dtype, b200dp_stream_t s);
extern "C" int b200dp_bn_bwd_reduce_from_out_tma_ex(const void* out, const void* dout, const float* gamma,
                                                    const float* beta, float* sums, long long Q, int C, int dtype,
                                                    b200dp_stream_t s);

extern "C" int b200dp_bn_stats_ex(const void* y, float* stats, long long P, int C, int dtype, b200dp_stream_t s) {
  if (C % 8 != 0 || C > 2048) {
    set_error("bn_stats: C=%d unsupported", C);
    return B200DP_EINVAL;
  }
  const int threads = 256;
  const int rows_per_block = threads / (C / 8);
  if (rows_per_block < 1) {
    set_error("bn_stats: C=%d too wide for a 256-thread block", C);
    return B200DP_EINVAL;
  }
  const size_t smem = 2 * rows_per_block * C * sizeof(float);
  int blocks = grid_for(P, rows_per_block * 8, 148 * 4);
  if (dtype == 1)
    bn_stats_kernel<float><<<blocks, threads, smem, static_cast<cudaStream_t>(s)>>>(static_cast<const float*>(y), stats,
                                                                                   P, C);
  else
    bn_stats_kernel<__nv_bfloat16><<<blocks, threads, smem, static_cast<cudaStream_t>(s)>>>(
        static_cast<const __nv_bfloat16*>(y), stats, P, C);
  B200DP_CUDA_TRY(cudaGetLastError());
  return 0;
}
extern "C" int b200dp_bn_stats(const void* y, float* stats, long long P, int C, b200dp_stream_t s) {
  return b200dp_bn_stats_ex(y, stats, P, C, 0, s);
}

extern "C" int b200dp_bn_relu_pool_fwd_ex(const void* y, const float* stats, const float* gamma, const float* beta,
                                          float* running_mean, float* running_var, long long* num_batches_tracked,
                                          float* save_mean, float* save_invstd, void* out, int B, int H, int W, int C,
                                          int pool, int training, float momentum, float eps, int dtype,
                                          b200dp_stream_t s) {
  if (C % 8 != 0 || (pool && ((H | W) & 1))) {
    set_error("bn_relu_pool_fwd: unsupported shape H=%d W=%d C=%d pool=%d", H, W, C, pool);
    return B200DP_EINVAL;
  }
  if (bn_tma_shape_ok(H, W, C, pool))
    return b200dp_bn_relu_pool_fwd_tma_ex(y, stats, gamma, beta, running_mean, running_var, num_batches_tracked,
                                          save_mean, save_invstd, out, B, H, W, C, pool, training, momentum, eps, dtype,
                                          s);
  BnFwdParams p{y, stats, gamma, beta, running_mean, running_var, num_batches_tracked, save_mean, save_invstd, out,
                B, H, W, C, pool, training, momentum, eps};
  if (C > 2048) {
    set_error("bn_relu_pool_fwd: C=%d too wide for a 256-thread block", C);
    return B200DP_EINVAL;
  }
  const long Q = static_cast<long>(B) * (pool ? H / 2 : H) * (pool ? W / 2 : W);
  const int rows_per_block = 256 / (C / 8);
  // two output pixels per thread per iteration; at most one resident wave (3 blocks / SM)
  const dim3 grid(grid_for(Q, rows_per_block * 2, 148 * 3));
  if (dtype == 1)
    B200DP_CUDA_TRY(launch_pdl(bn_relu_pool_fwd_kernel<float>, grid, dim3(256), 0, static_cast<cudaStream_t>(s), p));
  else
    B200DP_CUDA_TRY(launch_pdl(bn_relu_pool_fwd_kernel<__nv_bfloat16>, grid, dim3(256), 0,
                               static_cast<cudaStream_t>(s), p));
  return 0;
}
extern "C" int b200dp_bn_relu_pool_fwd(const void* y, const float* stats, const float* gamma, const float* beta,
                                       float* running_mean, float* running_var, long long* num_batches_tracked,
                                       float* save_mean, float* save_invstd, void* out, int B, int H, int W, int C,
                                       int pool, int training, float momentum, float eps, b200dp_stream_t s) {
  return b200dp_bn_relu_pool_fwd_ex(y, stats, gamma, beta, running_mean, running_var, num_batches_tracked, save_mean,
                                    save_invstd, out, B, H, W, C, pool, training, momentum, eps, 0, s);
}

template <class T>
static bool bn_tma_supported(const BnBwdParams& p) {
  constexpr int kItems = bn_tma_items<T>();
  if (!p.pool || p.C % 8 != 0 || ((p.H | p.W) & 1)) return false;
  const int items_per_tile = (p.W / 2) * (p.C / 8);
  return items_per_tile >= 1 && items_per_tile <= kItems && kItems % items_per_tile == 0;
}

template <class T>
static int launch_bn_apply_pool_tma(const BnBwdParams& p, cudaStream_t s) {
  constexpr int kItems = bn_tma_items<T>();
  static bool configured = false;
  if (!configured) {
    B200DP_CUDA_TRY(cudaFuncSetAttribute(bn_bwd_apply_pool_tma_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kBnTmaSmem));
    configured = true;
  }
  const int G = kItems / ((p.W / 2) * (p.C / 8));
  const long tiles = static_cast<long>(p.B) * (p.H / 2);
  const long num_chunks = (tiles + G - 1) / G;
  const int grid = static_cast<int>(num_chunks < 148 * 2 ? num_chunks : 148 * 2);
  B200DP_CUDA_TRY(launch_pdl(bn_bwd_apply_pool_tma_kernel<T>, dim3(grid), dim3(kItems), kBnTmaSmem, s, p));
  return 0;
}

template <class T>
static int launch_bn_bwd(bool apply, const BnBwdParams& p, cudaStream_t s) {
  if (p.C % 8 != 0 || p.C > 2048 || (p.pool && ((p.H | p.W) & 1))) {
    set_error("bn_relu_pool_bwd: unsupported shape H=%d W=%d C=%d pool=%d", p.H, p.W, p.C, p.pool);
    return B200DP_EINVAL;
  }
  constexpr int kDtype = sizeof(T) == 4 ? 1 : 0;
  const int threads = 256;
  const int rows_per_block = threads / (p.C / 8);
  const size_t smem = apply ? 0 : 2 * rows_per_block * p.C * sizeof(float);
  const long Q = static_cast<long>(p.B) * (p.pool ? p.H / 2 : p.H) * (p.pool ? p.W / 2 : p.W);
  // reduce pass: at most one resident wave (3 blocks / SM), a grid-stride loop balances better than a ragged tail
  const int blocks = grid_for(Q, rows_per_block, 148 * 3);
  if (apply && p.pool && bn_tma_supported<T>(p)) return launch_bn_apply_pool_tma<T>(p, s);
  if (apply && !p.pool && bn_tma_shape_ok(p.H, p.W, p.C, 0))
    return b200dp_bn_bwd_apply_nopool_tma_ex(p.y, p.dout, p.mean, p.invstd, p.gamma, p.beta, p.sums, p.dy, p.dgamma,
                                             p.dbeta, p.B, p.H, p.W, p.C, kDtype, s);
  if (apply && !p.pool)
    B200DP_CUDA_TRY(launch_pdl(bn_bwd_apply_nopool_kernel<T>, dim3(grid_for(Q, rows_per_block * 4, 148 * 3)),
                               dim3(threads), 0, s, p));
  else if (apply && p.pool)
    B200DP_CUDA_TRY(launch_pdl(bn_bwd_apply_pool_kernel<T>, dim3(grid_for(Q, rows_per_block * 2, 148 * 4)),
                               dim3(threads), 0, s, p));
  else
    B200DP_CUDA_TRY(launch_pdl(bn_relu_pool_bwd_kernel<T, false>, dim3(blocks), dim3(threads), smem, s, p));
  return 0;
}
static int launch_bn_bwd_dt(int dtype, bool apply, const BnBwdParams& p, cudaStream_t s) {
  return dtype == 1 ? launch_bn_bwd<float>(apply, p, s) : launch_bn_bwd<__nv_bfloat16>(apply, p, s);
}

extern "C" int b200dp_bn_relu_pool_bwd_reduce_ex(const void* y, const void* dout, const float* mean,
                                                 const float* invstd, const float* gamma, const float* beta,
                                                 float* sums, int B, int H, int W, int C, int pool, int dtype,
                                                 b200dp_stream_t s) {
  BnBwdParams p{y, dout, mean, invstd, gamma, beta, sums, nullptr, nullptr, nullptr, B, H, W, C, pool};
  return launch_bn_bwd_dt(dtype, false, p, static_cast<cudaStream_t>(s));
}
extern "C" int b200dp_bn_relu_pool_bwd_reduce(const void* y, const void* dout, const float* mean,
                                              const float* invstd, const float* gamma, const float* beta,
                                              float* sums, int B, int H, int W, int C, int pool,
                                              b200dp_stream_t s) {
  return b200dp_bn_relu_pool_bwd_reduce_ex(y, dout, mean, invstd, gamma, beta, sums, B, H, W, C, pool, 0, s);
}

/* Pass 1 computed from the block's stored OUTPUT (`out`, the same shape as `dout`) instead of the conv
 * output: fewer bytes for pooled layers, no arg-max search.  Same `sums` contract as _reduce. */
extern "C" int b200dp_bn_relu_pool_bwd_reduce_from_out_ex(const void* out, const void* dout, const float* gamma,
                                                          const float* beta, float* sums, long long Q, int C,
                                                          int dtype, b200dp_stream_t s) {
  if (C % 8 != 0 || C > 2048 || Q <= 0) {
    set_error("bn_relu_pool_bwd_reduce_from_out: unsupported shape Q=%lld C=%d", Q, C);
    return B200DP_EINVAL;
  }
  if (bn_tma_shape_ok(2, 2, C, 0))
    return b200dp_bn_bwd_reduce_from_out_tma_ex(out, dout, gamma, beta, sums, Q, C, dtype, s);
  const int threads = 256;
  const int rows_per_block = threads / (C / 8);
  const size_t smem = 2 * rows_per_block * C * sizeof(float);
  const int blocks = grid_for(Q, rows_per_block * 2, 148 * 4);
  if (dtype == 1)
    B200DP_CUDA_TRY(launch_pdl(bn_bwd_reduce_from_out_kernel<float>, dim3(blocks), dim3(threads), smem,
                               static_cast<cudaStream_t>(s), static_cast<const float*>(out),
                               static_cast<const float*>(dout), gamma, beta, sums, static_cast<long>(Q), C));
  else
    B200DP_CUDA_TRY(launch_pdl(bn_bwd_reduce_from_out_kernel<__nv_bfloat16>, dim3(blocks), dim3(threads), smem,
                               static_cast<cudaStream_t>(s), static_cast<const __nv_bfloat16*>(out),
                               static_cast<const __nv_bfloat16*>(dout), gamma, beta, sums, static_cast<long>(Q), C));
  return 0;
}
extern "C" int b200dp_bn_relu_pool_bwd_reduce_from_out(const void* out, const void* dout, const float* gamma,
                                                       const float* beta, float* sums, long long Q, int C,
                                                       b200dp_stream_t s) {
  return b200dp_bn_relu_pool_bwd_reduce_from_out_ex(out, dout, gamma, beta, sums, Q, C, 0, s);
}

extern "C" int b200dp_bn_relu_pool_bwd_apply_ex(const void* y, const void* dout, const float* mean,
                                                const float* invstd, const float* gamma, const float* beta,
                                                const float* sums, void* dy, float* dgamma, float* dbeta, int B,
                                                int H, int W, int C, int pool, int dtype, b200dp_stream_t s) {
  BnBwdParams p{y, dout, mean, invstd, gamma, beta, const_cast<float*>(sums), dy, dgamma, dbeta, B, H, W, C, pool};
  return launch_bn_bwd_dt(dtype, true, p, static_cast<cudaStream_t>(s));
}
extern "C" int b200dp_bn_relu_pool_bwd_apply(const void* y, const void* dout, const float* mean,
                                             const float* invstd, const float* gamma, const float* beta,
                                             const float* sums, void* dy, float* dgamma, float* dbeta, int B, int H,
                                             int W, int C, int pool, b200dp_stream_t s) {
  return b200dp_bn_relu_pool_bwd_apply_ex(y, dout, mean, invstd, gamma, beta, sums, dy, dgamma, dbeta, B, H, W, C, pool,
                                          0, s);
}

/* The pooled apply pass as a bulk-copy (TMA engine) pipeline: same contract as b200dp_bn_relu_pool_bwd_apply
 * with pool = 1 (which routes here whenever the shape fits).  Needs (W/2) * (C/8) to divide 256 (bf16). */
extern "C" int b200dp_bn_bwd_apply_pool_tma(const void* y, const void* dout, const float* mean, const float* invstd,
                                            const float* gamma, const float* beta, const float* sums, void* dy,
                                            float* dgamma, float* dbeta, int B, int H, int W, int C,
                                            b200dp_stream_t s) {
  BnBwdParams p{y, dout, mean, invstd, gamma, beta, const_cast<float*>(sums), dy, dgamma, dbeta, B, H, W, C, 1};
  if (B <= 0 || !bn_tma_supported<__nv_bfloat16>(p)) {
    set_error("bn_bwd_apply_pool_tma: unsupported shape B=%d H=%d W=%d C=%d ((W/2)*(C/8) must divide 256)", B, H, W, C);
    return B200DP_EINVAL;
  }
  return launch_bn_apply_pool_tma<__nv_bfloat16>(p, static_cast<cudaStream_t>(s));
}

extern "C" int b200dp_linear_fwd_ex(const void* x, const float* w, const float* b, float* logits, int B, int F, int O,
                                    int dtype, b200dp_stream_t s) {
  if (F % 8 != 0) {
    set_error("linear_fwd: F=%d must be a multiple of 8", F);
    return B200DP_EINVAL;
  }
  if (O > kLinMaxO) {
    set_error("linear_fwd: O=%d exceeds the supported %d outputs", O, kLinMaxO);
    return B200DP_EINVAL;
  }
  const int threads = 128;
  const int blocks = (B * 32 + threads - 1) / threads;
  if (dtype == 1)
    B200DP_CUDA_TRY(launch_pdl(linear_fwd_kernel<float>, dim3(blocks), dim3(threads), 0, static_cast<cudaStream_t>(s),
                               static_cast<const float*>(x), w, b, logits, B, F, O));
  else
    B200DP_CUDA_TRY(launch_pdl(linear_fwd_kernel<__nv_bfloat16>, dim3(blocks), dim3(threads), 0,
                               static_cast<cudaStream_t>(s), static_cast<const __nv_bfloat16*>(x), w, b, logits, B, F,
                               O));
  return 0;
}
extern "C" int b200dp_linear_fwd(const void* x, const float* w, const float* b, float* logits, int B, int F, int O,
                                 b200dp_stream_t s) {
  return b200dp_linear_fwd_ex(x, w, b, logits, B, F, O, 0, s);
}

extern "C" int b200dp_cross_entropy(const float* logits, const long long* target, float* loss, float* dlogits,
                                    int* correct, int B, int O, float grad_scale, b200dp_stream_t s) {
  B200DP_CUDA_TRY(launch_pdl(cross_entropy_kernel, dim3(1), dim3(256), 0, static_cast<cudaStream_t>(s), logits, target,
                             loss, dlogits, correct, B, O, grad_scale));
  return 0;
}

extern "C" int b200dp_linear_bwd_ex(const void* x, const float* w, const float* dlogits, float* dw, float* db,
                                    void* dx, int B, int F, int O, int dtype, b200dp_stream_t s) {
  if (O > kLinMaxO) {
    set_error("linear_bwd: O=%d exceeds the supported %d outputs", O, kLinMaxO);
    return B200DP_EINVAL;
  }
  const int threads = 256;
  const int nslices = (B + 31) / 32 < 16 ? (B + 31) / 32 : 16;   // <= kLinTileB rows per slice
  if ((B + nslices - 1) / nslices > kLinTileB) {
    set_error("linear_bwd: batch %d too large for the staged dlogits tile", B);
    return B200DP_EINVAL;
  }
  const int nW = ((F + 63) / 64) * nslices;
  const int nX = static_cast<int>((static_cast<long>(B) * F + threads - 1) / threads);
  cudaStream_t st = static_cast<cudaStream_t>(s);
  // dw / db are accumulated with atomics: cleared by a kernel of this library (no memset nodes in a captured step)
  B200DP_CUDA_TRY(launch_pdl(fill_zero_f32_pair_kernel, dim3(grid_for(static_cast<long>(O) * F, 256, 64)), dim3(256), 0,
                             st, dw, static_cast<long>(O) * F, db, static_cast<long>(O)));
  if (dtype == 1)
    B200DP_CUDA_TRY(launch_pdl(linear_bwd_kernel<float>, dim3(nW + nX), dim3(threads), 0, st,
                               static_cast<const float*>(x), w, dlogits, dw, db, static_cast<float*>(dx), B, F, O, nW));
  else
    B200DP_CUDA_TRY(launch_pdl(linear_bwd_kernel<__nv_bfloat16>, dim3(nW + nX), dim3(threads), 0, st,
                               static_cast<const __nv_bfloat16*>(x), w, dlogits, dw, db,
                               static_cast<__nv_bfloat16*>(dx), B, F, O, nW));
  return 0;
}
extern "C" int b200dp_linear_bwd(const void* x, const float* w, const float* dlogits, float* dw, float* db, void* dx,
                                 int B, int F, int O, b200dp_stream_t s) {
  return b200dp_linear_bwd_ex(x, w, dlogits, dw, db, dx, B, F, O, 0, s);
}

template <class T>
static int head_fwd_bwd_impl(const void* x, const float* w, const float* bias, const long long* target, float* logits,
                             float* loss, int* correct, float* dlogits, void* dx, float* dw, float* db, int B, int F,
                             int O, float grad_scale, cudaStream_t s) {
  const int blocks = (B + kHeadWarps - 1) / kHeadWarps;
  const size_t smem = (static_cast<size_t>(O + kHeadWarps) * F + kHeadWarps * kLinMaxO) * sizeof(float);
  static bool configured = false;
  if (!configured) {
    B200DP_CUDA_TRY(cudaFuncSetAttribute(head_fwd_bwd_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(((kLinMaxO + kHeadWarps) * 2048 + kHeadWarps * kLinMaxO) *
                                                          sizeof(float))));
    configured = true;
  }
  B200DP_CUDA_TRY(launch_pdl(head_fwd_bwd_kernel<T>, dim3(blocks), dim3(kHeadWarps * 32), smem, s,
                             static_cast<const T*>(x), w, bias, target, logits, loss, correct, dlogits,
                             static_cast<T*>(dx), dw, db, B, F, O, grad_scale));
  return 0;
}

/* Training-step classifier head in one launch: logits = x W^T + b, mean softmax cross-entropy into *loss,
 * top-1 count into *correct, dlogits (may be NULL), dx = dlogits W (activation type), and
 * dw += dlogits^T x, db += sum dlogits (dw / db must be zero or hold the value to accumulate onto --
 * unlike b200dp_linear_bwd this entry point does not clear them).  F % 8 == 0, F <= 2048, O <= 16,
 * B <= 4096, x / dx / w / dw 16-byte aligned.  Launches on one stream only (it keeps its block counter in a device global). */
extern "C" int b200dp_head_fwd_bwd_ex(const void* x, const float* w, const float* bias, const long long* target,
                                      float* logits, float* loss, int* correct, float* dlogits, void* dx, float* dw,
                                      float* db, int B, int F, int O, float grad_scale, int dtype, b200dp_stream_t s) {
  if (F % 8 != 0 || F > 2048 || O > kLinMaxO || O < 1 || B < 1 || B > kHeadWarps * kHeadMaxBlocks) {
    set_error("head_fwd_bwd: unsupported shape B=%d F=%d O=%d", B, F, O);
    return B200DP_EINVAL;
  }
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (dtype == 1)
    return head_fwd_bwd_impl<float>(x, w, bias, target, logits, loss, correct, dlogits, dx, dw, db, B, F, O, grad_scale,
                                    st);
  return head_fwd_bwd_impl<__nv_bfloat16>(x, w, bias, target, logits, loss, correct, dlogits, dx, dw, db, B, F, O,
                                          grad_scale, st);
}
extern "C" int b200dp_head_fwd_bwd(const void* x, const float* w, const float* bias, const long long* target,
                                   float* logits, float* loss, int* correct, float* dlogits, void* dx, float* dw,
                                   float* db, int B, int F, int O, float grad_scale, b200dp_stream_t s) {
  return b200dp_head_fwd_bwd_ex(x, w, bias, target, logits, loss, correct, dlogits, dx, dw, db, B, F, O, grad_scale, 0,
                                s);
}

extern "C" int b200dp_sgd_momentum(float* p, const float* g, float* buf, long long n, float lr, float momentum,
                                   float wd, float grad_scale, b200dp_stream_t s) {
  if ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(buf)) & 15) {
    set_error("sgd_momentum: buffers must be 16-byte aligned");
    return B200DP_EINVAL;
  }
  B200DP_CUDA_TRY(launch_pdl(sgd_momentum_kernel, dim3(grid_for(n / 4 + 1, 256, 148 * 8)), dim3(256), 0,
                             static_cast<cudaStream_t>(s), p, g, buf, static_cast<long>(n), lr, momentum, wd,
                             grad_scale));
  return 0;
}

extern "C" int b200dp_fill_zero(void* dst, long long nbytes, b200dp_stream_t s) {
  if ((reinterpret_cast<uintptr_t>(dst) & 15) || (nbytes & 15) || nbytes < 0) {
    set_error("fill_zero: pointer and byte count (%lld) must be multiples of 16", nbytes);
    return B200DP_EINVAL;
  }
  if (nbytes == 0) return 0;
  B200DP_CUDA_TRY(launch_pdl(fill_zero_kernel, dim3(grid_for(nbytes / 16, 256 * 4, 148 * 8)), dim3(256), 0,
                             static_cast<cudaStream_t>(s), static_cast<uint4*>(dst), static_cast<long>(nbytes / 16)));
  return 0;
}

extern "C" int b200dp_copy_pair(void* dst0, const void* src0, long long nbytes0, void* dst1, const void* src1,
                                long long nbytes1, b200dp_stream_t s) {
  const uintptr_t bits = reinterpret_cast<uintptr_t>(dst0) | reinterpret_cast<uintptr_t>(src0) |
                         reinterpret_cast<uintptr_t>(dst1) | reinterpret_cast<uintptr_t>(src1);
  if ((bits & 15) || (nbytes0 & 15) || (nbytes1 & 15) || nbytes0 < 0 || nbytes1 < 0) {
    set_error("copy_pair: pointers and byte counts (%lld, %lld) must be multiples of 16", nbytes0, nbytes1);
    return B200DP_EINVAL;
  }
  const long n = (nbytes0 > nbytes1 ? nbytes0 : nbytes1) / 16;
  if (n == 0) return 0;
  B200DP_CUDA_TRY(launch_pdl(copy_pair_kernel, dim3(grid_for(n, 256 * 2, 148 * 4)), dim3(256), 0,
                             static_cast<cudaStream_t>(s), static_cast<uint4*>(dst0), static_cast<const uint4*>(src0),
                             static_cast<long>(nbytes0 / 16), static_cast<uint4*>(dst1),
                             static_cast<const uint4*>(src1), static_cast<long>(nbytes1 / 16)));
  return 0;
}

/* dtype 0: bf16 operand copies; 1: fp32 copies for the tf32 kernels (wf may be NULL there: the master tensor is
 * the forward operand whenever kpad == taps*Cin). */
extern "C" int b200dp_prep_conv_weights_ex(const float* master, void* wf, void* wd, int Cout, int Cin, int taps,
                                           int kpad, int dtype, b200dp_stream_t s) {
  if (kpad < taps * Cin || (wf == nullptr && wd == nullptr)) {
    set_error("prep_conv_weights: kpad=%d < taps*Cin=%d, or nothing to write", kpad, taps * Cin);
    return B200DP_EINVAL;
  }
  const long n = (wf ? static_cast<long>(Cout) * kpad : 0) + (wd ? static_cast<long>(Cin) * taps * Cout : 0);
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (dtype == 1)
    prep_conv_weights_kernel<float><<<grid_for(n, 256), 256, 0, st>>>(master, static_cast<float*>(wf),
                                                                      static_cast<float*>(wd), Cout, Cin, taps, kpad);
  else
    prep_conv_weights_kernel<__nv_bfloat16><<<grid_for(n, 256), 256, 0, st>>>(
        master, static_cast<__nv_bfloat16*>(wf), static_cast<__nv_bfloat16*>(wd), Cout, Cin, taps, kpad);
  B200DP_CUDA_TRY(cudaGetLastError());
  return 0;
}

extern "C" int b200dp_prep_conv_weights(const float* master, void* wf, void* wd, int Cout, int Cin, int taps,
                                        int kpad, b200dp_stream_t s) {
  return b200dp_prep_conv_weights_ex(master, wf, wd, Cout, Cin, taps, kpad, 0, s);
}

// ---------------------------------------------------------------------------
// BatchNorm + ReLU (+ pool) backward for callers that must not allocate.
//
// b200dp_bn_relu_pool_bwd_reduce / _apply need a caller-owned [2][C] fp32 buffer for the
// per-channel sums.  Some hosts cannot provide one at the point of the call: an application that
// handed the whole device to its own arena at start-up, a plug-in running under a fixed memory
// plan, or a loop where cudaMalloc's device-wide synchronisation is not acceptable.  For them this
// entry point carries its own scratch: one file-scope __device__ array, resident on every GPU
// that loads the library until the process exits.
// ---------------------------------------------------------------------------
#include <mutex>

namespace b200dp {

constexpr int kNoAllocMaxChannels = 2048;  // 2 * 2048 floats = 16 KiB of device memory per GPU
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) float g_bn_bwd_scratch[2 * kNoAllocMaxChannels];

// The array's name is only a host-side handle; kernels must be given the device address that
// cudaGetSymbolAddress returns for the device that is current at the time of the query.
static float* g_bn_bwd_scratch_addr[kNoAllocMaxDevices] = {nullptr};
static std::mutex g_bn_bwd_scratch_mutex;

static int bn_bwd_scratch_for_current_device(float** out) {
  int dev = -1;
  B200DP_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) {
    set_error("bn_relu_pool_bwd_noalloc: device ordinal %d is outside the supported range [0, %d)", dev,
              kNoAllocMaxDevices);
    return B200DP_EINVAL;
  }
  std::lock_guard<std::mutex> lock(g_bn_bwd_scratch_mutex);
  if (g_bn_bwd_scratch_addr[dev] == nullptr) {
    void* p = nullptr;
    B200DP_CUDA_TRY(cudaGetSymbolAddress(&p, g_bn_bwd_scratch));
    g_bn_bwd_scratch_addr[dev] = static_cast<float*>(p);
  }
  *out = g_bn_bwd_scratch_addr[dev];
  return 0;
}

}  // namespace b200dp

extern "C" int b200dp_bn_relu_pool_bwd_noalloc(const void* y, const void* dout, const float* mean,
                                               const float* invstd, const float* gamma, const float* beta,
                                               void* dy, float* dgamma, float* dbeta, int B, int H, int W, int C,
                                               int pool, b200dp_stream_t s) {
  if (C <= 0 || C > kNoAllocMaxChannels) {
    set_error("bn_relu_pool_bwd_noalloc: C=%d exceeds the built-in scratch, which holds at most %d channels; "
              "use b200dp_bn_relu_pool_bwd_reduce/_apply with your own [2][C] buffer instead",
              C, kNoAllocMaxChannels);
    return B200DP_EINVAL;
  }
  float* scratch = nullptr;
  int rc = bn_bwd_scratch_for_current_device(&scratch);
  if (rc) return rc;
  cudaStream_t stream = static_cast<cudaStream_t>(s);
  // The array still holds the sums of whoever called last; the reduction adds onto it.
  B200DP_CUDA_TRY(cudaMemsetAsync(scratch, 0, 2 * static_cast<size_t>(C) * sizeof(float), stream));
  BnBwdParams p{y, dout, mean, invstd, gamma, beta, scratch, dy, dgamma, dbeta, B, H, W, C, pool};
  rc = launch_bn_bwd<__nv_bfloat16>(false, p, stream);
  if (rc) return rc;
  return launch_bn_bwd<__nv_bfloat16>(true, p, stream);
}

/* Batched form of b200dp_prep_conv_weights_ex: all layers in one launch.  Arrays have n entries;
 * layers need Cout % 64 == 0 and Cin % 64 == 0 (others: use the single-layer entry point). */
extern "C" int b200dp_prep_conv_weights_batched_ex(int n, const float* const* master, void* const* wf,
                                                   void* const* wd, const int* cout, const int* cin, const int* taps,
                                                   const int* kpad, int dtype, b200dp_stream_t s) {
  if (n < 1 || n > kPrepMaxLayers) {
    set_error("prep_conv_weights_batched: n=%d outside [1, %d]", n, kPrepMaxLayers);
    return B200DP_EINVAL;
  }
  PrepTable t{};
  t.n = n;
  int tiles = 0;
  for (int i = 0; i < n; ++i) {
    if (cout[i] % 64 || cin[i] % 64 || kpad[i] != taps[i] * cin[i]) {
      set_error("prep_conv_weights_batched: layer %d (Cout=%d Cin=%d kpad=%d) needs multiples of 64 and no padding",
                i, cout[i], cin[i], kpad[i]);
      return B200DP_EINVAL;
    }
    t.l[i] = PrepLayer{master[i], wf[i], wd[i], cout[i], cin[i], taps[i], kpad[i], tiles, cin[i] / 64};
    tiles += (cout[i] / 64) * taps[i] * (cin[i] / 64);
  }
  if (dtype == 1)
    B200DP_CUDA_TRY(launch_pdl(prep_conv_weights_batched_kernel<float>, dim3(tiles), dim3(256), 0,
                               static_cast<cudaStream_t>(s), t));
  else
    B200DP_CUDA_TRY(launch_pdl(prep_conv_weights_batched_kernel<__nv_bfloat16>, dim3(tiles), dim3(256), 0,
                               static_cast<cudaStream_t>(s), t));
  return 0;
}

extern "C" int b200dp_prep_conv_weights_batched(int n, const float* const* master, void* const* wf, void* const* wd,
                                                const int* cout, const int* cin, const int* taps, const int* kpad,
                                                b200dp_stream_t s) {
  return b200dp_prep_conv_weights_batched_ex(n, master, wf, wd, cout, cin, taps, kpad, 0, s);
}
