// tcgen05 / TMEM / TMA implicit-GEMM kernels for the VGG conv layers.
//
//   conv_igemm_kernel  : forward conv3x3(pad 1) and dgrad.  GEMM view
//        D[pixels, Cout] = sum_{tap, cin-block}  A_tap[pixels, 64] * W[Cout, tap*Cin + 64-block]^T
//     The A tile of a tap is fetched by ONE 4-D TMA box over the NHWC activation
//     tensor, shifted by (dr, ds); rows/columns that fall outside the image are
//     zero-filled by the TMA unit, which is exactly the pad-1 halo.  The box lands
//     in shared memory as 128 rows x 128 bytes with the 128-byte swizzle, i.e.
//     already in the canonical K-major UMMA operand layout -- no im2col buffer.
//   conv_wgrad_kernel  : dW[Cout, tap, Cin] = sum_pixels dY[p, Cout] * X[p+tap, Cin].
//     Both operands are consumed MN-major directly from NHWC (the contraction runs
//     over pixels), split-K over pixel blocks fills the 148 SMs, partial tiles are
//     combined with vector red.global.add.
//
// Warp roles per CTA (192 or 320 threads): warp 0 = TMA producer, warp 1 = TMEM owner +
// single-thread tcgen05.mma issuer, warps 2.. = epilogue, 4 or 8 of them (tcgen05.ld -> registers
// -> bias / BatchNorm statistics -> global).
//
// Replaces what the reference gets from ATen/oneDNN through nn.Conv2d
// (ref: master/part1/model.py:18-23) and its autograd (ref: master/part1/part1.py:37).
#include <cuda.h>
#include <cudaTypedefs.h>
#include <stdio.h>
#include <stdlib.h>

#include "../include/b200dp.h"
#include "common.cuh"
#include "host_util.h"

namespace b200dp {

constexpr int kBlockM = 128;
// A K block is one 128-byte SW128 swizzle span per row: 64 bf16 or 32 tf32 elements (Prec::kBlockK, common.cuh)
constexpr int kRowBytes = 128;
constexpr int kAStageBytes = kBlockM * kRowBytes;  // 16 KiB
// Warp roles: warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer, warps 2.. = epilogue.  A warp may only
// read the TMEM lane quadrant (warp % 4), so epilogue warps come in groups of four; tiles of 128 columns and
// more get TWO groups, each draining half of the columns (ncu on the one-group kernel: the epilogue was 22 %
// of a CTA's life with the tensor pipe idle behind it).
constexpr int epi_warps(int block_n) { return block_n >= 128 ? 8 : 4; }
constexpr int cta_threads(int block_n) { return 64 + 32 * epi_warps(block_n); }
// bytes after the stage ring: barriers + TMEM slot (256) and the CTA's bias slice (BLOCK_N floats)
constexpr int tail_bytes(int block_n) { return 256 + 4 * block_n; }

struct IgemmParams {
  int H, W, TH, TN;  // image dims; an M tile is TN images x TH rows x W columns = 128 pixels
  int taps;          // 9 (3x3 pad 1) or 1
  int cin_blocks;    // Cin / Prec::kBlockK
  int m_total;       // valid output rows (B*H*W)
  int n_total;       // Cout
  const float* bias;
  void* out;     // [M][Cout] of Prec::elem
  float* stats;  // optional [2][Cout]: sum, sum of squares
  // split-K: gridDim.z CTAs share one output tile, each reduces kb_per_split K blocks and adds its
  // fp32 partial tile into `partial` ([M][Cout], zero on entry); a finalize kernel produces `out`.
  float* partial;
  int kb_per_split;
  // 1: the four epilogue warps combine their statistics in shared memory (the drained stage-0 operand
  // buffer) and issue one atomic per channel per CTA instead of one per warp
  // dgrad only (conv_dgrad_bnr kernels): the tile just computed is dL/da for a = relu(bn(z)) of the
  // PREVIOUS block (max-pooled or not).  With `bnr_a` = that block's stored output, the epilogue also
  // accumulates its BatchNorm-backward sums  S1 = sum_{a>0} g,  S2 = sum_{a>0} g*(a-beta)/gamma  into
  // bnr_sums[0..N) / [N..2N) -- what bn_bwd_reduce_from_out_kernel computes in a separate pass.
  const void* bnr_a;   // [M][Cout] of Prec::elem
  const float* bnr_gamma;
  const float* bnr_beta;
  float* bnr_sums;
};

// Sum v[j] over the 32 lanes of the warp for every j; lane l returns the total of column l.
// 31 shuffles instead of 160 (butterfly that halves the live set each stage).
B200DP_DEVINL float warp_transpose_sum(float (&v)[32], int lane) {
#pragma unroll
  for (int off = 16, n = 32; off >= 1; off >>= 1, n >>= 1) {
    const bool upper = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < n / 2; ++i) {
      const float send = upper ? v[i] : v[i + n / 2];
      const float keep = upper ? v[i + n / 2] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return v[0];
}

B200DP_DEVINL void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

// 32 consecutive activations of one row, as they sit in memory: 4 x 16 bytes of bf16 or 8 x 16 bytes of fp32.
template <class Prec>
struct Row32 {
  uint4 v[Prec::kTf32 ? 8 : 4];
};
template <class Prec>
B200DP_DEVINL Row32<Prec> row32_load(const typename Prec::elem* p, bool pred) {
  Row32<Prec> r;
#pragma unroll
  for (int j = 0; j < (Prec::kTf32 ? 8 : 4); ++j)
    r.v[j] = pred ? __ldg(reinterpret_cast<const uint4*>(p) + j) : make_uint4(0u, 0u, 0u, 0u);
  return r;
}
template <class Prec>
B200DP_DEVINL void row32_unpack(const Row32<Prec>& r, float (&a)[32]) {
  if constexpr (Prec::kTf32) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      a[4 * j] = __uint_as_float(r.v[j].x); a[4 * j + 1] = __uint_as_float(r.v[j].y);
      a[4 * j + 2] = __uint_as_float(r.v[j].z); a[4 * j + 3] = __uint_as_float(r.v[j].w);
    }
  } else {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t w[4] = {r.v[j].x, r.v[j].y, r.v[j].z, r.v[j].w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 t = unpack_bf16x2(w[k]);
        a[8 * j + 2 * k] = t.x;
        a[8 * j + 2 * k + 1] = t.y;
      }
    }
  }
}
// Store 32 values of one row in the activation type; on return f[] holds the values AS STORED (bf16: rounded),
// so that statistics taken from f[] are exactly centred on what the next kernel reads.
template <class Prec>
B200DP_DEVINL void row32_store(typename Prec::elem* dst, bool pred, float (&f)[32]) {
  if constexpr (Prec::kTf32) {
    if (pred) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        reinterpret_cast<float4*>(dst)[j] = make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
    }
  } else {
    uint32_t packed[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      packed[j] = pack_bf16x2(f[2 * j], f[2 * j + 1]);
      const float2 r = unpack_bf16x2(packed[j]);
      f[2 * j] = r.x;
      f[2 * j + 1] = r.y;
    }
    if (pred) {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        reinterpret_cast<uint4*>(dst)[j] = make_uint4(packed[4 * j], packed[4 * j + 1], packed[4 * j + 2], packed[4 * j + 3]);
    }
  }
}

// Epilogue shared by the 1-CTA and 2-CTA kernels: TMEM -> registers -> (+bias, activation type, statistics) ->
// global, or fp32 red.add into the split-K workspace.
template <class Prec, int BLOCK_N, bool BNR = false>
B200DP_DEVINL void igemm_epilogue(const IgemmParams& p, uint32_t tmem_base, uint64_t* tmem_full_bar, int warp,
                                  int lane, int m_tile, int n_tile, float* scratch, float* bias_smem) {
  using elem = typename Prec::elem;
  // Warp w may only touch TMEM lanes [32*(w%4), 32*(w%4)+32); group (w-2)/4 takes its share of the columns.
  constexpr int kEpiThreads = 32 * epi_warps(BLOCK_N);
  constexpr int kColsPerGroup = BLOCK_N / (epi_warps(BLOCK_N) / 4);
  const int q = warp & 3;
  const int grp = (warp - 2) >> 2;
  const int row = q * 32 + lane;
  const long m = static_cast<long>(m_tile) * kBlockM + row;
  const bool valid = m < p.m_total;
  if (p.bias != nullptr && p.partial == nullptr) {
    // the bias slice of this tile goes to shared memory while the main loop runs (one L2 round trip per
    // 32-column chunk sat on the epilogue's critical path before)
    for (int i = static_cast<int>(threadIdx.x) - 64; i < BLOCK_N; i += kEpiThreads)
      bias_smem[i] = __ldg(p.bias + n_tile * BLOCK_N + i);
    asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");
  }
  mbar_wait(tmem_full_bar, 0);
  tc_fence_after();
  elem* out_row = static_cast<elem*>(p.out) + m * p.n_total + n_tile * BLOCK_N;
#pragma unroll 1
  for (int c = grp * kColsPerGroup; c < (grp + 1) * kColsPerGroup; c += 32) {
    uint32_t v[32];
    tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c, v);
    Row32<Prec> act;
    if constexpr (BNR) {
      // the previous block's output row, same [M][N] layout as the gradient being written; in flight
      // while the TMEM load completes
      act = row32_load<Prec>(static_cast<const elem*>(p.bnr_a) + m * p.n_total + n_tile * BLOCK_N + c,
                             valid && p.partial == nullptr);
    }
    tmem_ld_wait();
    if (p.partial != nullptr) {
      if (valid) {
        float* dst = p.partial + m * p.n_total + n_tile * BLOCK_N + c;
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          red_add_v4(dst + j, __uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                     __uint_as_float(v[j + 3]));
      }
      continue;
    }
    float f[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
    if (p.bias != nullptr) {
      const float4* b = reinterpret_cast<const float4*>(bias_smem + c);   // broadcast reads
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 bv = b[j];
        f[4 * j] += bv.x; f[4 * j + 1] += bv.y; f[4 * j + 2] += bv.z; f[4 * j + 3] += bv.w;
      }
    }
    row32_store<Prec>(out_row + c, valid, f);     // f[] now holds the stored (rounded) values
    if constexpr (BNR) {
      // T1 = sum_{a>0} g and T2 = sum_{a>0} g*a over the 32 rows of this warp; beta and gamma enter once
      // per channel afterwards: S2 = (T2 - beta*T1) / gamma.  g is the value as stored.
      float a[32], s2[32];
      row32_unpack<Prec>(act, a);
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        f[j] = (valid && a[j] > 0.f) ? f[j] : 0.f;
        s2[j] = f[j] * a[j];
      }
      const float t1 = warp_transpose_sum(f, lane);
      const float t2 = warp_transpose_sum(s2, lane);
      const int ch = n_tile * BLOCK_N + c + lane;
      const float gam = __ldg(p.bnr_gamma + ch);
      const float ig = fabsf(gam) > 1e-20f ? 1.f / gam : 0.f;
      const float u2 = (t2 - __ldg(p.bnr_beta + ch) * t1) * ig;
      scratch[q * 2 * BLOCK_N + c + lane] = t1;
      scratch[q * 2 * BLOCK_N + BLOCK_N + c + lane] = u2;
    } else if (p.stats != nullptr) {
      float s2[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        f[j] = valid ? f[j] : 0.f;
        s2[j] = f[j] * f[j];
      }
      const float t1 = warp_transpose_sum(f, lane);
      const float t2 = warp_transpose_sum(s2, lane);
      // combined across the epilogue warps in shared memory below: one atomic per channel per CTA.  scratch is
      // operand memory: every MMA that read it has completed (tmem_full_bar fired)
      scratch[q * 2 * BLOCK_N + c + lane] = t1;
      scratch[q * 2 * BLOCK_N + BLOCK_N + c + lane] = t2;
    }
  }
  if constexpr (BNR) {
    if (p.partial == nullptr) {
      asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");  // the epilogue warps only
      for (int i = static_cast<int>(threadIdx.x) - 64; i < 2 * BLOCK_N; i += kEpiThreads) {
        const float t = scratch[i] + scratch[2 * BLOCK_N + i] + scratch[4 * BLOCK_N + i] + scratch[6 * BLOCK_N + i];
        const int half = i / BLOCK_N;
        atomicAdd(p.bnr_sums + half * p.n_total + n_tile * BLOCK_N + (i - half * BLOCK_N), t);
      }
    }
  } else if (p.partial == nullptr && p.stats != nullptr) {
    asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");  // the epilogue warps only
    for (int i = static_cast<int>(threadIdx.x) - 64; i < 2 * BLOCK_N; i += kEpiThreads) {
      const float t = scratch[i] + scratch[2 * BLOCK_N + i] + scratch[4 * BLOCK_N + i] + scratch[6 * BLOCK_N + i];
      const int half = i / BLOCK_N;
      atomicAdd(p.stats + half * p.n_total + n_tile * BLOCK_N + (i - half * BLOCK_N), t);
    }
  }
  tc_fence_before();
}

// STAGES == 1 is the single-K-block case (layer 0: one 64-wide K block per tile).  Such a CTA is a pure
// latency chain (TMA -> one MMA group -> epilogue), so it is compiled for 5 CTAs per SM (25 KiB of
// shared memory, 64 TMEM columns each) instead of the 2 that a 4-stage ring allows.
template <class Prec, int BLOCK_N, int STAGES, bool BNR>
B200DP_DEVINL void conv_igemm_body(const CUtensorMap& tmA, const CUtensorMap& tmB, const IgemmParams& p) {
  constexpr int kBStageBytes = BLOCK_N * kRowBytes;
  constexpr int kStageBytes = kAStageBytes + kBStageBytes;
  constexpr uint32_t kTmemCols = BLOCK_N < 32 ? 32 : BLOCK_N;
  constexpr uint32_t kIdesc = umma_idesc(kBlockM, BLOCK_N, 0, 0, Prec::kFmt);

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * kStageBytes);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  pdl_launch_dependents();  // let the next kernel's CTAs start their prologue as SMs free up
  const int m_tile = blockIdx.x;
  const int n_tile = blockIdx.y;
  const int total_kb = p.taps * p.cin_blocks;
  const int kb_begin = blockIdx.z * p.kb_per_split;
  const int kb_end = min(kb_begin + p.kb_per_split, total_kb);
  const int num_kb = kb_end - kb_begin;  // >= 1 by construction of the grid

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full_bar, 1);
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // everything above overlapped the previous kernel's tail; global memory only from here on

  if (warp == 0) {
    if (elect_one()) {
      int n0, h0;
      if (p.TN == 1) {
        const int tiles_per_img = p.H / p.TH;
        n0 = m_tile / tiles_per_img;
        h0 = (m_tile - n0 * tiles_per_img) * p.TH;
      } else {
        n0 = m_tile * p.TN;
        h0 = 0;
      }
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        mbar_arrive_expect_tx(&full_bar[stage], kStageBytes);
        const int tap = kb / p.cin_blocks;
        const int cb = kb - tap * p.cin_blocks;
        int dr = 0, ds = 0;
        if (p.taps == 9) {
          dr = tap / 3 - 1;
          ds = tap % 3 - 1;
        }
        uint8_t* a_dst = smem + stage * kStageBytes;
        tma_load_4d(a_dst, &tmA, &full_bar[stage], cb * Prec::kBlockK, ds, h0 + dr, n0);
        tma_load_2d(a_dst + kAStageBytes, &tmB, &full_bar[stage], kb * Prec::kBlockK, n_tile * BLOCK_N);
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + stage * kStageBytes);
        const uint32_t b_addr = a_addr + kAStageBytes;
#pragma unroll
        for (int k = 0; k < Prec::kBlockK / Prec::kUmmaK; ++k) {
          // K-major SW128: 8-row groups are 1024 B apart; one MMA's K extent (16 bf16 / 8 tf32) is +32 B in the span.
          const uint64_t da = umma_smem_desc(a_addr + k * 32, 16, 1024);
          const uint64_t db = umma_smem_desc(b_addr + k * 32, 16, 1024);
          umma_ss<Prec>(tmem_base, da, db, kIdesc, (kb | k) != 0);
        }
        umma_commit(&empty_bar[stage]);
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
      umma_commit(tmem_full_bar);
    }
  } else {
    igemm_epilogue<Prec, BLOCK_N, BNR>(p, tmem_base, tmem_full_bar, warp, lane, m_tile, n_tile, reinterpret_cast<float*>(smem),
                                 reinterpret_cast<float*>(smem + STAGES * kStageBytes + 256));
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

template <class Prec, int BLOCK_N, int STAGES>
__global__ void __launch_bounds__(cta_threads(BLOCK_N), STAGES == 1 ? 5 : 1)
    conv_igemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                      const IgemmParams p) {
  conv_igemm_body<Prec, BLOCK_N, STAGES, false>(tmA, tmB, p);
}

// dgrad whose epilogue also accumulates the previous block's BatchNorm-backward sums (IgemmParams::bnr_*)
template <class Prec, int BLOCK_N, int STAGES>
__global__ void __launch_bounds__(cta_threads(BLOCK_N))
    conv_dgrad_bnr_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                          const IgemmParams p) {
  conv_igemm_body<Prec, BLOCK_N, STAGES, true>(tmA, tmB, p);
}

// ---------------------------------------------------------------------------
// CTA-pair variant: tcgen05.mma.cta_group::2, one 256(M) x BLOCK_N tile per pair of SMs.
// Each CTA stages its own 128 pixel rows of A and only HALF of the weight tile (BLOCK_N/2 rows);
// the tensor core reads the other half from the partner SM.  Per k-block an SM therefore ingests
// 16 KiB + BLOCK_N*64 B instead of 16 KiB + BLOCK_N*128 B -- these shapes are bound by L2->SM
// delivery (ncu: xbar2l1tex ~8 TB/s, tensor pipe ~40%), so fewer bytes per FLOP is the lever.
// Protocol: both CTAs run a TMA producer whose loads complete on the LEADER's full barrier
// (expect_tx counts both halves); the leader's single MMA thread issues for the pair and commits
// with a cluster multicast to the empty / tmem-full barriers of both CTAs; each CTA drains its own
// TMEM half in its epilogue warps.
// ---------------------------------------------------------------------------
template <class Prec, int BLOCK_N, int STAGES, bool BNR>
B200DP_DEVINL void conv_igemm_2cta_body(const CUtensorMap& tmA, const CUtensorMap& tmB, const IgemmParams& p) {
  constexpr int kBHalfBytes = (BLOCK_N / 2) * kRowBytes;
  constexpr int kStageBytes = kAStageBytes + kBHalfBytes;
  constexpr uint32_t kTmemCols = BLOCK_N < 32 ? 32 : BLOCK_N;
  constexpr uint32_t kIdesc = umma_idesc(2 * kBlockM, BLOCK_N, 0, 0, Prec::kFmt);

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * kStageBytes);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  pdl_launch_dependents();  // let the next kernel's CTAs start their prologue as SMs free up
  const uint32_t cta_rank = cluster_ctarank();
  const bool leader = cta_rank == 0;
  const int m_tile = blockIdx.x;  // the pair owns m tiles (2i, 2i+1)
  const int n_tile = blockIdx.y;
  const int total_kb = p.taps * p.cin_blocks;
  const int kb_begin = blockIdx.z * p.kb_per_split;
  const int kb_end = min(kb_begin + p.kb_per_split, total_kb);
  const int num_kb = kb_end - kb_begin;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);   // leader: its own arrive.expect_tx; tx bytes of both CTAs
      mbar_init(&empty_bar[s], 1);  // one multicast commit from the leader's MMA thread
    }
    mbar_init(tmem_full_bar, 1);
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc_2cta(tmem_slot, kTmemCols);
    tmem_relinquish_2cta();
  }
  tc_fence_before();
  cluster_sync_all();  // the partner's barriers must be initialised before anything can arrive on them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // everything above overlapped the previous kernel's tail; global memory only from here on

  if (warp == 0) {
    if (elect_one()) {
      int n0, h0;
      if (p.TN == 1) {
        const int tiles_per_img = p.H / p.TH;
        n0 = m_tile / tiles_per_img;
        h0 = (m_tile - n0 * tiles_per_img) * p.TH;
      } else {
        n0 = m_tile * p.TN;
        h0 = 0;
      }
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2 * kStageBytes);
        const int tap = kb / p.cin_blocks;
        const int cb = kb - tap * p.cin_blocks;
        int dr = 0, ds = 0;
        if (p.taps == 9) {
          dr = tap / 3 - 1;
          ds = tap % 3 - 1;
        }
        uint8_t* a_dst = smem + stage * kStageBytes;
        tma_load_4d_2cta(a_dst, &tmA, &full_bar[stage], cb * Prec::kBlockK, ds, h0 + dr, n0);
        tma_load_2d_2cta(a_dst + kAStageBytes, &tmB, &full_bar[stage], kb * Prec::kBlockK,
                         n_tile * BLOCK_N + static_cast<int>(cta_rank) * (BLOCK_N / 2));
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    if (leader && elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + stage * kStageBytes);
        const uint32_t b_addr = a_addr + kAStageBytes;
#pragma unroll
        for (int k = 0; k < Prec::kBlockK / Prec::kUmmaK; ++k) {
          const uint64_t da = umma_smem_desc(a_addr + k * 32, 16, 1024);
          const uint64_t db = umma_smem_desc(b_addr + k * 32, 16, 1024);
          umma_ss_2cta<Prec>(tmem_base, da, db, kIdesc, (kb | k) != 0);
        }
        umma_commit_2cta(&empty_bar[stage], 0b11);
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
      umma_commit_2cta(tmem_full_bar, 0b11);
    }
  } else {
    igemm_epilogue<Prec, BLOCK_N, BNR>(p, tmem_base, tmem_full_bar, warp, lane, m_tile, n_tile, reinterpret_cast<float*>(smem),
                                 reinterpret_cast<float*>(smem + STAGES * kStageBytes + 256));
  }
  __syncthreads();
  tc_fence_before();
  cluster_sync_all();  // neither CTA may free TMEM or exit while the pair's MMAs / arrives are in flight
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_2cta(tmem_base, kTmemCols);
  }
}

template <class Prec, int BLOCK_N, int STAGES>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(cta_threads(BLOCK_N))
    conv_igemm_2cta_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                           const IgemmParams p) {
  conv_igemm_2cta_body<Prec, BLOCK_N, STAGES, false>(tmA, tmB, p);
}

template <class Prec, int BLOCK_N, int STAGES>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(cta_threads(BLOCK_N))
    conv_dgrad_bnr_2cta_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                               const IgemmParams p) {
  conv_igemm_2cta_body<Prec, BLOCK_N, STAGES, true>(tmA, tmB, p);
}

// Split-K epilogue: partial fp32 [M][N] -> (+bias) -> activation type, fused BatchNorm statistics of the
// stored values, and the workspace is cleared again for its next user.
template <class T>
__global__ void __launch_bounds__(256) splitk_finalize_kernel(float* __restrict__ partial,
                                                              const float* __restrict__ bias, T* __restrict__ out,
                                                              float* __restrict__ stats, long M, int N) {
  extern __shared__ float red[];  // [rows_per_block][N] x 2 (only with stats)
  pdl_launch_dependents();
  pdl_wait();
  const int chunks = N / 8;
  const int rows_per_block = blockDim.x / chunks;
  const int chunk = threadIdx.x % chunks, rl = threadIdx.x / chunks, c0 = chunk * 8;
  float b[8], s1[8], s2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    b[j] = bias != nullptr ? bias[c0 + j] : 0.f;
    s1[j] = s2[j] = 0.f;
  }
  if (rl < rows_per_block) {
    for (long r = static_cast<long>(blockIdx.x) * rows_per_block + rl; r < M;
         r += static_cast<long>(gridDim.x) * rows_per_block) {
      float4* src = reinterpret_cast<float4*>(partial + r * N + c0);
      const float4 lo = src[0], hi = src[1];
      src[0] = make_float4(0.f, 0.f, 0.f, 0.f);
      src[1] = make_float4(0.f, 0.f, 0.f, 0.f);
      F8 o = {{lo.x + b[0], lo.y + b[1], lo.z + b[2], lo.w + b[3], hi.x + b[4], hi.y + b[5], hi.z + b[6], hi.w + b[7]}};
      st8(out + r * N + c0, o);          // o now holds the values as stored
      if (stats != nullptr) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          s1[j] += o.v[j];
          s2[j] = fmaf(o.v[j], o.v[j], s2[j]);
        }
      }
    }
  }
  if (stats != nullptr) {
    float* r1 = red;
    float* r2 = red + rows_per_block * N;
    if (rl < rows_per_block) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        r1[rl * N + c0 + j] = s1[j];
        r2[rl * N + c0 + j] = s2[j];
      }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < N; c += blockDim.x) {
      float a = 0.f, t = 0.f;
      for (int r = 0; r < rows_per_block; ++r) {
        a += r1[r * N + c];
        t += r2[r * N + c];
      }
      atomicAdd(stats + c, a);
      atomicAdd(stats + N + c, t);
    }
  }
}

// Split-K epilogue of a dgrad that also feeds the previous block's BatchNorm backward: partial fp32
// [M][N] -> dL/da in the activation type, workspace cleared, and S1 / S2 of IgemmParams::bnr_* from the
// stored values.
template <class T>
__global__ void __launch_bounds__(256) splitk_finalize_bnr_kernel(float* __restrict__ partial, T* __restrict__ out,
                                                                  const T* __restrict__ act,
                                                                  const float* __restrict__ gamma,
                                                                  const float* __restrict__ beta,
                                                                  float* __restrict__ sums, long M, int N) {
  extern __shared__ float red[];  // [rows_per_block][N] x 2
  pdl_launch_dependents();
  pdl_wait();
  const int chunks = N / 8;
  const int rows_per_block = blockDim.x / chunks;
  const int chunk = threadIdx.x % chunks, rl = threadIdx.x / chunks, c0 = chunk * 8;
  float s1[8], s2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s1[j] = s2[j] = 0.f;
  if (rl < rows_per_block) {
    for (long r = static_cast<long>(blockIdx.x) * rows_per_block + rl; r < M;
         r += static_cast<long>(gridDim.x) * rows_per_block) {
      float4* src = reinterpret_cast<float4*>(partial + r * N + c0);
      const float4 lo = src[0], hi = src[1];
      const F8 av = ld8(act + r * N + c0);
      src[0] = make_float4(0.f, 0.f, 0.f, 0.f);
      src[1] = make_float4(0.f, 0.f, 0.f, 0.f);
      F8 g = {{lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w}};
      st8(out + r * N + c0, g);          // g now holds the values as stored
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float d = av.v[j] > 0.f ? g.v[j] : 0.f;
        s1[j] += d;
        s2[j] = fmaf(d, av.v[j], s2[j]);
      }
    }
  }
  float* r1 = red;
  float* r2 = red + rows_per_block * N;
  if (rl < rows_per_block) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      r1[rl * N + c0 + j] = s1[j];
      r2[rl * N + c0 + j] = s2[j];
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < N; c += blockDim.x) {
    float t1 = 0.f, t2 = 0.f;
    for (int r = 0; r < rows_per_block; ++r) {
      t1 += r1[r * N + c];
      t2 += r2[r * N + c];
    }
    const float gam = gamma[c];
    const float ig = fabsf(gam) > 1e-20f ? 1.f / gam : 0.f;
    atomicAdd(sums + c, t1);
    atomicAdd(sums + N + c, (t2 - beta[c] * t1) * ig);
  }
}

// ---------------------------------------------------------------------------
// weight gradient
// ---------------------------------------------------------------------------
struct WgradParams {
  int H, W, TH, TN;  // a K block is TN images x TH rows x W columns = 64 pixels (bf16) or 32 pixels (tf32)
  int taps;
  int cin, cout;
  int co_tiles, ci_tiles;
  int a_boxes;  // dY boxes (Prec::kBlockK channels each) that really exist in this M tile: min(Cout, 128) / kBlockK
  int kblocks_total, kblocks_per_split;
  int ld_out, cin_valid;
  int use_red;
  float* out;
};

struct WgradPushParams : WgradParams {
  // conv_wgrad_push kernels only (GEMM -> reduce-scatter in one kernel): when the last of the `splits`
  // CTAs of an output tile has added its partial, that CTA forwards the finished tile to the rank that
  // owns it, with plain vector stores over NVLink into the owner's staging slot for this rank.  The
  // owner later sums its own gradient and the W-1 staged copies inside the fused SGD kernel -- no
  // cross-GPU atomics, no reduce-scatter phase in the collective.
  unsigned int* tile_counters;  // one per output tile (gridDim.x), zero on entry, left zero
  int splits;
  int world, rank;
  long long flat_off;    // element offset of out[0] in the flat gradient buffer (multiple of 4)
  long long own_start;   // first element of the bucket this tensor belongs to (multiple of 4)
  long long per4;        // float4s of the bucket owned by each rank: owner(e) = ((e - own_start) / 4) / per4
  long long stage_base;  // element offset of the bucket's staging area inside `peer_stage[r]`
  float* peer_stage[8];  // every rank's staging buffer: slot of source rank s starts at stage_base + s*per4*4
};

// pixels (GEMM K) per pipeline stage of the weight-gradient kernel: 64 for bf16, 32 for tf32 -- the same bytes
constexpr int wgrad_pixels(bool tf32) { return tf32 ? 32 : 64; }

template <class Prec, int BLOCK_N, int STAGES, bool PUSH, typename ParamsT>
B200DP_DEVINL void conv_wgrad_body(const CUtensorMap& tmDY, const CUtensorMap& tmX, const ParamsT& p) {
  constexpr int kPix = wgrad_pixels(Prec::kTf32);
  constexpr int kCh = Prec::kBlockK;                  // channels per box = one 128-byte row
  constexpr int kSubTileBytes = kPix * kRowBytes;     // one (kCh channels x kPix pixels) box: 8 KiB / 4 KiB
  constexpr int kABoxes = kBlockM / kCh;
  constexpr int kABytes = kABoxes * kSubTileBytes;    // 16 KiB
  constexpr int kBBytes = (BLOCK_N / kCh) * kSubTileBytes;
  constexpr int kStageBytes = kABytes + kBBytes;
  constexpr uint32_t kTmemCols = BLOCK_N < 32 ? 32 : BLOCK_N;
  constexpr uint32_t kIdesc = umma_idesc(kBlockM, BLOCK_N, 1, 1, Prec::kFmt);

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * kStageBytes);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  pdl_launch_dependents();  // let the next kernel's CTAs start their prologue as SMs free up

  int t = blockIdx.x;
  const int ci_tile = t % p.ci_tiles;
  t /= p.ci_tiles;
  const int co_tile = t % p.co_tiles;
  const int tap = t / p.co_tiles;
  const int kb_begin = blockIdx.y * p.kblocks_per_split;
  const int kb_end = min(kb_begin + p.kblocks_per_split, p.kblocks_total);
  const int num_kb = kb_end - kb_begin;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmDY);
    tma_prefetch_desc(&tmX);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full_bar, 1);
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // everything above overlapped the previous kernel's tail; global memory only from here on

  if (num_kb > 0) {
    if (warp == 0) {
      if (elect_one()) {
        int dr = 0, ds = 0;
        if (p.taps == 9) {
          dr = tap / 3 - 1;
          ds = tap % 3 - 1;
        }
        const int blocks_per_img = (p.TN == 1) ? (p.H / p.TH) : 1;
        int stage = 0;
        uint32_t phase = 0;
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full_bar[stage], kStageBytes);
          int n0, h0;
          if (p.TN == 1) {
            n0 = kb / blocks_per_img;
            h0 = (kb - n0 * blocks_per_img) * p.TH;
          } else {
            n0 = kb * p.TN;
            h0 = 0;
          }
          uint8_t* a_dst = smem + stage * kStageBytes;
          // dY tile: [kPix pixels][kCh channels] per box, pixels kb*kPix .. (boxes past Cout repeat box 0:
          // their output rows are never stored)
#pragma unroll
          for (int j = 0; j < kABoxes; ++j) {
            const int co = co_tile * kBlockM + (j < p.a_boxes ? j * kCh : 0);
            tma_load_2d(a_dst + j * kSubTileBytes, &tmDY, &full_bar[stage], co, kb * kPix);
          }
          uint8_t* b_dst = a_dst + kABytes;
#pragma unroll
          for (int j = 0; j < BLOCK_N / kCh; ++j)
            tma_load_4d(b_dst + j * kSubTileBytes, &tmX, &full_bar[stage], ci_tile * BLOCK_N + j * kCh, ds,
                        h0 + dr, n0);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    } else if (warp == 1) {
      if (elect_one()) {
        int stage = 0;
        uint32_t phase = 0;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + stage * kStageBytes);
          const uint32_t b_addr = a_addr + kABytes;
#pragma unroll
          for (int k = 0; k < kPix / Prec::kUmmaK; ++k) {
            // MN-major: rows are pixels (K), 128 B each; column blocks of one 128-byte row (64 bf16 / 32 tf32
            // channels) are one sub-tile apart (LBO); the pixels of one MMA (16 / 8) are +2048 / +1024 B.
            // bf16: SWIZZLE_128B, swizzle atoms of 8 pixel rows, 1024 B apart (SBO).  tf32: MN-major 32-bit
            // operands only exist as SWIZZLE_128B_BASE32B, whose atoms are 4 pixel rows = 512 B (SBO).
            constexpr uint32_t kSbo = Prec::kTf32 ? 512 : 1024;
            constexpr uint32_t kLayout = Prec::kTf32 ? 1 : 2;
            const uint64_t da = umma_smem_desc(a_addr + k * Prec::kUmmaK * kRowBytes, kSubTileBytes, kSbo, kLayout);
            const uint64_t db = umma_smem_desc(b_addr + k * Prec::kUmmaK * kRowBytes, kSubTileBytes, kSbo, kLayout);
            umma_ss<Prec>(tmem_base, da, db, kIdesc, (kb | k) != 0);
          }
          umma_commit(&empty_bar[stage]);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(tmem_full_bar);
      }
    } else {
      constexpr int kEpiThreads = 32 * epi_warps(BLOCK_N);
      constexpr int kGroups = epi_warps(BLOCK_N) / 4;
      constexpr int kColsPerGroup = BLOCK_N / kGroups;
      const int q = warp & 3;
      const int grp = (warp - 2) >> 2;
      const int row = q * 32 + lane;
      const int co = co_tile * kBlockM + row;
      const bool valid = co < p.cout;
      mbar_wait(tmem_full_bar, 0);
      tc_fence_after();
      float* out_row = p.out + static_cast<long>(co) * p.ld_out + tap * p.cin_valid + ci_tile * BLOCK_N;
      const int col_limit = p.cin_valid - ci_tile * BLOCK_N;  // columns >= this are padding
#pragma unroll 1
      for (int c = grp * kColsPerGroup; c < (grp + 1) * kColsPerGroup; c += 32) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c, v);
        tmem_ld_wait();
        if (valid) {
          if (c + 32 <= col_limit && (p.ld_out & 3) == 0 && (p.cin_valid & 3) == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              float* dst = out_row + c + j;
              if (p.use_red) {
                red_add_v4(dst, __uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                           __uint_as_float(v[j + 3]));
              } else {
                *reinterpret_cast<float4*>(dst) =
                    make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                                __uint_as_float(v[j + 3]));
              }
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              if (c + j < col_limit) {
                if (p.use_red)
                  atomicAdd(out_row + c + j, __uint_as_float(v[j]));
                else
                  out_row[c + j] = __uint_as_float(v[j]);
              }
            }
          }
        }
      }
      if constexpr (PUSH) {
        // Tile hand-off.  Every split has added (or stored) its partial above; the fence orders those
        // writes before the counter increment, so whoever draws the last ticket sees the complete tile.
        uint32_t* last_flag = tmem_slot + 1;
        __threadfence();
        asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");
        if (threadIdx.x == 64) {
          const unsigned int ticket = atomicAdd(p.tile_counters + blockIdx.x, 1u);
          const bool last = ticket == static_cast<unsigned int>(p.splits) - 1u;
          if (last) p.tile_counters[blockIdx.x] = 0u;  // nobody else touches it before the next launch
          *last_flag = last ? 1u : 0u;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");
        if (*last_flag != 0u) {
          __threadfence();
          // warp-cooperative: the 32 lanes walk one output row at a time, so every store instruction
          // sends 512 contiguous bytes over NVLink
          const int ncols4 = min(BLOCK_N, col_limit) >> 2;  // the host enables PUSH only for multiples of 4
          for (int rr = 0; rr < 32; ++rr) {
            const int co_r = co_tile * kBlockM + q * 32 + rr;
            if (co_r >= p.cout) break;
            const long long row_off = static_cast<long long>(co_r) * p.ld_out + tap * p.cin_valid + ci_tile * BLOCK_N;
            const float* src_row = p.out + row_off;
            const long long i4_row = (p.flat_off + row_off - p.own_start) >> 2;
            for (int c4 = lane + 32 * grp; c4 < ncols4; c4 += 32 * kGroups) {
              const long long i4 = i4_row + c4;
              const int owner = static_cast<int>(i4 / p.per4);
              if (owner != p.rank) {
                const float4 v = __ldcg(reinterpret_cast<const float4*>(src_row) + c4);
                float* dst = p.peer_stage[owner] + p.stage_base +
                             ((static_cast<long long>(p.rank) * p.per4 + (i4 - owner * p.per4)) << 2);
                *reinterpret_cast<float4*>(dst) = v;
              }
            }
          }
        }
      }
      tc_fence_before();
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

template <class Prec, int BLOCK_N, int STAGES>
__global__ void __launch_bounds__(cta_threads(BLOCK_N))
    conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmX,
                      const WgradParams p) {
  conv_wgrad_body<Prec, BLOCK_N, STAGES, false, WgradParams>(tmDY, tmX, p);
}

// Weight gradient + the transfer half of the gradient reduce-scatter (WgradParams::peer_stage ...).
template <int BLOCK_N, int STAGES>
__global__ void __launch_bounds__(cta_threads(BLOCK_N))
    conv_wgrad_push_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmX,
                           const WgradPushParams p) {
  conv_wgrad_body<PrecBf16, BLOCK_N, STAGES, true, WgradPushParams>(tmDY, tmX, p);
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
static PFN_cuTensorMapEncodeTiled_v12000 get_encode_fn() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  if (fn == nullptr) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || ptr == nullptr) {
      set_error("cuTensorMapEncodeTiled is not available from the CUDA driver (%s)", cudaGetErrorString(e));
      return nullptr;
    }
    fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(ptr);
  }
  return fn;
}

// Tensor map over 2-byte (bf16) or 4-byte elements; 4-byte maps use the TFLOAT32 type, i.e. the TMA unit rounds the
// fp32 data to tf32 (round to nearest) on its way into shared memory.  128-byte swizzle, out-of-bounds = zero.
// atom32: the 128-byte swizzle with 32-byte atoms (MN-major tf32 operands of the weight-gradient kernel).
static int make_tmap(CUtensorMap* tm, const void* base, int rank, const uint64_t* dims, const uint32_t* box,
                     int elem_bytes = 2, bool atom32 = false) {
  auto fn = get_encode_fn();
  if (fn == nullptr) return B200DP_ENOSYS;
  cuuint64_t gdim[4];
  cuuint64_t gstride[3];
  cuuint32_t bdim[4];
  cuuint32_t estr[4] = {1, 1, 1, 1};
  uint64_t stride = static_cast<uint64_t>(elem_bytes);
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bdim[i] = box[i];
    stride *= dims[i];
    if (i < rank - 1) gstride[i] = stride;
  }
  CUresult r = fn(tm, elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_TFLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), gdim, gstride, bdim, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %d dims %llu %llu box %u %u)", (int)r, rank,
              (unsigned long long)dims[0], (unsigned long long)dims[1], box[0], box[1]);
    return B200DP_EINVAL;
  }
  return 0;
}

// Split `pixels_per_tile` output pixels into TN images x TH rows x W columns.
static int tile_geometry(int H, int W, int pixels_per_tile, int* TH, int* TN) {
  if (W <= 0 || H <= 0 || pixels_per_tile % W != 0) return -1;
  const int rows = pixels_per_tile / W;
  if (rows <= H) {
    if (H % rows != 0) return -1;
    *TH = rows;
    *TN = 1;
  } else {
    if (rows % H != 0) return -1;
    *TH = H;
    *TN = rows / H;
  }
  return 0;
}

template <class Prec, int BLOCK_N, int STAGES>
static int launch_igemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const IgemmParams& p, int m_tiles,
                        int splits, cudaStream_t stream) {
  constexpr int smem = STAGES * (kAStageBytes + BLOCK_N * kRowBytes) + 1024 + tail_bytes(BLOCK_N);
  static bool configured = false;
  auto kern = conv_igemm_kernel<Prec, BLOCK_N, STAGES>;
  if (!configured) {
    B200DP_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  dim3 grid(m_tiles, p.n_total / BLOCK_N, splits);
  B200DP_CUDA_TRY(launch_pdl(kern, grid, dim3(cta_threads(BLOCK_N)), smem, stream, tmA, tmB, p));
  return 0;
}

template <class Prec, int BLOCK_N, int STAGES>
static int launch_igemm_2cta(const CUtensorMap& tmA, const CUtensorMap& tmB, const IgemmParams& p, int m_tiles,
                             int splits, cudaStream_t stream) {
  constexpr int smem = STAGES * (kAStageBytes + (BLOCK_N / 2) * kRowBytes) + 1024 + tail_bytes(BLOCK_N);
  static bool configured = false;
  auto kern = conv_igemm_2cta_kernel<Prec, BLOCK_N, STAGES>;
  if (!configured) {
    B200DP_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  dim3 grid(m_tiles, p.n_total / BLOCK_N, splits);  // m_tiles is even; __cluster_dims__(2,1,1) pairs them
  B200DP_CUDA_TRY(launch_pdl(kern, grid, dim3(cta_threads(BLOCK_N)), smem, stream, tmA, tmB, p));
  return 0;
}

template <class Prec, int BLOCK_N, int STAGES>
static int launch_dgrad_bnr(const CUtensorMap& tmA, const CUtensorMap& tmB, const IgemmParams& p, int m_tiles,
                            int splits, cudaStream_t stream) {
  constexpr int smem = STAGES * (kAStageBytes + BLOCK_N * kRowBytes) + 1024 + tail_bytes(BLOCK_N);
  static bool configured = false;
  auto kern = conv_dgrad_bnr_kernel<Prec, BLOCK_N, STAGES>;
  if (!configured) {
    B200DP_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  dim3 grid(m_tiles, p.n_total / BLOCK_N, splits);
  B200DP_CUDA_TRY(launch_pdl(kern, grid, dim3(cta_threads(BLOCK_N)), smem, stream, tmA, tmB, p));
  return 0;
}

template <class Prec, int BLOCK_N, int STAGES>
static int launch_dgrad_bnr_2cta(const CUtensorMap& tmA, const CUtensorMap& tmB, const IgemmParams& p, int m_tiles,
                                 int splits, cudaStream_t stream) {
  constexpr int smem = STAGES * (kAStageBytes + (BLOCK_N / 2) * kRowBytes) + 1024 + tail_bytes(BLOCK_N);
  static bool configured = false;
  auto kern = conv_dgrad_bnr_2cta_kernel<Prec, BLOCK_N, STAGES>;
  if (!configured) {
    B200DP_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  dim3 grid(m_tiles, p.n_total / BLOCK_N, splits);
  B200DP_CUDA_TRY(launch_pdl(kern, grid, dim3(cta_threads(BLOCK_N)), smem, stream, tmA, tmB, p));
  return 0;
}

template <class Prec, int BLOCK_N, int STAGES>
static int launch_wgrad(const CUtensorMap& tmDY, const CUtensorMap& tmX, const WgradParams& p, int tiles,
                        int splits, cudaStream_t stream) {
  constexpr int smem = STAGES * (kBlockM + BLOCK_N) * kRowBytes + 1024 + tail_bytes(BLOCK_N);
  static bool configured = false;
  auto kern = conv_wgrad_kernel<Prec, BLOCK_N, STAGES>;
  if (!configured) {
    B200DP_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  dim3 grid(tiles, splits);
  B200DP_CUDA_TRY(launch_pdl(kern, grid, dim3(cta_threads(BLOCK_N)), smem, stream, tmDY, tmX, p));
  return 0;
}

template <int BLOCK_N, int STAGES>
static int launch_wgrad_push(const CUtensorMap& tmDY, const CUtensorMap& tmX, const WgradPushParams& p, int tiles,
                             int splits, cudaStream_t stream) {
  constexpr int smem = STAGES * (kBlockM + BLOCK_N) * kRowBytes + 1024 + tail_bytes(BLOCK_N);
  static bool configured = false;
  auto kern = conv_wgrad_push_kernel<BLOCK_N, STAGES>;
  if (!configured) {
    B200DP_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  dim3 grid(tiles, splits);
  B200DP_CUDA_TRY(launch_pdl(kern, grid, dim3(cta_threads(BLOCK_N)), smem, stream, tmDY, tmX, p));
  return 0;
}

// What the push flavour of the weight-gradient kernel needs to know about the data-parallel group.
struct WgradPushArgs {
  unsigned int* tile_counters;
  int n_counters;
  int world, rank;
  void* const* peer_stage;  // `world` device pointers: every rank's staging buffer
  long long flat_off, own_start, own_len, stage_base;
};

}  // namespace b200dp

using namespace b200dp;

// Optional second job of a dgrad launch: BatchNorm-backward sums of the block that produced `act_out`.
struct BnReduceArgs {
  const void* act_out;
  const float* gamma;
  const float* beta;
  float* sums;
};

template <class Prec>
static int conv_igemm_impl(const void* act, const void* wgt, const float* bias, void* out, float* stats,
                           float* workspace, int B, int H, int W, int Cin, int Cout, int taps, int block_n, int split_k,
                           int cta_pair, cudaStream_t stream, const BnReduceArgs* bnr = nullptr) {
  using elem = typename Prec::elem;
  if ((taps != 9 && taps != 1) || Cin % Prec::kBlockK != 0 || Cout % 32 != 0 || B <= 0) {
    set_error("conv_igemm: unsupported shape B=%d H=%d W=%d Cin=%d Cout=%d taps=%d (Cin must be a multiple of %d)", B,
              H, W, Cin, Cout, taps, Prec::kBlockK);
    return B200DP_EINVAL;
  }
  IgemmParams p{};
  p.H = H;
  p.W = W;
  if (tile_geometry(H, W, kBlockM, &p.TH, &p.TN) != 0) {
    set_error("conv_igemm: H=%d W=%d cannot be tiled into 128-pixel blocks", H, W);
    return B200DP_EINVAL;
  }
  p.taps = taps;
  p.cin_blocks = Cin / Prec::kBlockK;
  p.m_total = B * H * W;
  p.n_total = Cout;
  p.bias = bias;
  p.out = out;
  p.stats = stats;
  const int m_tiles = (p.m_total + kBlockM - 1) / kBlockM;
  if (block_n == 0) {
    // largest N tile that still yields about one CTA per SM
    const int cands[4] = {256, 128, 64, 32};
    block_n = 32;
    for (int c : cands) {
      if (Cout % c == 0 && static_cast<long>(m_tiles) * (Cout / c) >= 120) {
        block_n = c;
        break;
      }
    }
    if (Cout % block_n != 0) block_n = (Cout % 64 == 0) ? 64 : 32;
  }
  if (Cout % block_n != 0) {
    set_error("conv_igemm: Cout=%d is not a multiple of block_n=%d", Cout, block_n);
    return B200DP_EINVAL;
  }
  const int total_kb = taps * p.cin_blocks;
  if (split_k < 1) split_k = 1;
  if (split_k > total_kb) split_k = total_kb;
  p.kb_per_split = (total_kb + split_k - 1) / split_k;
  split_k = (total_kb + p.kb_per_split - 1) / p.kb_per_split;
  if (split_k > 1) {
    if (workspace == nullptr || Cout % 8 != 0 || Cout > 2048) {
      set_error("conv_igemm: split_k=%d needs an fp32 workspace of B*H*W*Cout floats (zeroed) and Cout <= 2048",
                split_k);
      return B200DP_EINVAL;
    }
    p.partial = workspace;
  }
  CUtensorMap tmA, tmB;
  {
    const uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)B};
    const uint32_t box[4] = {(uint32_t)Prec::kBlockK, (uint32_t)W, (uint32_t)p.TH, (uint32_t)p.TN};
    int rc = make_tmap(&tmA, act, 4, dims, box, Prec::kBytes);
    if (rc) return rc;
  }
  if (cta_pair == 1 && ((m_tiles & 1) || block_n < 64)) {
    set_error("conv_igemm: the CTA-pair kernel needs an even number of 128-pixel tiles (got %d) and block_n >= 64",
              m_tiles);
    return B200DP_EINVAL;
  }
  {
    const uint64_t dims[2] = {(uint64_t)taps * Cin, (uint64_t)Cout};
    const uint32_t box[2] = {(uint32_t)Prec::kBlockK, (uint32_t)(cta_pair == 1 ? block_n / 2 : block_n)};
    int rc = make_tmap(&tmB, wgt, 2, dims, box, Prec::kBytes);
    if (rc) return rc;
  }
  int rc;
  if (bnr != nullptr) {
    if (bias != nullptr || stats != nullptr || Cout % 32 != 0) {
      set_error("conv_dgrad_bn_reduce: bias / forward statistics cannot be combined with the BatchNorm-backward sums");
      return B200DP_EINVAL;
    }
    p.bnr_a = bnr->act_out;
    p.bnr_gamma = bnr->gamma;
    p.bnr_beta = bnr->beta;
    p.bnr_sums = bnr->sums;
    if (cta_pair == 1) {
      switch (block_n) {
        case 256: rc = launch_dgrad_bnr_2cta<Prec, 256, 6>(tmA, tmB, p, m_tiles, split_k, stream); break;
        case 128: rc = launch_dgrad_bnr_2cta<Prec, 128, 6>(tmA, tmB, p, m_tiles, split_k, stream); break;
        case 64: rc = launch_dgrad_bnr_2cta<Prec, 64, 6>(tmA, tmB, p, m_tiles, split_k, stream); break;
        default: set_error("conv_dgrad_bn_reduce: block_n=%d not in {64,128,256} for the CTA-pair kernel", block_n); return B200DP_EINVAL;
      }
    } else {
      switch (block_n) {
        case 256: rc = launch_dgrad_bnr<Prec, 256, 4>(tmA, tmB, p, m_tiles, split_k, stream); break;
        case 128: rc = launch_dgrad_bnr<Prec, 128, 3>(tmA, tmB, p, m_tiles, split_k, stream); break;
        case 64: rc = launch_dgrad_bnr<Prec, 64, 4>(tmA, tmB, p, m_tiles, split_k, stream); break;
        case 32: rc = launch_dgrad_bnr<Prec, 32, 4>(tmA, tmB, p, m_tiles, split_k, stream); break;
        default: set_error("conv_dgrad_bn_reduce: block_n=%d not in {32,64,128,256}", block_n); return B200DP_EINVAL;
      }
    }
    if (rc || split_k == 1) return rc;
    const int threads = 256;
    const int rows_per_block = threads / (Cout / 8);
    const size_t smem = 2 * rows_per_block * Cout * sizeof(float);
    long blocks = (static_cast<long>(p.m_total) + rows_per_block * 2 - 1) / (rows_per_block * 2);
    if (blocks > 148 * 4) blocks = 148 * 4;
    if (blocks < 1) blocks = 1;
    B200DP_CUDA_TRY(launch_pdl(splitk_finalize_bnr_kernel<elem>, dim3(static_cast<unsigned>(blocks)), dim3(threads),
                               smem, stream, workspace, static_cast<elem*>(out), static_cast<const elem*>(p.bnr_a),
                               p.bnr_gamma, p.bnr_beta, p.bnr_sums, static_cast<long>(p.m_total), Cout));
    return 0;
  }
  if (cta_pair == 1) {
    switch (block_n) {
      case 256: rc = launch_igemm_2cta<Prec, 256, 6>(tmA, tmB, p, m_tiles, split_k, stream); break;
      case 128: rc = launch_igemm_2cta<Prec, 128, 6>(tmA, tmB, p, m_tiles, split_k, stream); break;
      case 64: rc = launch_igemm_2cta<Prec, 64, 6>(tmA, tmB, p, m_tiles, split_k, stream); break;
      default: set_error("conv_igemm: block_n=%d not in {64,128,256} for the CTA-pair kernel", block_n); return B200DP_EINVAL;
    }
  } else
  switch (block_n) {
    case 256: rc = launch_igemm<Prec, 256, 4>(tmA, tmB, p, m_tiles, split_k, stream); break;
    case 128: rc = launch_igemm<Prec, 128, 3>(tmA, tmB, p, m_tiles, split_k, stream); break;
    case 64:
      rc = (p.kb_per_split == 1) ? launch_igemm<Prec, 64, 1>(tmA, tmB, p, m_tiles, split_k, stream)
                                                      : launch_igemm<Prec, 64, 4>(tmA, tmB, p, m_tiles, split_k, stream);
      break;
    case 32: rc = launch_igemm<Prec, 32, 4>(tmA, tmB, p, m_tiles, split_k, stream); break;
    default: set_error("conv_igemm: block_n=%d not in {32,64,128,256}", block_n); return B200DP_EINVAL;
  }
  if (rc || split_k == 1) return rc;
  const int threads = 256;
  const int rows_per_block = threads / (Cout / 8);
  const size_t smem = stats != nullptr ? 2 * rows_per_block * Cout * sizeof(float) : 0;
  long blocks = (static_cast<long>(p.m_total) + rows_per_block * 2 - 1) / (rows_per_block * 2);
  if (blocks > 148 * 4) blocks = 148 * 4;
  if (blocks < 1) blocks = 1;
  B200DP_CUDA_TRY(launch_pdl(splitk_finalize_kernel<elem>, dim3(static_cast<unsigned>(blocks)), dim3(threads), smem,
                             stream, workspace, bias, static_cast<elem*>(out), stats, static_cast<long>(p.m_total),
                             Cout));
  return 0;
}

extern "C" int b200dp_conv_igemm(const void* act, const void* wgt, const float* bias, void* out, float* stats,
                                 int B, int H, int W, int Cin, int Cout, int taps, int block_n,
                                 b200dp_stream_t stream_) {
  return conv_igemm_impl<PrecBf16>(act, wgt, bias, out, stats, nullptr, B, H, W, Cin, Cout, taps, block_n, 1, 0,
                         static_cast<cudaStream_t>(stream_));
}

extern "C" int b200dp_conv_igemm_splitk(const void* act, const void* wgt, const float* bias, void* out, float* stats,
                                        float* workspace, int B, int H, int W, int Cin, int Cout, int taps,
                                        int block_n, int split_k, b200dp_stream_t stream_) {
  return conv_igemm_impl<PrecBf16>(act, wgt, bias, out, stats, workspace, B, H, W, Cin, Cout, taps, block_n, split_k, 0,
                         static_cast<cudaStream_t>(stream_));
}

/* CTA-pair (tcgen05 cta_group::2) variant: 256 x block_n tile per pair of SMs, each SM stages half of
 * the weight tile.  Needs an even number of 128-pixel tiles and block_n in {64,128,256}; split_k and
 * workspace as for b200dp_conv_igemm_splitk (split_k = 1: workspace may be NULL). */
extern "C" int b200dp_conv_igemm_pair(const void* act, const void* wgt, const float* bias, void* out, float* stats,
                                      float* workspace, int B, int H, int W, int Cin, int Cout, int taps,
                                      int block_n, int split_k, b200dp_stream_t stream_) {
  return conv_igemm_impl<PrecBf16>(act, wgt, bias, out, stats, workspace, B, H, W, Cin, Cout, taps, block_n, split_k, 1,
                         static_cast<cudaStream_t>(stream_));
}

/* Data gradient of a 3x3 conv (act = dY, wgt = tap-flipped transposed weights, as for the other entry
 * points) that ALSO accumulates the BatchNorm-backward sums of the block below it: `bn_out` is that
 * block's stored output a = relu(bn(z)) (max-pooled or not), same [B,H,W,Cout] shape as `out`; on return
 * sums[0..Cout) += sum_{a>0} g and sums[Cout..2Cout) += sum_{a>0} g*(a-beta)/gamma with g the gradient just
 * written to `out` -- the contract of b200dp_bn_relu_pool_bwd_reduce_from_out, without its extra pass.
 * cta_pair / split_k / workspace as for b200dp_conv_igemm_pair / _splitk.  This is the data gradient the executor
 * runs for every block but the first (0.700 -> 0.689 ms per step against dgrad + a separate reduction pass). */
extern "C" int b200dp_conv_dgrad_bn_reduce(const void* act, const void* wgt, void* out, float* workspace,
                                           const void* bn_out, const float* gamma, const float* beta, float* sums,
                                           int B, int H, int W, int Cin, int Cout, int taps, int block_n, int split_k,
                                           int cta_pair, b200dp_stream_t stream_) {
  if (bn_out == nullptr || gamma == nullptr || beta == nullptr || sums == nullptr) {
    set_error("conv_dgrad_bn_reduce: bn_out, gamma, beta and sums are required");
    return B200DP_EINVAL;
  }
  const BnReduceArgs bnr{bn_out, gamma, beta, sums};
  return conv_igemm_impl<PrecBf16>(act, wgt, nullptr, out, nullptr, workspace, B, H, W, Cin, Cout, taps, block_n,
                                   split_k, cta_pair, static_cast<cudaStream_t>(stream_), &bnr);
}

template <class Prec>
static int conv_wgrad_impl(const void* dy, const void* x, float* dw, int B, int H, int W, int Cin, int Cout, int taps,
                           int cin_valid, int split_k, int accumulate, int block_n_req, cudaStream_t stream,
                           const WgradPushArgs* push = nullptr) {
  constexpr int kPix = wgrad_pixels(Prec::kTf32);
  constexpr int kCh = Prec::kBlockK;
  if ((taps != 9 && taps != 1) || Cin % kCh != 0 || (Cout % 128 != 0 && Cout != 64) || B <= 0 || cin_valid > Cin ||
      cin_valid <= 0 || (push != nullptr && Prec::kTf32)) {
    set_error("conv_wgrad: unsupported shape B=%d H=%d W=%d Cin=%d Cout=%d taps=%d", B, H, W, Cin, Cout, taps);
    return B200DP_EINVAL;
  }
  WgradParams p{};
  p.H = H;
  p.W = W;
  if (tile_geometry(H, W, kPix, &p.TH, &p.TN) != 0) {
    set_error("conv_wgrad: B=%d H=%d W=%d cannot be tiled into %d-pixel blocks", B, H, W, kPix);
    return B200DP_EINVAL;
  }
  int block_n = (Cin % 128 == 0) ? 128 : (Cin % 64 == 0 ? 64 : 32);
  if (block_n_req == 256 && Cin % 256 == 0) block_n = 256;
  if (block_n_req == 64 && Cin % 64 == 0) block_n = 64;
  p.taps = taps;
  p.cin = Cin;
  p.cout = Cout;
  p.co_tiles = (Cout + kBlockM - 1) / kBlockM;
  p.ci_tiles = Cin / block_n;
  p.a_boxes = (Cout >= kBlockM ? kBlockM : Cout) / kCh;
  // A K block is kPix pixels; when a block spans whole images (TN > 1) the last one may hang over
  // the end of the batch -- the tensor maps are sized to B, so TMA zero-fills the overhang.
  p.kblocks_total = (p.TN > 1) ? (B + p.TN - 1) / p.TN : (B * H * W) / kPix;
  const int tiles = taps * p.co_tiles * p.ci_tiles;
  if (split_k <= 0) {
    split_k = (148 + tiles - 1) / tiles;
    if (split_k > p.kblocks_total) split_k = p.kblocks_total;
    // keep at least 4 K blocks per CTA so the pipeline prologue is amortised
    while (split_k > 1 && p.kblocks_total / split_k < 4) --split_k;
  }
  if (split_k > p.kblocks_total) split_k = p.kblocks_total;
  p.kblocks_per_split = (p.kblocks_total + split_k - 1) / split_k;
  split_k = (p.kblocks_total + p.kblocks_per_split - 1) / p.kblocks_per_split;
  p.cin_valid = cin_valid;
  p.ld_out = taps * cin_valid;
  p.use_red = (split_k > 1 || accumulate) ? 1 : 0;
  p.out = dw;
  CUtensorMap tmDY, tmX;
  {
    const uint64_t dims[2] = {(uint64_t)Cout, (uint64_t)B * H * W};
    const uint32_t box[2] = {(uint32_t)kCh, (uint32_t)kPix};
    int rc = make_tmap(&tmDY, dy, 2, dims, box, Prec::kBytes, Prec::kTf32);
    if (rc) return rc;
  }
  {
    const uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)B};
    const uint32_t box[4] = {(uint32_t)kCh, (uint32_t)W, (uint32_t)p.TH, (uint32_t)p.TN};
    int rc = make_tmap(&tmX, x, 4, dims, box, Prec::kBytes, Prec::kTf32);
    if (rc) return rc;
  }
  // 3 stages (96 / 144 / 72 KiB): the 128-wide kernel then fits on an SM next to a 3-stage dgrad CTA
  // (99 KiB), which matters because weight and data gradients run on two streams concurrently
  if constexpr (!Prec::kTf32) if (push != nullptr) {
    if (push->world < 2 || push->world > 8 || push->rank < 0 || push->rank >= push->world ||
        push->peer_stage == nullptr || push->tile_counters == nullptr || push->n_counters < tiles) {
      set_error("conv_wgrad_push: need 2..8 ranks, their staging pointers and >= %d tile counters", tiles);
      return B200DP_EINVAL;
    }
    if (cin_valid != Cin || (p.ld_out & 3) != 0 || (push->flat_off & 3) != 0 || (push->own_start & 3) != 0 ||
        (push->own_len & 3) != 0 || (push->stage_base & 3) != 0 || push->flat_off < push->own_start ||
        push->flat_off + static_cast<long long>(Cout) * p.ld_out > push->own_start + push->own_len) {
      set_error("conv_wgrad_push: the weight tensor must be unpadded, 16-byte aligned and inside its bucket");
      return B200DP_EINVAL;
    }
    WgradPushParams pp{};
    static_cast<WgradParams&>(pp) = p;
    pp.tile_counters = push->tile_counters;
    pp.splits = split_k;
    pp.world = push->world;
    pp.rank = push->rank;
    pp.flat_off = push->flat_off;
    pp.own_start = push->own_start;
    const long long n4 = push->own_len >> 2;
    pp.per4 = (n4 + push->world - 1) / push->world;
    pp.stage_base = push->stage_base;
    for (int r = 0; r < 8; ++r)
      pp.peer_stage[r] = r < push->world ? static_cast<float*>(push->peer_stage[r]) : nullptr;
    if (block_n == 256) return launch_wgrad_push<256, 4>(tmDY, tmX, pp, tiles, split_k, stream);
    if (block_n == 128) return launch_wgrad_push<128, 6>(tmDY, tmX, pp, tiles, split_k, stream);
    return launch_wgrad_push<64, 8>(tmDY, tmX, pp, tiles, split_k, stream);
  }
  // one CTA per SM: the pipeline takes all of the shared memory (8 / 6 / 4 stages = 192 KB); with the 3-4 stage
  // pipelines of round 1 the step was 10 us slower (latency-bound split-K slices, profiles/round2/ab_results.md)
  if (block_n == 256) return launch_wgrad<Prec, 256, 4>(tmDY, tmX, p, tiles, split_k, stream);
  if (block_n == 128) return launch_wgrad<Prec, 128, 6>(tmDY, tmX, p, tiles, split_k, stream);
  if (block_n == 64) return launch_wgrad<Prec, 64, 8>(tmDY, tmX, p, tiles, split_k, stream);
  if constexpr (Prec::kTf32) return launch_wgrad<Prec, 32, 8>(tmDY, tmX, p, tiles, split_k, stream);
  set_error("conv_wgrad: Cin=%d needs a 32-wide tile, which only the tf32 kernel has", Cin);
  return B200DP_EINVAL;
}

extern "C" int b200dp_conv_wgrad(const void* dy, const void* x, float* dw, int B, int H, int W, int Cin,
                                 int Cout, int taps, int cin_valid, int split_k, int accumulate,
                                 b200dp_stream_t stream_) {
  return conv_wgrad_impl<PrecBf16>(dy, x, dw, B, H, W, Cin, Cout, taps, cin_valid, split_k, accumulate, 0,
                         static_cast<cudaStream_t>(stream_));
}

/* Weight gradient fused with the transfer half of the data-parallel reduce-scatter (optional path,
 * B200DP_WGRAD_PUSH=1: correct, but measured slower than the bucket reducer at 2 and 8 GPUs).
 * As b200dp_conv_wgrad_tiled (dw zero on entry, full Cin columns), plus: the tensor lives at element
 * flat_off of the flat gradient buffer, inside the bucket [own_start, own_start + own_len) that the
 * ranks own in `world` equal float4 slices (the slicing of the two-shot all-reduce).  Whenever an output
 * tile is complete on this rank, its float4s owned by another rank r are stored -- over NVLink, as the
 * GEMM tiles finish -- into peer_stage[r] at  stage_base + (rank * per4 + j) * 4,  j = index inside
 * r's slice, per4 = ceil(own_len / 4 / world).  The owner adds the W-1 staged copies to its own gradient
 * in b200dp_reduce_staged_sgd.  tile_counters: >= taps * ceil(Cout/128) * Cin/block_n zeroed uints (left
 * zero).  peer_stage: host array of `world` device pointers. */
extern "C" int b200dp_conv_wgrad_push(const void* dy, const void* x, float* dw, int B, int H, int W, int Cin,
                                      int Cout, int taps, int split_k, int block_n, unsigned int* tile_counters,
                                      int n_counters, int world, int rank, void* const* peer_stage,
                                      long long flat_off, long long own_start, long long own_len,
                                      long long stage_base, b200dp_stream_t stream_) {
  const WgradPushArgs push{tile_counters, n_counters, world, rank, peer_stage, flat_off, own_start, own_len,
                           stage_base};
  return conv_wgrad_impl<PrecBf16>(dy, x, dw, B, H, W, Cin, Cout, taps, Cin, split_k, 0, block_n,
                         static_cast<cudaStream_t>(stream_), &push);
}

/* As b200dp_conv_wgrad with an explicit input-channel tile (64, 128 or 256; 0 = default). */
extern "C" int b200dp_conv_wgrad_tiled(const void* dy, const void* x, float* dw, int B, int H, int W, int Cin,
                                       int Cout, int taps, int cin_valid, int split_k, int accumulate, int block_n,
                                       b200dp_stream_t stream_) {
  return conv_wgrad_impl<PrecBf16>(dy, x, dw, B, H, W, Cin, Cout, taps, cin_valid, split_k, accumulate, block_n,
                         static_cast<cudaStream_t>(stream_));
}

/* ------------------------------------------------------------------------------------------------------
 * Precision-generic forms (dtype 0 = bf16 operands and activations, 1 = tf32: fp32 activations and operand
 * copies in memory, tcgen05.mma kind::tf32).  With dtype 1: act / out / bn_out / dy / x are fp32, wgt is fp32
 * [Cout][taps*Cin], Cin must be a multiple of 32 (bf16: 64).  Everything else as for the bf16 entry points.
 * ------------------------------------------------------------------------------------------------------ */
extern "C" int b200dp_conv_igemm_ex(const void* act, const void* wgt, const float* bias, void* out, float* stats,
                                    float* workspace, int B, int H, int W, int Cin, int Cout, int taps, int block_n,
                                    int split_k, int cta_pair, int dtype, b200dp_stream_t stream_) {
  cudaStream_t st = static_cast<cudaStream_t>(stream_);
  if (dtype == 1)
    return conv_igemm_impl<PrecTf32>(act, wgt, bias, out, stats, workspace, B, H, W, Cin, Cout, taps, block_n, split_k,
                                     cta_pair ? 1 : 0, st);
  return conv_igemm_impl<PrecBf16>(act, wgt, bias, out, stats, workspace, B, H, W, Cin, Cout, taps, block_n, split_k,
                                   cta_pair ? 1 : 0, st);
}

extern "C" int b200dp_conv_dgrad_bn_reduce_ex(const void* act, const void* wgt, void* out, float* workspace,
                                              const void* bn_out, const float* gamma, const float* beta, float* sums,
                                              int B, int H, int W, int Cin, int Cout, int taps, int block_n,
                                              int split_k, int cta_pair, int dtype, b200dp_stream_t stream_) {
  if (bn_out == nullptr || gamma == nullptr || beta == nullptr || sums == nullptr) {
    set_error("conv_dgrad_bn_reduce: bn_out, gamma, beta and sums are required");
    return B200DP_EINVAL;
  }
  const BnReduceArgs bnr{bn_out, gamma, beta, sums};
  cudaStream_t st = static_cast<cudaStream_t>(stream_);
  if (dtype == 1)
    return conv_igemm_impl<PrecTf32>(act, wgt, nullptr, out, nullptr, workspace, B, H, W, Cin, Cout, taps, block_n,
                                     split_k, cta_pair ? 1 : 0, st, &bnr);
  return conv_igemm_impl<PrecBf16>(act, wgt, nullptr, out, nullptr, workspace, B, H, W, Cin, Cout, taps, block_n,
                                   split_k, cta_pair ? 1 : 0, st, &bnr);
}

extern "C" int b200dp_conv_wgrad_ex(const void* dy, const void* x, float* dw, int B, int H, int W, int Cin, int Cout,
                                    int taps, int cin_valid, int split_k, int accumulate, int block_n, int dtype,
                                    b200dp_stream_t stream_) {
  cudaStream_t st = static_cast<cudaStream_t>(stream_);
  if (dtype == 1)
    return conv_wgrad_impl<PrecTf32>(dy, x, dw, B, H, W, Cin, Cout, taps, cin_valid, split_k, accumulate, block_n, st);
  return conv_wgrad_impl<PrecBf16>(dy, x, dw, B, H, W, Cin, Cout, taps, cin_valid, split_k, accumulate, block_n, st);
}
