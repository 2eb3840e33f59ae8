// In-kernel NVLink collectives over a symmetric heap (one process per GPU).
//
// Memory model: every rank cudaMalloc's one heap of identical size, exports it with CUDA IPC
// through the rendezvous store and maps every peer's heap; an object allocated at byte offset
// `o` on one rank lives at `o` on all of them, so a kernel reaches rank r's copy at
// peer_base[r] + o with plain ld/st.global over NVLink 5 / NVSwitch.  A small signal pad at
// the start of the heap holds per-(channel, slot, source rank, block) 32-bit flags that are
// written with st.release.sys and polled with ld.acquire.sys; values are launch epochs that
// only grow, so the pad never has to be cleared.
//
// Kernels:
//   allreduce_twoshot  reduce-scatter by P2P loads of my 1/W slice from every rank, 1/W scale,
//                      optional SGD(momentum, weight decay) epilogue on the slice, all-gather by
//                      P2P stores of the mean gradient or of the UPDATED PARAMETERS.
//                      (ref: master/part2b/part2b.py:43-45 all_reduce; part3 DDP buckets)
//   allreduce_oneshot  every rank loads all W copies and reduces locally (latency-optimal for
//                      small buckets), optional fused SGD.
//   ps_reduce_bcast    parameter server: rank 0 pulls every rank's gradient, sums in rank order
//                      starting from zero, divides by W and pushes the mean to every rank
//                      (ref: master/part2a/part2a.py:42-52); `push` variant: workers store their
//                      gradients into rank 0's staging area first (the isend/irecv flavour,
//                      ref: master/part2a/part2a_extra.py:41-58).
//   broadcast          rank `root` pushes a range to every rank (DDP buffer / module-state sync,
//                      SURVEY.md 2.6 C11-C12).
//
// Every wait is bounded: after `timeout_ns` a block reports which peer / slot / block it was
// waiting for and traps, so a dead rank surfaces as a CUDA error instead of a hang.
#include <cuda.h>
#include <cudaTypedefs.h>
#include <stdio.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include <string>
#include <vector>

#include "../include/b200dp.h"
#include "common.cuh"
#include "host_util.h"

extern "C" int b200dp_store_set(void* h, const char* key, const void* data, int len);
extern "C" int b200dp_store_get(void* h, const char* key, void* buf, int cap, int* out_len);
extern "C" int b200dp_store_barrier(void* h, const char* name);
extern "C" int b200dp_store_timeout_ms(void* h);

namespace b200dp {
int fd_listen(const std::string& name);
int fd_send(const std::string& name, int fd, int tag, int timeout_ms);
int fd_recv(int listen_fd, int* tag);
int fd_recv_timeout(int listen_fd, int* tag, int timeout_ms);

constexpr int kMaxWorld = 8;
constexpr int kMaxBlocks = 256;
constexpr int kNumChannels = 4;
constexpr int kSlotsPerChannel = 2;  // 0 = entry barrier, 1 = exit barrier
constexpr size_t kSignalPadBytes =
    sizeof(uint32_t) * kNumChannels * kSlotsPerChannel * kMaxWorld * kMaxBlocks + 4096;

struct PeerTable {
  char* base[kMaxWorld];  // peer-mapped heap base addresses, index = rank
};

struct Comm {
  int rank = 0, world = 1, device = 0;
  void* store = nullptr;
  char* local_base = nullptr;
  size_t heap_bytes = 0;
  PeerTable peers{};
  uint32_t* epochs = nullptr;       // [kNumChannels] launch epoch per channel (device memory, local)
  uint32_t* done_blocks = nullptr;  // [kNumChannels] finished-block counters
  unsigned long long timeout_ns = 20ull * 1000 * 1000 * 1000;
  bool ipc_opened[kMaxWorld] = {false};
  // virtual-memory-management heap (cuMemCreate + POSIX fd exchange) with an NVSwitch multicast
  // mapping: mc_base + offset addresses the copies of ALL ranks at once (multimem.* instructions)
  bool vmm = false;
  CUmemGenericAllocationHandle mem_handle = 0;
  CUmemGenericAllocationHandle peer_handle[kMaxWorld] = {0};
  CUmemGenericAllocationHandle mc_handle = 0;
  char* mc_base = nullptr;
  bool prefer_nvls = false;  // algo 0 ("choose") picks the NVLS kernel for large ranges
};

struct CommDev {
  PeerTable peers;
  char* mc_base;  // multicast address of the heap (NULL without NVLS)
  int rank, world;
  uint32_t* epochs;
  uint32_t* done_blocks;
  unsigned long long timeout_ns;
};

B200DP_DEVINL uint32_t* flag_addr(char* heap_base, int channel, int slot, int src_rank, int block) {
  return reinterpret_cast<uint32_t*>(heap_base) +
         ((static_cast<size_t>(channel) * kSlotsPerChannel + slot) * kMaxWorld + src_rank) * kMaxBlocks + block;
}

// All-to-all barrier between block `blockIdx.x` of every rank.  Thread t < world talks to peer t.
B200DP_DEVINL void rank_barrier(const CommDev& c, int channel, int slot, uint32_t epoch) {
  __syncthreads();
  if (threadIdx.x < c.world) {
    const int peer = threadIdx.x;
    __threadfence_system();
    st_release_sys(flag_addr(c.peers.base[peer], channel, slot, c.rank, blockIdx.x), epoch);
    const uint32_t* mine = flag_addr(c.peers.base[c.rank], channel, slot, peer, blockIdx.x);
    const uint64_t t0 = globaltimer_ns();
    while (static_cast<int32_t>(ld_acquire_sys(mine) - epoch) < 0) {
      if (globaltimer_ns() - t0 > c.timeout_ns) {
        printf("b200dp: rank %d block %d timed out waiting for rank %d (channel %d slot %d epoch %u, saw %u)\n",
               c.rank, blockIdx.x, peer, channel, slot, epoch, ld_relaxed_sys(mine));
        __trap();
      }
    }
  }
  __syncthreads();
}

B200DP_DEVINL uint32_t begin_collective(const CommDev& c, int channel) {
  return *reinterpret_cast<volatile uint32_t*>(c.epochs + channel) + 1;
}

// The last block to finish publishes the new epoch for the next launch on this channel.
B200DP_DEVINL void end_collective(const CommDev& c, int channel, uint32_t epoch) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const uint32_t prev = atomicAdd(c.done_blocks + channel, 1u);
    if (prev == gridDim.x - 1) {
      c.done_blocks[channel] = 0;
      __threadfence();
      *reinterpret_cast<volatile uint32_t*>(c.epochs + channel) = epoch;
    }
  }
}

B200DP_DEVINL float4 ld_peer_f4(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }
B200DP_DEVINL void st_peer_f4(float* p, const float4& v) { __stcg(reinterpret_cast<float4*>(p), v); }

struct SgdHyper {
  float lr, momentum, weight_decay;
};

B200DP_DEVINL float4 sgd4(float4 g, float4& p, float4& m, const SgdHyper& h) {
  m.x = fmaf(h.momentum, m.x, fmaf(h.weight_decay, p.x, g.x));
  m.y = fmaf(h.momentum, m.y, fmaf(h.weight_decay, p.y, g.y));
  m.z = fmaf(h.momentum, m.z, fmaf(h.weight_decay, p.z, g.z));
  m.w = fmaf(h.momentum, m.w, fmaf(h.weight_decay, p.w, g.w));
  p.x = fmaf(-h.lr, m.x, p.x);
  p.y = fmaf(-h.lr, m.y, p.y);
  p.z = fmaf(-h.lr, m.z, p.z);
  p.w = fmaf(-h.lr, m.w, p.w);
  return p;
}

struct RangeArgs {
  size_t grad_off;   // byte offset of the flat gradient buffer inside the heap
  size_t param_off;  // byte offset of the flat parameter buffer (fused SGD)
  size_t mom_off;    // byte offset of the flat momentum buffer (fused SGD)
  size_t stage_off;  // byte offset of the parameter-server staging area (push variant)
  long start, end;   // element range [start, end) of the flat buffers; multiples of 4
  int channel;
  int root;
  SgdHyper sgd;
};

// ---------------------------------------------------------------------------
// two-shot all-reduce: reduce-scatter -> (scale, SGD) -> all-gather
// ---------------------------------------------------------------------------
template <bool FUSE_SGD, bool PRESCALE, bool EXIT_BARRIER>
B200DP_DEVINL void allreduce_twoshot_body(const CommDev& c, const RangeArgs& a) {
  const uint32_t epoch = begin_collective(c, a.channel);
  rank_barrier(c, a.channel, 0, epoch);
  const int W = c.world;
  const long n4 = (a.end - a.start) >> 2;           // float4 elements in the bucket
  const long per = (n4 + W - 1) / W;                // slice length per rank
  const long lo = min(per * c.rank, n4), hi = min(lo + per, n4);
  const float inv_w = 1.f / static_cast<float>(W);
  constexpr int U = 4;  // independent float4 streams per thread: U x W 16-byte peer loads in flight
  const long tstride = static_cast<long>(gridDim.x) * blockDim.x;
  for (long i0 = lo + blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i0 < hi; i0 += tstride * U) {
    float4 acc[U];
    size_t eoff[U];
    bool live[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long i = i0 + u * tstride;
      live[u] = i < hi;
      eoff[u] = static_cast<size_t>(a.start + ((live[u] ? i : lo) << 2)) * sizeof(float);
      acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r) {
      if (r < W) {
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
          v[u] = ld_peer_f4(reinterpret_cast<const float*>(c.peers.base[r] + a.grad_off + eoff[u]));
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (PRESCALE) {  // divide-then-sum, the order of part2b / DDP
            v[u].x *= inv_w; v[u].y *= inv_w; v[u].z *= inv_w; v[u].w *= inv_w;
          }
          acc[u].x += v[u].x; acc[u].y += v[u].y; acc[u].z += v[u].z; acc[u].w += v[u].w;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (!live[u]) continue;
      float4 g = acc[u];
      if (!PRESCALE) {
        g.x *= inv_w; g.y *= inv_w; g.z *= inv_w; g.w *= inv_w;
      }
      if (FUSE_SGD) {
        char* me = c.peers.base[c.rank];
        float4 p = *reinterpret_cast<const float4*>(me + a.param_off + eoff[u]);
        float4 m = *reinterpret_cast<const float4*>(me + a.mom_off + eoff[u]);
        const float4 np = sgd4(g, p, m, a.sgd);
        *reinterpret_cast<float4*>(me + a.mom_off + eoff[u]) = m;
        *reinterpret_cast<float4*>(me + a.grad_off + eoff[u]) = g;  // owner keeps the mean gradient
#pragma unroll
        for (int r = 0; r < kMaxWorld; ++r)
          if (r < W) st_peer_f4(reinterpret_cast<float*>(c.peers.base[r] + a.param_off + eoff[u]), np);
      } else {
#pragma unroll
        for (int r = 0; r < kMaxWorld; ++r)
          if (r < W) st_peer_f4(reinterpret_cast<float*>(c.peers.base[r] + a.grad_off + eoff[u]), g);
      }
    }
  }
  if (EXIT_BARRIER) rank_barrier(c, a.channel, 1, epoch);
  end_collective(c, a.channel, epoch);
}

template <bool FUSE_SGD, bool PRESCALE>
__global__ void __launch_bounds__(512) allreduce_twoshot_kernel(const CommDev c, const RangeArgs a) {
  allreduce_twoshot_body<FUSE_SGD, PRESCALE, true>(c, a);
}

// Without the exit barrier (the bucket reducer's kernels 0..n-2 of a step).  What the exit barrier guarantees --
// every peer's stores into my buffers have landed, every peer has finished reading mine -- is also implied by
// the ENTRY barrier of the next collective on the same channel: a peer reaches it only after its previous
// kernel has completed, stores fenced.  So only the last bucket kernel of a step needs the exit barrier; at
// 8 ranks a barrier costs ~10 us and a step has eight bucket kernels queued back to back.
template <bool FUSE_SGD, bool PRESCALE>
__global__ void __launch_bounds__(512) allreduce_twoshot_lazy_kernel(const CommDev c, const RangeArgs a) {
  allreduce_twoshot_body<FUSE_SGD, PRESCALE, false>(c, a);
}

// ---------------------------------------------------------------------------
// one-shot all-reduce: every rank reduces the whole range locally.  The range must fit in one
// pass (gridDim.x * blockDim.x float4's) so that the read and write phases are separated by a
// single barrier.
// ---------------------------------------------------------------------------
template <bool FUSE_SGD, bool PRESCALE>
__global__ void __launch_bounds__(512) allreduce_oneshot_kernel(const CommDev c, const RangeArgs a) {
  const uint32_t epoch = begin_collective(c, a.channel);
  rank_barrier(c, a.channel, 0, epoch);
  const int W = c.world;
  const long n4 = (a.end - a.start) >> 2;
  const long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x;
  const float inv_w = 1.f / static_cast<float>(W);
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  const size_t eoff = static_cast<size_t>(a.start + (i << 2)) * sizeof(float);
  if (i < n4) {
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r) {
      if (r < W) {
        float4 v = ld_peer_f4(reinterpret_cast<const float*>(c.peers.base[r] + a.grad_off + eoff));
        if (PRESCALE) {
          v.x *= inv_w; v.y *= inv_w; v.z *= inv_w; v.w *= inv_w;
        }
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
    }
    if (!PRESCALE) {
      acc.x *= inv_w; acc.y *= inv_w; acc.z *= inv_w; acc.w *= inv_w;
    }
  }
  // every rank has finished reading block-partition `blockIdx.x` of every peer before anyone overwrites it
  rank_barrier(c, a.channel, 1, epoch);
  if (i < n4) {
    char* me = c.peers.base[c.rank];
    *reinterpret_cast<float4*>(me + a.grad_off + eoff) = acc;
    if (FUSE_SGD) {
      float4 p = *reinterpret_cast<const float4*>(me + a.param_off + eoff);
      float4 m = *reinterpret_cast<const float4*>(me + a.mom_off + eoff);
      sgd4(acc, p, m, a.sgd);
      *reinterpret_cast<float4*>(me + a.mom_off + eoff) = m;
      *reinterpret_cast<float4*>(me + a.param_off + eoff) = p;
    }
  }
  end_collective(c, a.channel, epoch);
}

// ---------------------------------------------------------------------------
// NVLS all-reduce: the reduction happens INSIDE the NVSwitch.  Each rank owns 1/W of the range;
// one multimem.ld_reduce returns the sum over all ranks' copies of 16 bytes, the 1/W scale (and the
// SGD update of the owned slice) runs in registers, and one multimem.st writes the result into every
// rank's copy.  Per rank that is S/W bytes in and S/W bytes out over NVLink instead of the two-shot
// kernel's (W-1)/W * S each way.  
// ---------------------------------------------------------------------------
B200DP_DEVINL float4 multimem_ld_reduce_add_f4(const float* mc_addr) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc_addr)
               : "memory");
  return v;
}
B200DP_DEVINL void multimem_st_f4(float* mc_addr, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}

template <bool FUSE_SGD, bool EXIT_BARRIER>
B200DP_DEVINL void allreduce_nvls_body(const CommDev& c, const RangeArgs& a) {
  const uint32_t epoch = begin_collective(c, a.channel);
  rank_barrier(c, a.channel, 0, epoch);
  const int W = c.world;
  const long n4 = (a.end - a.start) >> 2;
  const long per = (n4 + W - 1) / W;
  const long lo = min(per * c.rank, n4), hi = min(lo + per, n4);
  const float inv_w = 1.f / static_cast<float>(W);
  constexpr int U = 4;
  const long tstride = static_cast<long>(gridDim.x) * blockDim.x;
  for (long i0 = lo + blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i0 < hi; i0 += tstride * U) {
    float4 g[U];
    size_t eoff[U];
    bool live[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long i = i0 + u * tstride;
      live[u] = i < hi;
      eoff[u] = static_cast<size_t>(a.start + ((live[u] ? i : lo) << 2)) * sizeof(float);
      g[u] = multimem_ld_reduce_add_f4(reinterpret_cast<const float*>(c.mc_base + a.grad_off + eoff[u]));
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (!live[u]) continue;
      g[u].x *= inv_w; g[u].y *= inv_w; g[u].z *= inv_w; g[u].w *= inv_w;
      if (FUSE_SGD) {
        char* me = c.peers.base[c.rank];
        float4 p = *reinterpret_cast<const float4*>(me + a.param_off + eoff[u]);
        float4 m = *reinterpret_cast<const float4*>(me + a.mom_off + eoff[u]);
        const float4 np = sgd4(g[u], p, m, a.sgd);
        *reinterpret_cast<float4*>(me + a.mom_off + eoff[u]) = m;
        multimem_st_f4(reinterpret_cast<float*>(c.mc_base + a.param_off + eoff[u]), np);
      } else {
        multimem_st_f4(reinterpret_cast<float*>(c.mc_base + a.grad_off + eoff[u]), g[u]);
      }
    }
  }
  if (EXIT_BARRIER) rank_barrier(c, a.channel, 1, epoch);
  end_collective(c, a.channel, epoch);
}

template <bool FUSE_SGD>
__global__ void __launch_bounds__(512) allreduce_nvls_kernel(const CommDev c, const RangeArgs a) {
  allreduce_nvls_body<FUSE_SGD, true>(c, a);
}

template <bool FUSE_SGD>
__global__ void __launch_bounds__(512) allreduce_nvls_lazy_kernel(const CommDev c, const RangeArgs a) {
  allreduce_nvls_body<FUSE_SGD, false>(c, a);
}

// ---------------------------------------------------------------------------
// Owner-side half of the fused "weight-gradient GEMM -> reduce-scatter" path (optional, B200DP_WGRAD_PUSH=1).
// The conv_wgrad_push kernels of every rank have already stored their finished tiles of this bucket
// into the owner's staging slots (conv_igemm.cu), so the owner reads only LOCAL memory here: its own
// gradient plus W-1 staged copies.  Tensors that are not produced by a wgrad GEMM (biases, BatchNorm
// weights, the classifier) are outside [pushed_lo, pushed_hi) and are pulled from the peers as in the
// two-shot kernel.  Then 1/W, SGD on the owned slice, and the updated parameters go to every rank
// (P2P stores, or one multimem.st through the switch).
// ---------------------------------------------------------------------------
struct StagedArgs {
  RangeArgs r;
  size_t stage_buf_off;       // byte offset of the staging buffer inside the heap
  long pushed_lo, pushed_hi;  // absolute element range (multiples of 4) whose gradients arrive by push
  long stage_base;            // element offset of this bucket's area inside the staging buffer
};

template <bool NVLS>
__global__ void __launch_bounds__(512) reduce_staged_sgd_kernel(const CommDev c, const StagedArgs s) {
  const RangeArgs& a = s.r;
  const uint32_t epoch = begin_collective(c, a.channel);
  // every rank has launched this collective, i.e. its wgrad kernels of this bucket (earlier in stream
  // order) have completed and their staged tiles are visible
  rank_barrier(c, a.channel, 0, epoch);
  const int W = c.world;
  const long n4 = (a.end - a.start) >> 2;
  const long per = (n4 + W - 1) / W;
  const long lo = min(per * c.rank, n4), hi = min(lo + per, n4);
  const float inv_w = 1.f / static_cast<float>(W);
  char* me = c.peers.base[c.rank];
  const float* stage = reinterpret_cast<const float*>(me + s.stage_buf_off) + s.stage_base;
  const long tstride = static_cast<long>(gridDim.x) * blockDim.x;
  for (long i = lo + blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < hi; i += tstride) {
    const long e = a.start + (i << 2);
    const size_t eoff = static_cast<size_t>(e) * sizeof(float);
    float4 g;
    if (e >= s.pushed_lo && e < s.pushed_hi) {
      g = __ldcg(reinterpret_cast<const float4*>(me + a.grad_off + eoff));
#pragma unroll
      for (int r = 0; r < kMaxWorld; ++r) {
        if (r < W && r != c.rank) {
          const float4 v = __ldcg(reinterpret_cast<const float4*>(stage + ((r * per + (i - lo)) << 2)));
          g.x += v.x; g.y += v.y; g.z += v.z; g.w += v.w;
        }
      }
    } else {
      g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int r = 0; r < kMaxWorld; ++r) {
        if (r < W) {
          const float4 v = ld_peer_f4(reinterpret_cast<const float*>(c.peers.base[r] + a.grad_off + eoff));
          g.x += v.x; g.y += v.y; g.z += v.z; g.w += v.w;
        }
      }
    }
    g.x *= inv_w; g.y *= inv_w; g.z *= inv_w; g.w *= inv_w;
    float4 p = *reinterpret_cast<const float4*>(me + a.param_off + eoff);
    float4 m = *reinterpret_cast<const float4*>(me + a.mom_off + eoff);
    const float4 np = sgd4(g, p, m, a.sgd);
    *reinterpret_cast<float4*>(me + a.mom_off + eoff) = m;
    *reinterpret_cast<float4*>(me + a.grad_off + eoff) = g;  // the owner keeps the mean gradient
    if (NVLS) {
      multimem_st_f4(reinterpret_cast<float*>(c.mc_base + a.param_off + eoff), np);
    } else {
#pragma unroll
      for (int r = 0; r < kMaxWorld; ++r)
        if (r < W) st_peer_f4(reinterpret_cast<float*>(c.peers.base[r] + a.param_off + eoff), np);
    }
  }
  rank_barrier(c, a.channel, 1, epoch);
  end_collective(c, a.channel, epoch);
}

// ---------------------------------------------------------------------------
// parameter server (rank `root` reduces, everyone receives the mean)
// ---------------------------------------------------------------------------
template <bool PUSH>
__global__ void __launch_bounds__(512) ps_reduce_bcast_kernel(const CommDev c, const RangeArgs a) {
  const uint32_t epoch = begin_collective(c, a.channel);
  const int W = c.world;
  const long n4 = (a.end - a.start) >> 2;
  const long stride = static_cast<long>(gridDim.x) * blockDim.x;
  const long first = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x;
  if (PUSH && c.rank != a.root) {
    // worker "isend": store my gradient into my slot of the root's staging area
    const int slot = c.rank < a.root ? c.rank : c.rank - 1;
    char* dst = c.peers.base[a.root] + a.stage_off + static_cast<size_t>(slot) * (n4 << 4);
    const char* src = c.peers.base[c.rank] + a.grad_off + static_cast<size_t>(a.start) * sizeof(float);
    for (long i = first; i < n4; i += stride)
      st_peer_f4(reinterpret_cast<float*>(dst + (i << 4)), *reinterpret_cast<const float4*>(src + (i << 4)));
  }
  rank_barrier(c, a.channel, 0, epoch);
  if (c.rank == a.root) {
    const float wf = static_cast<float>(W);
    for (long i = first; i < n4; i += stride) {
      const size_t eoff = static_cast<size_t>(a.start + (i << 2)) * sizeof(float);
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int r = 0; r < kMaxWorld; ++r) {
        if (r < W) {
          float4 v;
          if (PUSH && r != a.root) {
            const int slot = r < a.root ? r : r - 1;
            v = *reinterpret_cast<const float4*>(c.peers.base[a.root] + a.stage_off +
                                                 static_cast<size_t>(slot) * (n4 << 4) + (i << 4));
          } else {
            v = ld_peer_f4(reinterpret_cast<const float*>(c.peers.base[r] + a.grad_off + eoff));
          }
          acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
      }
      // sum first, then a true division, exactly like `grad_sum / 4`
      acc.x /= wf; acc.y /= wf; acc.z /= wf; acc.w /= wf;
#pragma unroll
      for (int r = 0; r < kMaxWorld; ++r)
        if (r < W) st_peer_f4(reinterpret_cast<float*>(c.peers.base[r] + a.grad_off + eoff), acc);
    }
  }
  rank_barrier(c, a.channel, 1, epoch);
  end_collective(c, a.channel, epoch);
}

// ---------------------------------------------------------------------------
// broadcast of a byte range (multiple of 16) at heap offset grad_off from `root`
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(512) broadcast_kernel(const CommDev c, const RangeArgs a) {
  const uint32_t epoch = begin_collective(c, a.channel);
  rank_barrier(c, a.channel, 0, epoch);
  if (c.rank == a.root) {
    const long n4 = (a.end - a.start) >> 2;
    const long stride = static_cast<long>(gridDim.x) * blockDim.x;
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4; i += stride) {
      const size_t eoff = static_cast<size_t>(a.start + (i << 2)) * sizeof(float);
      const float4 v = *reinterpret_cast<const float4*>(c.peers.base[c.rank] + a.grad_off + eoff);
      for (int r = 0; r < c.world; ++r)
        if (r != c.rank) st_peer_f4(reinterpret_cast<float*>(c.peers.base[r] + a.grad_off + eoff), v);
    }
  }
  rank_barrier(c, a.channel, 1, epoch);
  end_collective(c, a.channel, epoch);
}

// ---------------------------------------------------------------------------
// all-gather of owner slices: the element range [start, end) of a flat buffer at heap byte offset
// grad_off is owned in W equal float4 slices (the slicing of the two-shot / NVLS / staged kernels);
// every rank stores ITS slice into every peer's copy.  Used before a checkpoint: the fused collectives
// keep SGD momentum only on the owner of a slice, so the full optimizer state has to be assembled.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(512) allgather_slices_kernel(const CommDev c, const RangeArgs a) {
  const uint32_t epoch = begin_collective(c, a.channel);
  rank_barrier(c, a.channel, 0, epoch);
  const int W = c.world;
  const long n4 = (a.end - a.start) >> 2;
  const long per = (n4 + W - 1) / W;
  const long lo = min(per * c.rank, n4), hi = min(lo + per, n4);
  const long stride = static_cast<long>(gridDim.x) * blockDim.x;
  for (long i = lo + blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < hi; i += stride) {
    const size_t eoff = static_cast<size_t>(a.start + (i << 2)) * sizeof(float);
    const float4 v = *reinterpret_cast<const float4*>(c.peers.base[c.rank] + a.grad_off + eoff);
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r)
      if (r < W && r != c.rank) st_peer_f4(reinterpret_cast<float*>(c.peers.base[r] + a.grad_off + eoff), v);
  }
  rank_barrier(c, a.channel, 1, epoch);
  end_collective(c, a.channel, epoch);
}

__global__ void barrier_kernel(const CommDev c, int channel) {
  const uint32_t epoch = begin_collective(c, channel);
  rank_barrier(c, channel, 0, epoch);
  end_collective(c, channel, epoch);
}

static long long globaltimer_host() {
  timespec ts;
  clock_gettime(CLOCK_REALTIME, &ts);
  return static_cast<long long>(ts.tv_sec) * 1000000000ll + ts.tv_nsec;
}

static CommDev dev_view(const Comm* c) {
  CommDev d;
  d.peers = c->peers;
  d.mc_base = c->mc_base;
  d.rank = c->rank;
  d.world = c->world;
  d.epochs = c->epochs;
  d.done_blocks = c->done_blocks;
  d.timeout_ns = c->timeout_ns;
  return d;
}

static int check_range(const Comm* c, long start, long end, int blocks, int channel, const char* what) {
  if (!c) {
    set_error("%s: null communicator", what);
    return B200DP_EINVAL;
  }
  if (start < 0 || end < start || (start & 3) || (end & 3)) {
    set_error("%s: element range [%ld, %ld) must be non-negative and a multiple of 4", what, start, end);
    return B200DP_EINVAL;
  }
  if (blocks < 1 || blocks > kMaxBlocks) {
    set_error("%s: blocks=%d outside [1, %d]", what, blocks, kMaxBlocks);
    return B200DP_EINVAL;
  }
  if (channel < 0 || channel >= kNumChannels) {
    set_error("%s: channel=%d outside [0, %d)", what, channel, kNumChannels);
    return B200DP_EINVAL;
  }
  return 0;
}

}  // namespace b200dp

using namespace b200dp;

// ---------------------------------------------------------------------------
// communicator / symmetric heap
// ---------------------------------------------------------------------------
namespace b200dp {

#define B200DP_CU_TRY(expr)                                                                     \
  do {                                                                                          \
    CUresult _r = (expr);                                                                       \
    if (_r != CUDA_SUCCESS) {                                                                   \
      set_error("%s:%d: %s -> CUresult %d", __FILE__, __LINE__, #expr, static_cast<int>(_r));   \
      return B200DP_EINVAL;                                                                     \
    }                                                                                           \
  } while (0)

template <typename Fn>
static Fn driver_fn(const char* name) {
  void* ptr = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &ptr, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
    return nullptr;
  return reinterpret_cast<Fn>(ptr);
}

struct DriverApi {
  PFN_cuMemCreate_v10020 memCreate;
  PFN_cuMemRelease_v10020 memRelease;
  PFN_cuMemAddressReserve_v10020 addrReserve;
  PFN_cuMemAddressFree_v10020 addrFree;
  PFN_cuMemMap_v10020 memMap;
  PFN_cuMemUnmap_v10020 memUnmap;
  PFN_cuMemSetAccess_v10020 setAccess;
  PFN_cuMemExportToShareableHandle_v10020 exportHandle;
  PFN_cuMemImportFromShareableHandle_v10020 importHandle;
  PFN_cuMulticastCreate_v12010 mcCreate;
  PFN_cuMulticastAddDevice_v12010 mcAddDevice;
  PFN_cuMulticastBindMem_v12010 mcBindMem;
  bool ok;
};

static const DriverApi& driver_api() {
  static DriverApi a = [] {
    DriverApi d{};
    d.memCreate = driver_fn<PFN_cuMemCreate_v10020>("cuMemCreate");
    d.memRelease = driver_fn<PFN_cuMemRelease_v10020>("cuMemRelease");
    d.addrReserve = driver_fn<PFN_cuMemAddressReserve_v10020>("cuMemAddressReserve");
    d.addrFree = driver_fn<PFN_cuMemAddressFree_v10020>("cuMemAddressFree");
    d.memMap = driver_fn<PFN_cuMemMap_v10020>("cuMemMap");
    d.memUnmap = driver_fn<PFN_cuMemUnmap_v10020>("cuMemUnmap");
    d.setAccess = driver_fn<PFN_cuMemSetAccess_v10020>("cuMemSetAccess");
    d.exportHandle = driver_fn<PFN_cuMemExportToShareableHandle_v10020>("cuMemExportToShareableHandle");
    d.importHandle = driver_fn<PFN_cuMemImportFromShareableHandle_v10020>("cuMemImportFromShareableHandle");
    d.mcCreate = driver_fn<PFN_cuMulticastCreate_v12010>("cuMulticastCreate");
    d.mcAddDevice = driver_fn<PFN_cuMulticastAddDevice_v12010>("cuMulticastAddDevice");
    d.mcBindMem = driver_fn<PFN_cuMulticastBindMem_v12010>("cuMulticastBindMem");
    d.ok = d.memCreate && d.memRelease && d.addrReserve && d.addrFree && d.memMap && d.memUnmap && d.setAccess &&
           d.exportHandle && d.importHandle && d.mcCreate && d.mcAddDevice && d.mcBindMem;
    return d;
  }();
  return a;
}

static int map_handle(const DriverApi& d, CUmemGenericAllocationHandle h, size_t bytes, int device, char** out) {
  CUdeviceptr va = 0;
  B200DP_CU_TRY(d.addrReserve(&va, bytes, 2u << 20, 0, 0));
  B200DP_CU_TRY(d.memMap(va, bytes, 0, h, 0));
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  B200DP_CU_TRY(d.setAccess(va, bytes, &acc, 1));
  *out = reinterpret_cast<char*>(va);
  return 0;
}

// Undo whatever create_vmm_heap() has set up so far (also the tail of b200dp_comm_destroy).
static void release_vmm_heap(Comm* c) {
  const DriverApi& d = driver_api();
  if (c->mc_base) {
    d.memUnmap(reinterpret_cast<CUdeviceptr>(c->mc_base), c->heap_bytes);
    d.addrFree(reinterpret_cast<CUdeviceptr>(c->mc_base), c->heap_bytes);
    c->mc_base = nullptr;
  }
  for (int r = 0; r < c->world; ++r) {
    if (r == c->rank) continue;
    if (c->peers.base[r] != nullptr) {
      d.memUnmap(reinterpret_cast<CUdeviceptr>(c->peers.base[r]), c->heap_bytes);
      d.addrFree(reinterpret_cast<CUdeviceptr>(c->peers.base[r]), c->heap_bytes);
      c->peers.base[r] = nullptr;
    }
    if (c->peer_handle[r]) {
      d.memRelease(c->peer_handle[r]);
      c->peer_handle[r] = 0;
    }
  }
  if (c->mc_handle) {
    d.memRelease(c->mc_handle);
    c->mc_handle = 0;
  }
  if (c->local_base) {
    d.memUnmap(reinterpret_cast<CUdeviceptr>(c->local_base), c->heap_bytes);
    d.addrFree(reinterpret_cast<CUdeviceptr>(c->local_base), c->heap_bytes);
    c->local_base = nullptr;
    c->peers.base[c->rank] = nullptr;
  }
  if (c->mem_handle) {
    d.memRelease(c->mem_handle);
    c->mem_handle = 0;
  }
  c->vmm = false;
}

// Every rank publishes whether `stage` worked for it and reads everybody's answer: either all ranks go on
// or all of them give up together (and fall back to the CUDA-IPC heap with their store counters still in
// step).  A rank that died instead of answering surfaces as a store timeout that names its key.
static int vmm_agree(Comm* c, const std::string& gen, const char* stage, int my_rc) {
  if (c->world == 1) return my_rc;
  char mine[kErrorBufferSize];
  strncpy(mine, error_buffer(), sizeof(mine) - 1);
  mine[sizeof(mine) - 1] = 0;
  const std::string prefix = "vmm/" + gen + "/" + stage + "/";
  const char ok = my_rc == 0 ? 1 : 0;
  int rc = b200dp_store_set(c->store, (prefix + std::to_string(c->rank)).c_str(), &ok, 1);
  if (rc) return rc;
  int failed_rank = -1;
  for (int r = 0; r < c->world; ++r) {
    char v = 0;
    int len = 0;
    rc = b200dp_store_get(c->store, (prefix + std::to_string(r)).c_str(), &v, 1, &len);
    if (rc) return rc;
    if (v == 0 && failed_rank < 0) failed_rank = r;
  }
  if (my_rc != 0) {
    set_error("%s", mine);
    return my_rc;
  }
  if (failed_rank >= 0) {
    set_error("comm_create: rank %d failed at stage '%s' of the multicast heap set-up; every rank falls back",
              failed_rank, stage);
    return B200DP_ENOSYS;
  }
  return 0;
}

struct VmmScratch {  // host resources of the handshake, closed on every path
  int lfd = -1, my_fd = -1, mc_fd = -1;
  ~VmmScratch() {
    if (lfd >= 0) ::close(lfd);
    if (my_fd >= 0) ::close(my_fd);
    if (mc_fd >= 0) ::close(mc_fd);
  }
};

static int vmm_stage_local(Comm* c, const DriverApi& d) {
  if (!d.ok) {
    set_error("comm_create: the CUDA driver lacks the VMM / multicast entry points");
    return B200DP_ENOSYS;
  }
  int mc_supported = 0;
  auto devGet = driver_fn<PFN_cuDeviceGet_v2000>("cuDeviceGet");
  auto getAttr = driver_fn<PFN_cuDeviceGetAttribute_v2000>("cuDeviceGetAttribute");
  CUdevice dv = 0;
  if (!devGet || !getAttr) return B200DP_ENOSYS;
  B200DP_CU_TRY(devGet(&dv, c->device));
  B200DP_CU_TRY(getAttr(&mc_supported, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dv));
  if (!mc_supported) {
    set_error("comm_create: device %d does not support NVSwitch multicast", c->device);
    return B200DP_ENOSYS;
  }
  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = c->device;
  B200DP_CU_TRY(d.memCreate(&c->mem_handle, c->heap_bytes, &prop, 0));
  int rc = map_handle(d, c->mem_handle, c->heap_bytes, c->device, &c->local_base);
  if (rc) return rc;
  B200DP_CUDA_TRY(cudaMemset(c->local_base, 0, c->heap_bytes));
  c->peers.base[c->rank] = c->local_base;
  c->vmm = true;
  return 0;
}

// exchange the heap descriptors (and rank 0's multicast descriptor), import and map the peers' heaps
static int vmm_stage_exchange(Comm* c, const DriverApi& d, VmmScratch& x, long long token, const std::string& gen,
                              int timeout_ms) {
  auto sock_name = [&](int r) { return "b200dp-" + std::to_string(token) + "-" + gen + "-" + std::to_string(r); };
  B200DP_CU_TRY(d.exportHandle(&x.my_fd, c->mem_handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  for (int r = 0; r < c->world; ++r)
    if (r != c->rank && fd_send(sock_name(r), x.my_fd, c->rank, timeout_ms) != 0) {
      set_error("comm_create: rank %d could not pass its heap descriptor to rank %d", c->rank, r);
      return B200DP_ETIMEDOUT;
    }
  if (c->rank == 0) {
    CUmulticastObjectProp mp{};
    mp.numDevices = c->world;
    mp.size = c->heap_bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    B200DP_CU_TRY(d.mcCreate(&c->mc_handle, &mp));
    B200DP_CU_TRY(d.exportHandle(&x.mc_fd, c->mc_handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    for (int r = 1; r < c->world; ++r)
      if (fd_send(sock_name(r), x.mc_fd, 100, timeout_ms) != 0) {
        set_error("comm_create: rank 0 could not pass the multicast descriptor to rank %d", r);
        return B200DP_ETIMEDOUT;
      }
  }
  const int expect = (c->world - 1) + (c->rank == 0 ? 0 : 1);
  for (int k = 0; k < expect; ++k) {
    int tag = -1;
    const int fd = fd_recv_timeout(x.lfd, &tag, timeout_ms);
    if (fd < 0) {
      set_error("comm_create: rank %d %s while waiting for descriptor %d of %d (a peer failed before sending?)",
                c->rank, fd == -2 ? "timed out" : "hit a socket error", k + 1, expect);
      return fd == -2 ? B200DP_ETIMEDOUT : B200DP_EINVAL;
    }
    int rc = 0;
    if (tag == 100) {
      CUresult r = d.importHandle(&c->mc_handle, reinterpret_cast<void*>(static_cast<intptr_t>(fd)),
                                  CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
      if (r != CUDA_SUCCESS) {
        set_error("comm_create: importing the multicast descriptor -> CUresult %d", static_cast<int>(r));
        rc = B200DP_EINVAL;
      }
    } else if (tag >= 0 && tag < c->world && tag != c->rank) {
      CUresult r = d.importHandle(&c->peer_handle[tag], reinterpret_cast<void*>(static_cast<intptr_t>(fd)),
                                  CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
      if (r != CUDA_SUCCESS) {
        set_error("comm_create: importing rank %d's heap descriptor -> CUresult %d", tag, static_cast<int>(r));
        rc = B200DP_EINVAL;
      } else {
        rc = map_handle(d, c->peer_handle[tag], c->heap_bytes, c->device, &c->peers.base[tag]);
      }
    }
    ::close(fd);
    if (rc) return rc;
  }
  return 0;
}

// Heap through the virtual-memory-management API so that it can be bound to an NVSwitch multicast
// object: cuMemCreate -> export as POSIX fd -> every peer imports and maps it; rank 0 creates the
// multicast object, everyone adds its device, binds its allocation and maps the multicast address.
// Every stage ends with vmm_agree(): a failure on one rank (or a rank that stops answering, bounded by the
// store / descriptor time-outs) makes ALL ranks release what they mapped and return an error together.
static int create_vmm_heap(Comm* c) {
  const DriverApi& d = driver_api();
  static int generation = 0;
  const std::string gen = std::to_string(generation++);
  const int timeout_ms = c->store ? b200dp_store_timeout_ms(c->store) : 30000;
  VmmScratch x;
  auto fail = [&](int rc) {
    release_vmm_heap(c);
    return rc;
  };

  int rc = vmm_agree(c, gen, "alloc", vmm_stage_local(c, d));
  if (rc) return fail(rc);
  if (c->world == 1) return 0;

  // ---- agree on a socket-name token, open the listening sockets
  long long token = 0;
  if (c->rank == 0) {
    token = (static_cast<long long>(getpid()) << 20) ^ static_cast<long long>(globaltimer_host());
    rc = b200dp_store_set(c->store, ("vmm/token/" + gen).c_str(), &token, sizeof(token));
    if (rc) return fail(rc);
  }
  int len = 0;
  rc = b200dp_store_get(c->store, ("vmm/token/" + gen).c_str(), &token, sizeof(token), &len);
  if (rc) return fail(rc);
  x.lfd = fd_listen("b200dp-" + std::to_string(token) + "-" + gen + "-" + std::to_string(c->rank));
  rc = 0;
  if (x.lfd < 0) {
    set_error("comm_create: rank %d cannot open its descriptor-passing socket", c->rank);
    rc = B200DP_EINVAL;
  }
  rc = vmm_agree(c, gen, "listen", rc);
  if (rc) return fail(rc);

  rc = vmm_agree(c, gen, "exchange", vmm_stage_exchange(c, d, x, token, gen, timeout_ms));
  if (rc) return fail(rc);

  // ---- multicast: add devices, then bind memory, then map (each step is collective)
  auto cu_stage = [&](const char* what, CUresult r) {
    if (r == CUDA_SUCCESS) return 0;
    set_error("comm_create: %s on rank %d -> CUresult %d", what, c->rank, static_cast<int>(r));
    return static_cast<int>(B200DP_EINVAL);
  };
  CUdevice cudev = 0;
  auto devGet = driver_fn<PFN_cuDeviceGet_v2000>("cuDeviceGet");
  rc = devGet ? cu_stage("cuDeviceGet", devGet(&cudev, c->device)) : static_cast<int>(B200DP_ENOSYS);
  if (rc == 0) rc = cu_stage("cuMulticastAddDevice", d.mcAddDevice(c->mc_handle, cudev));
  rc = vmm_agree(c, gen, "mc_add", rc);
  if (rc) return fail(rc);
  rc = vmm_agree(c, gen, "mc_bind",
                 cu_stage("cuMulticastBindMem", d.mcBindMem(c->mc_handle, 0, c->mem_handle, 0, c->heap_bytes, 0)));
  if (rc) return fail(rc);
  rc = vmm_agree(c, gen, "mc_map", map_handle(d, c->mc_handle, c->heap_bytes, c->device, &c->mc_base));
  if (rc) return fail(rc);
  return 0;
}

}  // namespace b200dp

// flags bit 0: allocate the heap with the VMM API and map an NVSwitch multicast view of it (NVLS)
extern "C" int b200dp_comm_create_ex(void* store, int rank, int world, int device, long long heap_bytes, int flags,
                                     void** out_comm) {
  if (world < 1 || world > kMaxWorld || rank < 0 || rank >= world || heap_bytes <= 0 || !out_comm) {
    set_error("comm_create: bad arguments (rank %d, world %d, heap %lld bytes; world <= %d)", rank, world,
              heap_bytes, kMaxWorld);
    return B200DP_EINVAL;
  }
  if (world > 1 && store == nullptr) {
    set_error("comm_create: a rendezvous store is required for world > 1");
    return B200DP_EINVAL;
  }
  B200DP_CUDA_TRY(cudaSetDevice(device));
  B200DP_CUDA_TRY(cudaFree(nullptr));  // make sure the primary context exists before driver-API calls
  auto* c = new Comm();
  c->rank = rank;
  c->world = world;
  c->device = device;
  c->store = store;
  c->heap_bytes = ((static_cast<size_t>(heap_bytes) + kSignalPadBytes + (2u << 20) - 1) >> 21) << 21;
  // every failure path below releases what was set up so far; the caller only sees the error code
  auto fail = [&](int rc) {
    for (int r = 0; r < c->world; ++r)
      if (c->ipc_opened[r]) cudaIpcCloseMemHandle(c->peers.base[r]);
    if (c->vmm) release_vmm_heap(c);
    else if (c->local_base) cudaFree(c->local_base);
    if (c->epochs) cudaFree(c->epochs);
    delete c;
    return rc;
  };
  auto cuda_rc = [](cudaError_t e, const char* what) {
    if (e == cudaSuccess) return 0;
    set_error("comm_create: %s -> %s", what, cudaGetErrorString(e));
    return static_cast<int>(e);
  };
  int rc = cuda_rc(cudaMalloc(reinterpret_cast<void**>(&c->epochs), 2 * kNumChannels * sizeof(uint32_t)), "cudaMalloc");
  if (rc) return fail(rc);
  rc = cuda_rc(cudaMemset(c->epochs, 0, 2 * kNumChannels * sizeof(uint32_t)), "cudaMemset");
  if (rc) return fail(rc);
  c->done_blocks = c->epochs + kNumChannels;
  if (flags & 1) {
    rc = create_vmm_heap(c);   // releases its own mappings on failure
    if (rc) return fail(rc);
    rc = cuda_rc(cudaDeviceSynchronize(), "cudaDeviceSynchronize");
    if (rc) return fail(rc);
    if (world > 1) {
      rc = b200dp_store_barrier(store, "comm_create");
      if (rc) return fail(rc);
    }
    *out_comm = c;
    return 0;
  }
  rc = cuda_rc(cudaMalloc(reinterpret_cast<void**>(&c->local_base), c->heap_bytes), "cudaMalloc(heap)");
  if (rc) return fail(rc);
  rc = cuda_rc(cudaMemset(c->local_base, 0, c->heap_bytes), "cudaMemset(heap)");
  if (rc == 0) rc = cuda_rc(cudaDeviceSynchronize(), "cudaDeviceSynchronize");
  if (rc) return fail(rc);
  c->peers.base[rank] = c->local_base;
  if (world > 1) {
    cudaIpcMemHandle_t mine;
    rc = cuda_rc(cudaIpcGetMemHandle(&mine, c->local_base), "cudaIpcGetMemHandle");
    if (rc) return fail(rc);
    static int generation = 0;
    const std::string prefix = "heap/" + std::to_string(generation++) + "/";
    rc = b200dp_store_set(store, (prefix + std::to_string(rank)).c_str(), &mine, sizeof(mine));
    if (rc) return fail(rc);
    for (int r = 0; r < world; ++r) {
      if (r == rank) continue;
      cudaIpcMemHandle_t h;
      int len = 0;
      rc = b200dp_store_get(store, (prefix + std::to_string(r)).c_str(), &h, sizeof(h), &len);
      if (rc) return fail(rc);
      void* p = nullptr;
      cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) {
        set_error("comm_create: rank %d cannot map rank %d's heap over CUDA IPC: %s", rank, r,
                  cudaGetErrorString(e));
        return fail(static_cast<int>(e));
      }
      c->peers.base[r] = static_cast<char*>(p);
      c->ipc_opened[r] = true;
    }
    rc = b200dp_store_barrier(store, "comm_create");
    if (rc) return fail(rc);
  }
  *out_comm = c;
  return 0;
}

extern "C" int b200dp_comm_create(void* store, int rank, int world, int device, long long heap_bytes,
                                  void** out_comm) {
  return b200dp_comm_create_ex(store, rank, world, device, heap_bytes, 0, out_comm);
}

/* Make "algo 0" choose the in-switch (NVLS) kernel for ranges above 256 KiB. */
extern "C" int b200dp_comm_prefer_nvls(void* comm, int on) {
  auto* c = static_cast<Comm*>(comm);
  if (!c) return B200DP_EINVAL;
  if (on && c->mc_base == nullptr) {
    set_error("comm_prefer_nvls: this communicator has no multicast mapping");
    return B200DP_EINVAL;
  }
  c->prefer_nvls = on != 0;
  return 0;
}

/* 1 if the communicator has an NVSwitch multicast mapping of its heap (NVLS collectives usable). */
extern "C" int b200dp_comm_has_multicast(void* comm) {
  auto* c = static_cast<Comm*>(comm);
  return (c != nullptr && c->mc_base != nullptr) ? 1 : 0;
}

extern "C" int b200dp_comm_destroy(void* comm) {
  auto* c = static_cast<Comm*>(comm);
  if (!c) return 0;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  if (c->world > 1 && c->store) b200dp_store_barrier(c->store, "comm_destroy");
  for (int r = 0; r < c->world; ++r)
    if (c->ipc_opened[r]) cudaIpcCloseMemHandle(c->peers.base[r]);
  if (c->vmm) {
    // peers unmap each other's heaps before anybody releases its own allocation
    const DriverApi& d = driver_api();
    if (c->mc_base) {
      d.memUnmap(reinterpret_cast<CUdeviceptr>(c->mc_base), c->heap_bytes);
      d.addrFree(reinterpret_cast<CUdeviceptr>(c->mc_base), c->heap_bytes);
      c->mc_base = nullptr;
    }
    for (int r = 0; r < c->world; ++r) {
      if (r == c->rank || c->peers.base[r] == nullptr) continue;
      d.memUnmap(reinterpret_cast<CUdeviceptr>(c->peers.base[r]), c->heap_bytes);
      d.addrFree(reinterpret_cast<CUdeviceptr>(c->peers.base[r]), c->heap_bytes);
      c->peers.base[r] = nullptr;
    }
    if (c->world > 1 && c->store) b200dp_store_barrier(c->store, "comm_destroy2");
    release_vmm_heap(c);
    cudaFree(c->epochs);
    delete c;
    return 0;
  }
  if (c->world > 1 && c->store) b200dp_store_barrier(c->store, "comm_destroy2");
  cudaFree(c->local_base);
  cudaFree(c->epochs);
  delete c;
  return 0;
}

/* Local address of the first usable byte of the heap (after the signal pad) and its size. */
extern "C" int b200dp_comm_heap(void* comm, void** base, long long* usable_bytes, long long* data_offset) {
  auto* c = static_cast<Comm*>(comm);
  if (!c) return B200DP_EINVAL;
  if (base) *base = c->local_base;
  if (data_offset) *data_offset = static_cast<long long>(kSignalPadBytes);
  if (usable_bytes) *usable_bytes = static_cast<long long>(c->heap_bytes - kSignalPadBytes);
  return 0;
}

extern "C" int b200dp_comm_peer_base(void* comm, int rank, void** base) {
  auto* c = static_cast<Comm*>(comm);
  if (!c || rank < 0 || rank >= c->world) return B200DP_EINVAL;
  *base = c->peers.base[rank];
  return 0;
}

extern "C" int b200dp_comm_set_timeout(void* comm, double seconds) {
  auto* c = static_cast<Comm*>(comm);
  if (!c || seconds <= 0) return B200DP_EINVAL;
  c->timeout_ns = static_cast<unsigned long long>(seconds * 1e9);
  return 0;
}

extern "C" int b200dp_comm_barrier(void* comm, int channel, b200dp_stream_t stream) {
  auto* c = static_cast<Comm*>(comm);
  int rc = check_range(c, 0, 0, 1, channel, "comm_barrier");
  if (rc) return rc;
  barrier_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(dev_view(c), channel);
  B200DP_CUDA_TRY(cudaGetLastError());
  return 0;
}

// algo: 0 = choose, 1 = one-shot, 2 = two-shot
static int allreduce_impl(void* comm, long long grad_off, long long param_off, long long mom_off, long long start,
                          long long end, int fuse_sgd, int prescale, float lr, float momentum, float weight_decay,
                          int algo, int blocks, int channel, b200dp_stream_t stream_, bool skip_exit_barrier) {
  auto* c = static_cast<Comm*>(comm);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int threads = 512;
  const long n4 = (end - start) >> 2;
  if (algo == 3 && c != nullptr && c->mc_base == nullptr) {
    set_error("allreduce: algo 3 (NVLS) needs a communicator created with the multicast heap");
    return B200DP_EINVAL;
  }
  if (algo == 0) algo = (n4 * 16 <= 256 * 1024) ? 1 : ((c != nullptr && c->prefer_nvls && c->mc_base) ? 3 : 2);
  const int requested_blocks = blocks;
  if (algo == 1) {
    blocks = static_cast<int>((n4 + threads - 1) / threads);
    if (blocks < 1) blocks = 1;
    if (blocks > kMaxBlocks) algo = 2;  // does not fit in one pass: use the two-shot kernel
  }
  if (algo == 2 || algo == 3) blocks = requested_blocks > 0 ? requested_blocks : 64;
  int rc = check_range(c, start, end, blocks, channel, "allreduce");
  if (rc) return rc;
  RangeArgs a{};
  a.grad_off = static_cast<size_t>(grad_off);
  a.param_off = static_cast<size_t>(param_off);
  a.mom_off = static_cast<size_t>(mom_off);
  a.start = start;
  a.end = end;
  a.channel = channel;
  a.sgd = SgdHyper{lr, momentum, weight_decay};
  const CommDev d = dev_view(c);
#define B200DP_LAUNCH_AR(KERN)                                                       \
  do {                                                                               \
    if (fuse_sgd && prescale) KERN<true, true><<<blocks, threads, 0, stream>>>(d, a); \
    else if (fuse_sgd) KERN<true, false><<<blocks, threads, 0, stream>>>(d, a);       \
    else if (prescale) KERN<false, true><<<blocks, threads, 0, stream>>>(d, a);       \
    else KERN<false, false><<<blocks, threads, 0, stream>>>(d, a);                    \
  } while (0)
  if (algo == 3 && skip_exit_barrier) {
    if (fuse_sgd)
      allreduce_nvls_lazy_kernel<true><<<blocks, threads, 0, stream>>>(d, a);
    else
      allreduce_nvls_lazy_kernel<false><<<blocks, threads, 0, stream>>>(d, a);
  } else if (algo == 3) {
    if (fuse_sgd)
      allreduce_nvls_kernel<true><<<blocks, threads, 0, stream>>>(d, a);
    else
      allreduce_nvls_kernel<false><<<blocks, threads, 0, stream>>>(d, a);
  } else if (algo == 1)
    B200DP_LAUNCH_AR(allreduce_oneshot_kernel);
  else if (skip_exit_barrier)
    B200DP_LAUNCH_AR(allreduce_twoshot_lazy_kernel);
  else
    B200DP_LAUNCH_AR(allreduce_twoshot_kernel);
#undef B200DP_LAUNCH_AR
  B200DP_CUDA_TRY(cudaGetLastError());
  return 0;
}

extern "C" int b200dp_allreduce(void* comm, long long grad_off, long long param_off, long long mom_off,
                                long long start, long long end, int fuse_sgd, int prescale, float lr,
                                float momentum, float weight_decay, int algo, int blocks, int channel,
                                b200dp_stream_t stream_) {
  return allreduce_impl(comm, grad_off, param_off, mom_off, start, end, fuse_sgd, prescale, lr, momentum,
                        weight_decay, algo, blocks, channel, stream_, false);
}

/* Owner-side half of the fused wgrad -> reduce-scatter path (optional; see conv_wgrad_push):
 * for the bucket [start, end): mean gradient of my slice = (own gradient + W-1 staged copies) / W inside
 * [pushed_lo, pushed_hi), pulled from the peers elsewhere; SGD; updated parameters to every rank.
 * stage_buf_off: byte offset of the staging buffer in the heap, stage_base: element offset of this
 * bucket's area in it (source rank s at stage_base + s * per4 * 4).  nvls = 1 broadcasts the parameters
 * with multimem.st (multicast heap required). */
extern "C" int b200dp_reduce_staged_sgd(void* comm, long long grad_off, long long param_off, long long mom_off,
                                        long long stage_buf_off, long long start, long long end,
                                        long long pushed_lo, long long pushed_hi, long long stage_base, float lr,
                                        float momentum, float weight_decay, int nvls, int blocks, int channel,
                                        b200dp_stream_t stream_) {
  auto* c = static_cast<Comm*>(comm);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (blocks <= 0) blocks = 32;
  int rc = check_range(c, start, end, blocks, channel, "reduce_staged_sgd");
  if (rc) return rc;
  if ((pushed_lo & 3) || (pushed_hi & 3) || (stage_base & 3) || pushed_lo < start || pushed_hi > end ||
      pushed_lo > pushed_hi) {
    set_error("reduce_staged_sgd: pushed range [%lld, %lld) must be 4-aligned and inside [%lld, %lld)", pushed_lo,
              pushed_hi, start, end);
    return B200DP_EINVAL;
  }
  if (nvls && c->mc_base == nullptr) {
    set_error("reduce_staged_sgd: nvls=1 needs a communicator created with the multicast heap");
    return B200DP_EINVAL;
  }
  StagedArgs s{};
  s.r.grad_off = static_cast<size_t>(grad_off);
  s.r.param_off = static_cast<size_t>(param_off);
  s.r.mom_off = static_cast<size_t>(mom_off);
  s.r.start = start;
  s.r.end = end;
  s.r.channel = channel;
  s.r.sgd = SgdHyper{lr, momentum, weight_decay};
  s.stage_buf_off = static_cast<size_t>(stage_buf_off);
  s.pushed_lo = pushed_lo;
  s.pushed_hi = pushed_hi;
  s.stage_base = stage_base;
  const CommDev d = dev_view(c);
  if (nvls)
    reduce_staged_sgd_kernel<true><<<blocks, 512, 0, stream>>>(d, s);
  else
    reduce_staged_sgd_kernel<false><<<blocks, 512, 0, stream>>>(d, s);
  B200DP_CUDA_TRY(cudaGetLastError());
  return 0;
}

extern "C" int b200dp_ps_reduce_bcast(void* comm, long long grad_off, long long stage_off, long long start,
                                      long long end, int root, int push, int blocks, int channel,
                                      b200dp_stream_t stream_) {
  auto* c = static_cast<Comm*>(comm);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (blocks <= 0) blocks = 64;
  int rc = check_range(c, start, end, blocks, channel, "ps_reduce_bcast");
  if (rc) return rc;
  if (root < 0 || root >= c->world) {
    set_error("ps_reduce_bcast: root %d outside the world of %d", root, c->world);
    return B200DP_EINVAL;
  }
  RangeArgs a{};
  a.grad_off = static_cast<size_t>(grad_off);
  a.stage_off = static_cast<size_t>(stage_off);
  a.start = start;
  a.end = end;
  a.channel = channel;
  a.root = root;
  const CommDev d = dev_view(c);
  if (push)
    ps_reduce_bcast_kernel<true><<<blocks, 512, 0, stream>>>(d, a);
  else
    ps_reduce_bcast_kernel<false><<<blocks, 512, 0, stream>>>(d, a);
  B200DP_CUDA_TRY(cudaGetLastError());
  return 0;
}

extern "C" int b200dp_broadcast(void* comm, long long byte_off, long long nbytes, int root, int blocks,
                                int channel, b200dp_stream_t stream_) {
  auto* c = static_cast<Comm*>(comm);
  if (blocks <= 0) blocks = nbytes > (1 << 20) ? 64 : 4;
  if (nbytes & 15) {
    set_error("broadcast: nbytes=%lld must be a multiple of 16", nbytes);
    return B200DP_EINVAL;
  }
  int rc = check_range(c, 0, nbytes / 4, blocks, channel, "broadcast");
  if (rc) return rc;
  RangeArgs a{};
  a.grad_off = static_cast<size_t>(byte_off);
  a.start = 0;
  a.end = nbytes / 4;
  a.channel = channel;
  a.root = root;
  broadcast_kernel<<<blocks, 512, 0, static_cast<cudaStream_t>(stream_)>>>(dev_view(c), a);
  B200DP_CUDA_TRY(cudaGetLastError());
  return 0;
}

extern "C" int b200dp_allgather_slices(void* comm, long long byte_off, long long start, long long end, int blocks,
                                       int channel, b200dp_stream_t stream_) {
  auto* c = static_cast<Comm*>(comm);
  if (blocks <= 0) blocks = (end - start) > (1 << 18) ? 32 : 4;
  int rc = check_range(c, start, end, blocks, channel, "allgather_slices");
  if (rc) return rc;
  RangeArgs a{};
  a.grad_off = static_cast<size_t>(byte_off);
  a.start = start;
  a.end = end;
  a.channel = channel;
  allgather_slices_kernel<<<blocks, 512, 0, static_cast<cudaStream_t>(stream_)>>>(dev_view(c), a);
  B200DP_CUDA_TRY(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------
// Bucket reducer: the framework-owned counterpart of DDP's C++ Reducer
// (ref: master/part3/part3.py:116).  Gradients already live in the flat
// symmetric buffer, so "mark ready" is just: record an event on the compute stream, make the
// high-priority side stream wait on it, and launch that bucket's fused all-reduce(+SGD) there
// with a capped number of CTAs so that the backward GEMMs keep most of the SMs.  finalize()
// joins the side stream back into the compute stream.  All of it is stream-ordered and
// capturable in a CUDA graph (fork = event record + wait, join = the reverse).
// ---------------------------------------------------------------------------
namespace b200dp {
struct Reducer {
  Comm* comm = nullptr;
  std::vector<long long> start, end;
  long long grad_off = 0, param_off = 0, mom_off = 0;
  int fuse_sgd = 0, prescale = 1, blocks = 16, channel = 1;
  float lr = 0.1f, momentum = 0.9f, weight_decay = 1e-4f;
  cudaStream_t side = nullptr;
  std::vector<cudaEvent_t> ready;
  cudaEvent_t done = nullptr;
  int launched = 0;
  // optional path: buckets whose conv weight gradient arrives through conv_wgrad_push staging
  // only the last bucket kernel of a step keeps its exit barrier (the entry barrier of the next bucket
  // kernel on the channel implies it); bucket_ready() must be called in index order
  bool lazy_exit = true;
  bool staged = false;
  long long stage_buf_off = 0;
  std::vector<long long> pushed_lo, pushed_hi, stage_base;
};
}  // namespace b200dp

extern "C" int b200dp_reducer_create(void* comm, int n_buckets, const long long* starts, const long long* ends,
                                     long long grad_off, long long param_off, long long mom_off, int fuse_sgd,
                                     int prescale, float lr, float momentum, float weight_decay, int blocks,
                                     int channel, void** out) {
  if (!comm || n_buckets < 1 || !starts || !ends || !out) {
    set_error("reducer_create: bad arguments");
    return B200DP_EINVAL;
  }
  auto* r = new Reducer();
  r->comm = static_cast<Comm*>(comm);
  r->start.assign(starts, starts + n_buckets);
  r->end.assign(ends, ends + n_buckets);
  r->grad_off = grad_off;
  r->param_off = param_off;
  r->mom_off = mom_off;
  r->fuse_sgd = fuse_sgd;
  r->prescale = prescale;
  r->lr = lr;
  r->momentum = momentum;
  r->weight_decay = weight_decay;
  r->blocks = blocks > 0 ? blocks : 16;
  r->channel = channel;
  int lo = 0, hi = 0;
  B200DP_CUDA_TRY(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  B200DP_CUDA_TRY(cudaStreamCreateWithPriority(&r->side, cudaStreamNonBlocking, hi));
  r->ready.resize(n_buckets);
  for (auto& e : r->ready) B200DP_CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  B200DP_CUDA_TRY(cudaEventCreateWithFlags(&r->done, cudaEventDisableTiming));
  *out = r;
  return 0;
}

extern "C" int b200dp_reducer_set_lr(void* reducer, float lr) {
  auto* r = static_cast<Reducer*>(reducer);
  if (!r) return B200DP_EINVAL;
  r->lr = lr;
  return 0;
}

/* The reducer's side stream (a cudaStream_t), so that the caller can queue work behind a bucket's
 * collective -- e.g. the bf16 re-layout of the weights that bucket just updated. */
extern "C" int b200dp_reducer_stream(void* reducer, void** stream_out) {
  auto* r = static_cast<Reducer*>(reducer);
  if (!r || !stream_out) return B200DP_EINVAL;
  *stream_out = r->side;
  return 0;
}

/* CTAs per bucket collective from the next bucket_ready() on (takes effect when the step is re-captured). */
extern "C" int b200dp_reducer_set_blocks(void* reducer, int blocks) {
  auto* r = static_cast<Reducer*>(reducer);
  if (!r || blocks < 1) {
    set_error("reducer_set_blocks: need a reducer and blocks >= 1");
    return B200DP_EINVAL;
  }
  r->blocks = blocks;
  return 0;
}

extern "C" int b200dp_reducer_bucket_ready(void* reducer, int k, b200dp_stream_t compute_stream) {
  auto* r = static_cast<Reducer*>(reducer);
  if (!r || k < 0 || k >= static_cast<int>(r->ready.size())) {
    set_error("reducer_bucket_ready: bucket %d out of range", k);
    return B200DP_EINVAL;
  }
  B200DP_CUDA_TRY(cudaEventRecord(r->ready[k], static_cast<cudaStream_t>(compute_stream)));
  B200DP_CUDA_TRY(cudaStreamWaitEvent(r->side, r->ready[k], 0));
  ++r->launched;
  if (r->staged && r->pushed_hi[k] > r->pushed_lo[k])
    return b200dp_reduce_staged_sgd(r->comm, r->grad_off, r->param_off, r->mom_off, r->stage_buf_off, r->start[k],
                                    r->end[k], r->pushed_lo[k], r->pushed_hi[k], r->stage_base[k], r->lr, r->momentum,
                                    r->weight_decay, (r->comm->prefer_nvls && r->comm->mc_base) ? 1 : 0, r->blocks,
                                    r->channel, r->side);
  const bool skip_exit = r->lazy_exit && k != static_cast<int>(r->ready.size()) - 1;
  return allreduce_impl(r->comm, r->grad_off, r->param_off, r->mom_off, r->start[k], r->end[k], r->fuse_sgd,
                        r->prescale, r->lr, r->momentum, r->weight_decay, 0, r->blocks, r->channel, r->side, skip_exit);
}

/* Bucket kernels 0 .. n-2 of a step run without their exit barrier (the entry barrier of the next bucket
 * kernel on the channel implies it; on by default); bucket_ready() must then be called in index order, which
 * is the gradient-ready order the executor uses.  on = 0 gives every bucket kernel both barriers. */
extern "C" int b200dp_reducer_set_lazy_exit(void* reducer, int on) {
  auto* r = static_cast<Reducer*>(reducer);
  if (!r) return B200DP_EINVAL;
  r->lazy_exit = on != 0;
  return 0;
}

/* Optional path: tell the reducer which part of every bucket is delivered by conv_wgrad_push
 * (pushed_lo[k] == pushed_hi[k]: none, the bucket keeps the plain all-reduce).  Needs fuse_sgd. */
extern "C" int b200dp_reducer_set_staged(void* reducer, long long stage_buf_off, const long long* pushed_lo,
                                         const long long* pushed_hi, const long long* stage_base) {
  auto* r = static_cast<Reducer*>(reducer);
  if (!r || !pushed_lo || !pushed_hi || !stage_base || !r->fuse_sgd) {
    set_error("reducer_set_staged: needs a reducer with the fused optimizer and three arrays of n_buckets entries");
    return B200DP_EINVAL;
  }
  const size_t n = r->start.size();
  r->pushed_lo.assign(pushed_lo, pushed_lo + n);
  r->pushed_hi.assign(pushed_hi, pushed_hi + n);
  r->stage_base.assign(stage_base, stage_base + n);
  r->stage_buf_off = stage_buf_off;
  r->staged = true;
  return 0;
}

extern "C" int b200dp_reducer_finalize(void* reducer, b200dp_stream_t compute_stream) {
  auto* r = static_cast<Reducer*>(reducer);
  if (!r) return B200DP_EINVAL;
  B200DP_CUDA_TRY(cudaEventRecord(r->done, r->side));
  B200DP_CUDA_TRY(cudaStreamWaitEvent(static_cast<cudaStream_t>(compute_stream), r->done, 0));
  r->launched = 0;
  return 0;
}

extern "C" int b200dp_reducer_destroy(void* reducer) {
  auto* r = static_cast<Reducer*>(reducer);
  if (!r) return 0;
  cudaStreamSynchronize(r->side);
  for (auto& e : r->ready) cudaEventDestroy(e);
  cudaEventDestroy(r->done);
  cudaStreamDestroy(r->side);
  delete r;
  return 0;
}
