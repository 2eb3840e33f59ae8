// Native batch loader: the role torch's C++ DataLoader workers play for the reference
// (ref: master/part2a/part2a.py:103-118 -- DataLoader(num_workers=2) over torchvision CIFAR10 with
// RandomCrop(32, padding=4), RandomHorizontalFlip, ToTensor, Normalize).
//
// A ~0.7 ms training step consumes 365 k images/s per GPU; the tutorial's per-sample Python transforms
// deliver a few thousand, a vectorised torch pipeline ~20 k.  Here a small pool of threads turns
// uint8 HWC images straight into normalised fp32 NCHW batches inside caller-provided (pinned) buffers:
//   * one pass per image: crop offset, flip, uint8 -> float and normalisation are a table lookup
//     (256 x 3 entries computed as ((v / 255) - mean) / std in float, i.e. bit-identical to
//     ToTensor + Normalize), written channel by channel;
//   * augmentation parameters come from a counter-based generator keyed by (seed, epoch, position in
//     the epoch), so a batch does not depend on which thread built it or on the thread count;
//   * batches are produced in order into a ring of slots; the consumer takes them in order with
//     next() and hands a slot back with release() once its host->device copy has completed.
// The sample order of an epoch is supplied by the caller (ShardedSampler = DistributedSampler
// semantics stay in one place, data.py).  Host-only code: tested on CPU.
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>

#include "../include/b200dp.h"
#include "host_util.h"

namespace b200dp {

constexpr int kImg = 32, kChan = 3, kPad = 4;
constexpr int kPixels = kImg * kImg;

struct Loader {
  const uint8_t* images = nullptr;  // [n][32][32][3]
  const int64_t* labels = nullptr;
  long long n = 0;
  int batch = 0, ring = 0, augment = 0;
  unsigned long long seed = 0;
  std::vector<float*> xs;    // ring x [batch][3][32][32]
  std::vector<int64_t*> ys;  // ring x [batch]
  float lut[kChan][256];

  // epoch state (guarded by mu)
  std::mutex mu;
  std::condition_variable cv;
  std::vector<int64_t> order;       // sample indices of this epoch, in consumption order
  long long epoch = 0;
  long long num_batches = 0;
  long long next_to_build = 0;      // next batch index a worker may claim
  long long next_to_take = 0;       // next batch index the consumer gets
  std::vector<long long> slot_batch;  // which batch a slot holds (-1: free)
  std::vector<char> slot_ready;
  unsigned long long generation = 0;  // bumped by start_epoch / shutdown to wake and reset workers
  int in_flight = 0;                  // workers currently writing into a slot
  bool paused = false;                // start_epoch is waiting for them: no new claims
  bool stop = false;
  std::vector<std::thread> workers;
};

static inline uint64_t mix64(uint64_t x) {  // splitmix64 finaliser
  x += 0x9e3779b97f4a7c15ull;
  x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
  x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
  return x ^ (x >> 31);
}

// crop offsets in [0, 8] (position of the 32x32 window inside the 40x40 zero-padded image) and flip
static inline void aug_params(const Loader& L, long long epoch, long long pos, int* di, int* dj, int* flip) {
  const uint64_t r = mix64(mix64(L.seed ^ (static_cast<uint64_t>(epoch) << 40)) + static_cast<uint64_t>(pos));
  *di = static_cast<int>(r % 9);
  *dj = static_cast<int>((r >> 16) % 9);
  *flip = static_cast<int>((r >> 40) & 1);
}

static void build_sample(const Loader& L, const uint8_t* src, float* dst, int di, int dj, int flip) {
  // dst[c][h][w] = lut[c][ padded[h + di][w' + dj] ],  padded = image at (4,4) in a 40x40 zero canvas,
  // w' = flip ? 31 - w : w  (flip of the cropped window, as RandomHorizontalFlip after RandomCrop)
  for (int h = 0; h < kImg; ++h) {
    const int sh = h + di - kPad;
    float* d0 = dst + h * kImg;
    float* d1 = d0 + kPixels;
    float* d2 = d1 + kPixels;
    if (sh < 0 || sh >= kImg) {
      for (int w = 0; w < kImg; ++w) {
        d0[w] = L.lut[0][0];
        d1[w] = L.lut[1][0];
        d2[w] = L.lut[2][0];
      }
      continue;
    }
    const uint8_t* row = src + sh * kImg * kChan;
    for (int w = 0; w < kImg; ++w) {
      const int sw = (flip ? (kImg - 1 - w) : w) + dj - kPad;
      if (sw < 0 || sw >= kImg) {
        d0[w] = L.lut[0][0];
        d1[w] = L.lut[1][0];
        d2[w] = L.lut[2][0];
      } else {
        const uint8_t* px = row + sw * kChan;
        d0[w] = L.lut[0][px[0]];
        d1[w] = L.lut[1][px[1]];
        d2[w] = L.lut[2][px[2]];
      }
    }
  }
}

static void worker_main(Loader* L) {
  std::unique_lock<std::mutex> lk(L->mu);
  for (;;) {
    // claim the next batch whose slot is free
    long long b = -1;
    int slot = -1;
    unsigned long long gen = L->generation;
    L->cv.wait(lk, [&] {
      if (L->stop) return true;
      if (L->paused || L->next_to_build >= L->num_batches) return false;
      const int s = static_cast<int>(L->next_to_build % L->ring);
      return L->slot_batch[s] < 0;
    });
    if (L->stop) return;
    b = L->next_to_build++;
    slot = static_cast<int>(b % L->ring);
    L->slot_batch[slot] = b;
    L->slot_ready[slot] = 0;
    ++L->in_flight;
    gen = L->generation;
    const long long epoch = L->epoch;
    const long long first = b * L->batch;
    const long long count = std::min<long long>(L->batch, static_cast<long long>(L->order.size()) - first);
    const int64_t* idx = L->order.data() + first;
    lk.unlock();

    float* x = L->xs[slot];
    int64_t* y = L->ys[slot];
    for (long long i = 0; i < count; ++i) {
      const int64_t s = idx[i];
      int di = kPad, dj = kPad, flip = 0;  // identity: window at (4,4), no flip
      if (L->augment) aug_params(*L, epoch, first + i, &di, &dj, &flip);
      build_sample(*L, L->images + s * (kPixels * kChan), x + i * (kChan * kPixels), di, dj, flip);
      y[i] = L->labels[s];
    }

    lk.lock();
    --L->in_flight;
    if (gen == L->generation) L->slot_ready[slot] = 1;  // otherwise the epoch was restarted: stale result
    L->cv.notify_all();
  }
}

}  // namespace b200dp

using namespace b200dp;

/* images: uint8 [n][32][32][3] (HWC), labels: int64 [n]; both must outlive the loader.  xs / ys: `ring`
 * caller-owned output buffers of batch*3*32*32 floats / batch int64 each (pinned memory for the GPU path).
 * mean / std: per-channel normalisation constants (of the [0,1]-scaled image).  augment: RandomCrop(32,
 * padding 4) + RandomHorizontalFlip.  threads >= 1, ring >= 2. */
extern "C" int b200dp_loader_create(const void* images, const long long* labels, long long n, int batch, int ring,
                                    void* const* xs, void* const* ys, const float* mean, const float* std_,
                                    int augment, unsigned long long seed, int threads, void** out) {
  if (!images || !labels || n <= 0 || batch <= 0 || ring < 2 || !xs || !ys || !mean || !std_ || threads < 1 || !out) {
    set_error("loader_create: bad arguments (n=%lld batch=%d ring=%d threads=%d)", n, batch, ring, threads);
    return B200DP_EINVAL;
  }
  auto* L = new Loader();
  L->images = static_cast<const uint8_t*>(images);
  L->labels = reinterpret_cast<const int64_t*>(labels);
  L->n = n;
  L->batch = batch;
  L->ring = ring;
  L->augment = augment;
  L->seed = seed;
  for (int i = 0; i < ring; ++i) {
    if (!xs[i] || !ys[i]) {
      delete L;
      set_error("loader_create: output buffer %d is NULL", i);
      return B200DP_EINVAL;
    }
    L->xs.push_back(static_cast<float*>(xs[i]));
    L->ys.push_back(static_cast<int64_t*>(ys[i]));
  }
  for (int c = 0; c < kChan; ++c)
    for (int v = 0; v < 256; ++v) L->lut[c][v] = ((static_cast<float>(v) / 255.0f) - mean[c]) / std_[c];
  L->slot_batch.assign(ring, -1);
  L->slot_ready.assign(ring, 0);
  for (int t = 0; t < threads; ++t) L->workers.emplace_back(worker_main, L);
  *out = L;
  return 0;
}

/* Start (or restart) an epoch: `order` lists the `count` sample indices in consumption order (already
 * sharded / shuffled by the caller); batches of `batch` samples are built ahead into the ring.  Slots still
 * held from an abandoned epoch are reclaimed: the caller must be done with their contents.  drop_last: do
 * not build the partial last batch. */
extern "C" int b200dp_loader_start_epoch(void* loader, const long long* order, long long count, long long epoch,
                                         int drop_last) {
  auto* L = static_cast<Loader*>(loader);
  if (!L || !order || count < 0) {
    set_error("loader_start_epoch: bad arguments");
    return B200DP_EINVAL;
  }
  for (long long i = 0; i < count; ++i) {
    if (order[i] < 0 || order[i] >= L->n) {
      set_error("loader_start_epoch: sample index %lld at position %lld is outside [0, %lld)", order[i], i, L->n);
      return B200DP_EINVAL;
    }
  }
  std::unique_lock<std::mutex> lk(L->mu);
  // nobody may be writing into a slot while the ring is reset (an abandoned epoch can still have builds
  // in progress): stop new claims, wait for the builds under way, then restart
  L->paused = true;
  ++L->generation;
  L->cv.wait(lk, [&] { return L->in_flight == 0; });
  L->order.assign(reinterpret_cast<const int64_t*>(order), reinterpret_cast<const int64_t*>(order) + count);
  L->epoch = epoch;
  L->num_batches = drop_last ? count / L->batch : (count + L->batch - 1) / L->batch;
  L->next_to_build = 0;
  L->next_to_take = 0;
  std::fill(L->slot_batch.begin(), L->slot_batch.end(), -1);
  std::fill(L->slot_ready.begin(), L->slot_ready.end(), 0);
  L->paused = false;
  L->cv.notify_all();
  return 0;
}

/* Block until the next batch (in order) is complete.  *slot: which output buffer holds it, *count: samples
 * in it (the last batch may be partial).  Returns 1 when the epoch is exhausted (nothing returned). */
extern "C" int b200dp_loader_next(void* loader, int* slot, int* count) {
  auto* L = static_cast<Loader*>(loader);
  if (!L || !slot || !count) return B200DP_EINVAL;
  std::unique_lock<std::mutex> lk(L->mu);
  if (L->next_to_take >= L->num_batches) return 1;
  const long long b = L->next_to_take;
  const int s = static_cast<int>(b % L->ring);
  L->cv.wait(lk, [&] { return L->slot_batch[s] == b && L->slot_ready[s]; });
  ++L->next_to_take;
  *slot = s;
  const long long first = b * L->batch;
  *count = static_cast<int>(std::min<long long>(L->batch, static_cast<long long>(L->order.size()) - first));
  return 0;
}

/* Hand a slot back once its contents are no longer needed (its host->device copy has completed). */
extern "C" int b200dp_loader_release(void* loader, int slot) {
  auto* L = static_cast<Loader*>(loader);
  if (!L || slot < 0 || slot >= L->ring) {
    set_error("loader_release: slot %d out of range", slot);
    return B200DP_EINVAL;
  }
  std::lock_guard<std::mutex> lk(L->mu);
  L->slot_batch[slot] = -1;
  L->slot_ready[slot] = 0;
  L->cv.notify_all();
  return 0;
}

extern "C" int b200dp_loader_destroy(void* loader) {
  auto* L = static_cast<Loader*>(loader);
  if (!L) return 0;
  {
    std::lock_guard<std::mutex> lk(L->mu);
    L->stop = true;
    ++L->generation;
    L->cv.notify_all();
  }
  for (auto& t : L->workers) t.join();
  delete L;
  return 0;
}
