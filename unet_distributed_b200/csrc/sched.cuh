// Dynamic tile scheduler for the persistent forward / dgrad kernels (igemm, conv_halo, conv_halo_rp).
//
// Those kernels run one CTA per SM and used to walk tiles blockIdx.x, blockIdx.x + gridDim.x, ...  A CTA that starts
// late -- its SM still holds a block of a kernel from another stream: a weight gradient, a bucket's Adam / re-pack, an
// NCCL all-reduce -- then still owns its full share of tiles and the whole kernel ends late by that delay (a CTA that
// cannot become resident until another CTA of the same kernel exits doubles the kernel's time). Here the producer
// thread of a CTA draws tile indices from a global counter and hands them to the CTA's other roles through a small
// shared-memory ring guarded by mbarriers, so a late CTA simply takes fewer tiles. Output tiles are independent:
// results do not depend on which CTA computes which tile.
//
// Counter protocol: ctr[0] = next tile, ctr[1] = CTAs that have drawn their last (out-of-range) index. Both are zero
// before the launch; the CTA that brings ctr[1] to gridDim.x resets both, so the slot is ready for its next launch.
#pragma once
#include <cstdint>

#include "ptx.cuh"

namespace ub {

constexpr int kSchedDepth = 4;

struct __align__(8) TileSched {
  uint64_t full[kSchedDepth];
  uint64_t empty[kSchedDepth];
  int ids[kSchedDepth];
};

// Per-thread cursor into the ring (every role keeps its own).
struct SchedCursor {
  int slot = 0;
  uint32_t phase = 0;
  int static_tile = -1;   // static mode: last tile handed out
  __device__ __forceinline__ void advance() {
    if (++slot == kSchedDepth) { slot = 0; phase ^= 1u; }
  }
};

// thread 0 of the CTA, before the CTA-wide barrier that publishes the other mbarriers
__device__ __forceinline__ void sched_init(TileSched* s, int consumer_arrivals) {
  for (int i = 0; i < kSchedDepth; ++i) {
    mbar_init(&s->full[i], 1);
    mbar_init(&s->empty[i], static_cast<uint32_t>(consumer_arrivals));
  }
}

// Producer (one thread): next tile of this CTA, or -1 when the kernel's tiles are exhausted. ctr == nullptr selects the
// static round-robin walk (every role then computes the same sequence by itself).
__device__ __forceinline__ int sched_produce(TileSched* s, SchedCursor& c, int* ctr, int num_tiles) {
  if (ctr == nullptr) {
    c.static_tile = c.static_tile < 0 ? static_cast<int>(blockIdx.x) : c.static_tile + static_cast<int>(gridDim.x);
    return c.static_tile < num_tiles ? c.static_tile : -1;
  }
  mbar_wait(&s->empty[c.slot], c.phase ^ 1u);
  int t = atomicAdd(ctr, 1);
  if (t >= num_tiles) {
    t = -1;
    if (atomicAdd(ctr + 1, 1) == static_cast<int>(gridDim.x) - 1) {   // every CTA is past its last draw: re-arm the slot
      ctr[0] = 0;
      ctr[1] = 0;
    }
  }
  s->ids[c.slot] = t;
  mbar_arrive(&s->full[c.slot]);   // release: the id is visible to whoever observes this phase
  c.advance();
  return t;
}

// Consumer: a whole warp (warp_collective) or a single thread. Returns the producer's next tile, -1 at the end.
__device__ __forceinline__ int sched_consume(TileSched* s, SchedCursor& c, int* ctr, int num_tiles, bool warp_collective,
                                             int lane) {
  if (ctr == nullptr) {
    c.static_tile = c.static_tile < 0 ? static_cast<int>(blockIdx.x) : c.static_tile + static_cast<int>(gridDim.x);
    return c.static_tile < num_tiles ? c.static_tile : -1;
  }
  mbar_wait(&s->full[c.slot], c.phase);
  const int t = s->ids[c.slot];
  if (warp_collective) {
    __syncwarp();
    if (lane == 0) mbar_arrive(&s->empty[c.slot]);
  } else {
    mbar_arrive(&s->empty[c.slot]);
  }
  c.advance();
  return t;
}

// Host: a zeroed {next, done} counter pair on the current device for one launch, or nullptr = static walk (the default;
// UNETB200_DYNAMIC_TILES=1 turns the scheduler on). Slots come round-robin from a static device array; a kernel re-arms its
// slot when it finishes, and far fewer kernels are ever in flight than there are slots.
int* sched_counter_slot();

}  // namespace ub
