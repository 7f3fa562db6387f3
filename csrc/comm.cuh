// Cross-GPU signalling shared by the fused collective kernels (comm.cu) and by the kernels that
// consume what they publish (the first kernel of a training step, elementwise.cu).
//
// Every rank owns one signal buffer of B200_SIGNAL_WORDS 32-bit words in symmetric memory; peers
// write into it with system-scope stores / reductions over NVLink, the owner polls it locally.
// Nothing is ever reset: all words are monotonically increasing counters compared with a signed
// difference, so a CUDA graph can replay the launches and a 32-bit wrap is harmless.
//
//   READY[block][peer]   epoch of the peer's block `block` (one-way store by the peer). A block
//                        that wrote epoch e to every peer and read >= e from every peer knows that
//                        all ranks have reached the same exchange ("barrier" = two one-way trips
//                        that overlap, no compare-and-swap round trip).
//   EPOCH[block]         this rank's own exchange counter per block (only this rank touches it).
//   PUB[peer]            number of block-level publishes of `peer` (red.add by the peer after a
//                        system-scope fence behind its parameter stores). PUB[own rank] counts this
//                        rank's own publishes, so "everything published so far has landed here" is
//                        PUB[p] >= PUB[rank] for every p: the end barrier of the update kernel is
//                        split-phase, its second half runs in the next step's first kernel.
//   PUB2[peer]           deferred variant (b200_comm_set_defer_publish): the update kernel ends without a
//                        fence; the next step's first kernel stores this rank's launch count into every
//                        peer's PUB2 word (the update's stores are complete at its kernel boundary) and
//                        waits for theirs.
//   METRICS[peer][8]     (epoch tag, value) pairs: per-step metric partial sums, one 8-byte store
//                        each, validated by the tag instead of by a fence.
#pragma once
#include "../include/b200dist.h"
#include "common.cuh"

namespace b200 {

constexpr int MAX_WORLD = 8;
constexpr int SIG_READY = 0;                                               // [MAX_BLOCKS][MAX_WORLD]
constexpr int SIG_EPOCH = SIG_READY + B200_COMM_MAX_BLOCKS * MAX_WORLD;    // [MAX_BLOCKS]
constexpr int SIG_PUB = SIG_EPOCH + B200_COMM_MAX_BLOCKS;                  // [MAX_WORLD]
constexpr int SIG_PUB2 = SIG_PUB + MAX_WORLD;                              // [MAX_WORLD]
constexpr int SIG_METRICS = SIG_PUB + 2 * MAX_WORLD;                       // [MAX_WORLD][8][2]
constexpr int SIG_END = SIG_METRICS + MAX_WORLD * 8 * 2;
static_assert(SIG_END <= B200_SIGNAL_WORDS, "signal buffer too small");
static_assert((SIG_METRICS & 1) == 0, "metric slots are 8-byte pairs");

// Host-side settings (comm.cu): how long a rank waits for a peer before it gives up, and an
// optional device buffer for per-phase globaltimer stamps.
uint64_t comm_timeout_ns();
bool comm_defer_publish();
struct SignalPeers { uint32_t* p[MAX_WORLD]; };

#ifdef __CUDACC__

__device__ __forceinline__ void st_relaxed_sys_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_add_relaxed_sys_u32(uint32_t* p, uint32_t v) {
  asm volatile("red.relaxed.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Polls a counter written by a peer until it has reached `want`. A peer that never arrives (a dead
// or hung rank) turns into a trap after `timeout_ns` instead of a GPU that spins forever: the
// launch fails, the next CUDA call on the host reports it and the job can abort.
__device__ __forceinline__ void spin_until_reached(const uint32_t* p, uint32_t want, uint64_t timeout_ns, int rank,
                                                   int peer, int what) {
  uint32_t spins = 0;
  uint64_t t0 = 0;
  while (static_cast<int32_t>(ld_acquire_sys_u32(p) - want) < 0) {
    if ((++spins & 0x3FFu) == 0) {
      const uint64_t now = globaltimer_ns();
      if (t0 == 0) {
        t0 = now;
      } else if (now - t0 > timeout_ns) {
        printf("b200: rank %d gave up waiting for rank %d (%s, block %d, want %u, have %u)\n", rank, peer,
               what == 0 ? "gradients ready" : what == 1 ? "parameters published" : "metrics", blockIdx.x, want,
               ld_relaxed_sys_u32(p));
        __trap();
      }
    }
  }
}

// Second half of the split-phase end barrier: returns once every publish that this rank has issued
// so far has also been issued by (and has landed from) every peer. Called by all threads of a block.
__device__ __forceinline__ void wait_published(const uint32_t* sig, int rank, int world, uint64_t timeout_ns) {
  if (world > 1 && static_cast<int>(threadIdx.x) < world) {
    const uint32_t want = ld_relaxed_sys_u32(sig + SIG_PUB + rank);
    spin_until_reached(sig + SIG_PUB + threadIdx.x, want, timeout_ns, rank, threadIdx.x, 1);
  }
  __syncthreads();
}
// Deferred variant: the publish itself happens here, one kernel boundary after the update's stores.
// The count of fused launches so far (block 0's exchange counter) is the value every rank expects.
__device__ __forceinline__ void publish_and_wait(const uint32_t* sig, const SignalPeers& peers, int rank, int world,
                                                 uint64_t timeout_ns) {
  if (world > 1 && static_cast<int>(threadIdx.x) < world) {
    const uint32_t want = ld_relaxed_sys_u32(sig + SIG_EPOCH);
    if (blockIdx.x == 0) st_relaxed_sys_u32(peers.p[threadIdx.x] + SIG_PUB2 + rank, want);
    spin_until_reached(sig + SIG_PUB2 + threadIdx.x, want, timeout_ns, rank, threadIdx.x, 1);
  }
  __syncthreads();
}

#endif  // __CUDACC__
}  // namespace b200
