// Fused collective + optimizer kernel over NVLink 5 / NVSwitch.
//
// Every rank launches this kernel on symmetric-memory buffers. Data moves with loads/stores
// issued from inside the kernel: `multimem.ld_reduce` lets the switch sum the copies of a
// gradient vector held by all GPUs (NVLS) and `multimem.st` stores the updated parameters to all
// GPUs at once; without a multicast mapping the same kernel uses plain loads/stores on the
// peer-mapped pointers. The optimizer runs between the reduce and the publish on the values in
// registers, so gradients cross NVLink exactly once and reduced gradients never touch HBM.
//
// Reference semantics: pmean + replicated AdamW (data_paral.py:210-217) and
// psum_scatter/axis_size + sharded AdamW (param_sharding.py:134-138,353).
//
// Cross-GPU synchronisation: compare-and-swap flags in a symmetric signal buffer, one word per
// (phase, block, sender). A sender flips the receiver's word 0->1 with release.sys, the receiver
// flips it back 1->0 with acquire.sys, so the protocol resets itself and a CUDA graph can replay
// the launch. All blocks of all ranks must be co-resident: the grid is capped at
// B200_COMM_MAX_BLOCKS CTAs.
#include "../include/b200dist.h"
#include "common.cuh"

namespace b200 {

constexpr int MAX_WORLD = 8;
// signal buffer layout (32-bit words)
constexpr int SIG_BARRIER = 0;                                            // [2][MAX_BLOCKS][MAX_WORLD]
constexpr int SIG_METRICS = 2 * B200_COMM_MAX_BLOCKS * MAX_WORLD;         // float [MAX_WORLD][8]
static_assert(SIG_METRICS + MAX_WORLD * 8 <= B200_SIGNAL_WORDS, "signal buffer too small");

struct CommDev {
  int rank, world;
  float* grad[MAX_WORLD];
  float* param[MAX_WORLD];
  __nv_bfloat16* bf16[MAX_WORLD];
  uint32_t* signal[MAX_WORLD];
  float* mc_grad;
  float* mc_param;
  __nv_bfloat16* mc_bf16;
};

__device__ __forceinline__ uint32_t cas_release_sys(uint32_t* addr, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.global.release.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(addr), "r"(cmp), "r"(val) : "memory");
  return old;
}
__device__ __forceinline__ uint32_t cas_acquire_sys(uint32_t* addr, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(addr), "r"(cmp), "r"(val) : "memory");
  return old;
}

// Barrier across ranks for this block index. `phase` selects one of two word sets so that the
// start and end barriers of one launch never alias.
__device__ __forceinline__ void cross_rank_barrier(const CommDev& c, int phase) {
  __syncthreads();
  if (c.world > 1 && threadIdx.x < c.world) {
    const int peer = threadIdx.x;
    const int slot = SIG_BARRIER + (phase * B200_COMM_MAX_BLOCKS + blockIdx.x) * MAX_WORLD;
    uint32_t* remote = c.signal[peer] + slot + c.rank;   // my word in the peer's buffer
    uint32_t* local = c.signal[c.rank] + slot + peer;    // the peer's word in my buffer
    uint32_t spins = 0;
    while (cas_release_sys(remote, 0u, 1u) != 0u) {
      if (++spins > (1u << 28)) { printf("b200: signal put timed out (rank %d -> %d)\n", c.rank, peer); __trap(); }
    }
    spins = 0;
    while (cas_acquire_sys(local, 1u, 0u) != 1u) {
      if (++spins > (1u << 28)) { printf("b200: signal wait timed out (rank %d <- %d)\n", c.rank, peer); __trap(); }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ float4 ld_sys_f32x4(const float* p) {
  float4 r;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ float4 mc_ld_reduce_f32x4(const float* mc) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(mc) : "memory");
  return r;
}
__device__ __forceinline__ void mc_st_f32x4(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void mc_st_b32x2(void* mc, uint32_t a, uint32_t b) {
  asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1,%2};" ::"l"(mc), "f"(__uint_as_float(a)), "f"(__uint_as_float(b)) : "memory");
}

__device__ __forceinline__ void adamw1(float& p, float& m, float& v, float g, const b200_adamw_hparams& hp, float c1, float c2) {
  m = hp.beta1 * m + (1.0f - hp.beta1) * g;
  v = hp.beta2 * v + (1.0f - hp.beta2) * g * g;
  p = p - hp.lr * ((m * c1) / (sqrtf(v * c2) + hp.eps) + hp.weight_decay * p);
}

__global__ void __launch_bounds__(256, 2)
fused_reduce_adamw_kernel(const CommDev c, const b200_seg* __restrict__ segs, int nseg, int broadcast,
                          float* __restrict__ p_local, __nv_bfloat16* __restrict__ bf16_local,
                          float* __restrict__ m, float* __restrict__ v, const b200_adamw_hparams hp,
                          const int32_t* __restrict__ step_count, const float* __restrict__ step_metrics,
                          float* __restrict__ metrics_acc, int nmetrics, const b200_step_end end) {
  pdl_sync();
  // metrics: every rank drops its partial sums into every peer's signal buffer before the barrier
  if (blockIdx.x == 0 && step_metrics != nullptr && c.world > 1) {
    for (int i = threadIdx.x; i < c.world * nmetrics; i += blockDim.x) {
      int peer = i / nmetrics, j = i - peer * nmetrics;
      float* dst = reinterpret_cast<float*>(c.signal[peer] + SIG_METRICS) + c.rank * 8 + j;
      asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(dst), "f"(step_metrics[j]) : "memory");
    }
  }
  cross_rank_barrier(c, 0);

  const float t = static_cast<float>(*step_count + 1);
  const float c1 = 1.0f / (1.0f - __powf(hp.beta1, t));
  const float c2 = 1.0f / (1.0f - __powf(hp.beta2, t));
  const float* my_grad = c.grad[c.rank];

  for (int s = 0; s < nseg; ++s) {
    const b200_seg sg = segs[s];
    const bool vec = ((sg.row_len | sg.src_pitch | sg.src_off | sg.dst_off | sg.loc_off) & 3) == 0;
    if (vec) {
      const int rl4 = sg.row_len >> 2;
      const int64_t total4 = static_cast<int64_t>(sg.rows) * rl4;   // multi-row segments: < 2^31 (host-checked)
      const int64_t nthreads = static_cast<int64_t>(gridDim.x) * blockDim.x;
      constexpr int U = 4;   // independent 16-byte vectors per thread per iteration (memory-level parallelism)
      for (int64_t i0 = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i0 < total4; i0 += nthreads * U) {
        // per vector only (row, col) are kept (32-bit); the three 64-bit offsets are rebuilt where used
        uint32_t vrow[U], vcol[U];
        bool ok[U];
        float4 g[U], pp[U], mm[U], vv[U];
        float* p_base = broadcast ? c.param[c.rank] : p_local;
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int64_t i = i0 + u * nthreads;
          ok[u] = i < total4;
          const uint32_t ii = ok[u] ? static_cast<uint32_t>(i) : 0u;
          // rows == 1 (every DP segment, contiguous FSDP shards) needs no division; otherwise the
          // segment is far below 2^31 vectors and a 32-bit division is enough
          vrow[u] = sg.rows == 1 ? 0u : ii / static_cast<uint32_t>(rl4);
          vcol[u] = (ii - vrow[u] * static_cast<uint32_t>(rl4)) << 2;
        }
        auto src_of = [&](int u) { return sg.src_off + static_cast<int64_t>(vrow[u]) * sg.src_pitch + vcol[u]; };
        auto loc_of = [&](int u) { return sg.loc_off + static_cast<int64_t>(vrow[u]) * sg.row_len + vcol[u]; };
        auto dst_of = [&](int u) {
          return (broadcast ? sg.dst_off : sg.loc_off) + static_cast<int64_t>(vrow[u]) * sg.row_len + vcol[u];
        };
        // ---- loads (all issued before any use) ----
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (!ok[u]) continue;
          if (c.world == 1) g[u] = *reinterpret_cast<const float4*>(my_grad + src_of(u));
          else if (c.mc_grad != nullptr) g[u] = mc_ld_reduce_f32x4(c.mc_grad + src_of(u));
          else g[u] = make_float4(0.f, 0.f, 0.f, 0.f);
          mm[u] = *reinterpret_cast<float4*>(m + loc_of(u));
          vv[u] = *reinterpret_cast<float4*>(v + loc_of(u));
          pp[u] = *reinterpret_cast<float4*>(p_base + dst_of(u));
        }
        if (c.world > 1 && c.mc_grad == nullptr) {
#pragma unroll 1
          for (int k = 0; k < c.world; ++k) {
            // DP slices are disjoint per rank: start at myself to spread the traffic. Replicated
            // FSDP leaves are reduced by every rank: a fixed order keeps the replicas bit-identical.
            int peer = broadcast ? c.rank + k : k;
            if (peer >= c.world) peer -= c.world;
            float4 t4[U];
#pragma unroll
            for (int u = 0; u < U; ++u)
              if (ok[u]) t4[u] = ld_sys_f32x4(c.grad[peer] + src_of(u));
#pragma unroll
            for (int u = 0; u < U; ++u)
              if (ok[u]) { g[u].x += t4[u].x; g[u].y += t4[u].y; g[u].z += t4[u].z; g[u].w += t4[u].w; }
          }
        }
        // ---- AdamW + publish ----
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (!ok[u]) continue;
          adamw1(pp[u].x, mm[u].x, vv[u].x, g[u].x * hp.grad_scale, hp, c1, c2);
          adamw1(pp[u].y, mm[u].y, vv[u].y, g[u].y * hp.grad_scale, hp, c1, c2);
          adamw1(pp[u].z, mm[u].z, vv[u].z, g[u].z * hp.grad_scale, hp, c1, c2);
          adamw1(pp[u].w, mm[u].w, vv[u].w, g[u].w * hp.grad_scale, hp, c1, c2);
          *reinterpret_cast<float4*>(m + loc_of(u)) = mm[u];
          *reinterpret_cast<float4*>(v + loc_of(u)) = vv[u];
          __nv_bfloat162 lo = __floats2bfloat162_rn(pp[u].x, pp[u].y), hi = __floats2bfloat162_rn(pp[u].z, pp[u].w);
          const uint32_t w0 = *reinterpret_cast<uint32_t*>(&lo), w1 = *reinterpret_cast<uint32_t*>(&hi);
          if (!broadcast || c.world == 1) {
            *reinterpret_cast<float4*>(p_base + dst_of(u)) = pp[u];
            __nv_bfloat16* b = broadcast ? c.bf16[c.rank] : bf16_local;
            if (b != nullptr) *reinterpret_cast<uint2*>(b + dst_of(u)) = make_uint2(w0, w1);
          } else if (c.mc_param != nullptr) {
            mc_st_f32x4(c.mc_param + dst_of(u), pp[u]);
            if (c.mc_bf16 != nullptr) mc_st_b32x2(c.mc_bf16 + dst_of(u), w0, w1);
          } else {
#pragma unroll 1
            for (int k = 0; k < c.world; ++k) {
              int peer = c.rank + k;
              if (peer >= c.world) peer -= c.world;
              *reinterpret_cast<float4*>(c.param[peer] + dst_of(u)) = pp[u];
              if (c.bf16[peer] != nullptr) *reinterpret_cast<uint2*>(c.bf16[peer] + dst_of(u)) = make_uint2(w0, w1);
            }
          }
        }
      }
    } else {
      // scalar path for odd shapes (never hit by the padded layouts the engines build)
      const int64_t total = static_cast<int64_t>(sg.rows) * sg.row_len;
      for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
           i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int64_t row = i / sg.row_len, col = i - row * sg.row_len;
        const int64_t src = sg.src_off + row * sg.src_pitch + col;
        const int64_t loc = sg.loc_off + i;
        float g = 0.f;
        for (int k = 0; k < c.world; ++k) g += *reinterpret_cast<const volatile float*>(c.grad[k] + src);
        const int64_t dst = broadcast ? sg.dst_off + i : loc;
        float* p_base = broadcast ? c.param[c.rank] : p_local;
        float pp = p_base[dst], mm = m[loc], vv = v[loc];
        adamw1(pp, mm, vv, g * hp.grad_scale, hp, c1, c2);
        m[loc] = mm; v[loc] = vv;
        if (broadcast) {
          for (int k = 0; k < c.world; ++k) {
            c.param[k][dst] = pp;
            if (c.bf16[k] != nullptr) c.bf16[k][dst] = __float2bfloat16_rn(pp);
          }
        } else {
          p_base[dst] = pp;
          if (bf16_local != nullptr) bf16_local[dst] = __float2bfloat16_rn(pp);
        }
      }
    }
  }

  if (blockIdx.x == 0 && step_metrics != nullptr && metrics_acc != nullptr && threadIdx.x < nmetrics) {
    float acc = 0.f;
    if (c.world == 1) {
      acc = step_metrics[threadIdx.x];
    } else {
      const volatile float* box = reinterpret_cast<const volatile float*>(c.signal[c.rank] + SIG_METRICS);
      for (int k = 0; k < c.world; ++k) acc += box[k * 8 + threadIdx.x];
    }
    metrics_acc[threadIdx.x] += acc;
  }
  __threadfence_system();
  cross_rank_barrier(c, 1);
  // End-of-step bookkeeping folded into the last update of the step: the block that finishes last
  // (self-resetting counter) advances the RNG key and the step counter. Every block read the step
  // counter long before it got here, and nothing in this kernel reads the key.
  if (end.done_word != nullptr && threadIdx.x == 0) {
    __threadfence();
    if (atomicInc(end.done_word, gridDim.x - 1) == gridDim.x - 1) {
      const uint64_t k = *end.rng_key;
      *end.rng_key = mix64(k ^ end.salt_next);
      *end.step_count += 1;
    }
  }
}

// All-gather by peer loads: for every segment and every rank p, `rows` rows of `row_bytes` bytes are
// copied from peer p (contiguous rows at src_off, pitch src_pitch) to
// dst + dst_off + p*peer_stride (+ r*dst_pitch per row). Row shards use rows = 1; column shards
// interleave the peers' column blocks into the full rows. No collective launch, no staging copy.
// The shards must already be published (the optimizer kernel's end barrier).
struct PeerPtrs { const uint8_t* p[MAX_WORLD]; };
__global__ void __launch_bounds__(256)
peer_gather_kernel(const PeerPtrs peers, int world, const b200_gather_seg* __restrict__ segs, int nseg,
                   uint8_t* __restrict__ dst) {
  pdl_sync();
  const int64_t nthreads = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  for (int s = 0; s < nseg; ++s) {
    const b200_gather_seg sg = segs[s];
    const bool vec = ((sg.src_off | sg.dst_off | sg.row_bytes | sg.src_pitch | sg.dst_pitch | sg.peer_stride) & 15) == 0;
    for (int p = blockIdx.y; p < world; p += gridDim.y) {
      const uint8_t* src = peers.p[p] + sg.src_off;
      uint8_t* d = dst + sg.dst_off + static_cast<int64_t>(p) * sg.peer_stride;
      if (vec) {
        const int64_t rb16 = sg.row_bytes >> 4;
        const int64_t n16 = rb16 * sg.rows;
        constexpr int U = 4;
        for (int64_t i0 = tid; i0 < n16; i0 += nthreads * U) {
          uint4 v[U];
          int64_t doff[U];
          bool ok[U];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int64_t i = i0 + u * nthreads;
            ok[u] = i < n16;
            const int64_t ii = ok[u] ? i : 0;
            const int64_t r = sg.rows == 1 ? 0 : static_cast<int64_t>(static_cast<uint32_t>(ii) / static_cast<uint32_t>(rb16));
            const int64_t c = ii - r * rb16;
            doff[u] = r * sg.dst_pitch + (c << 4);
            if (ok[u]) v[u] = *reinterpret_cast<const uint4*>(src + r * sg.src_pitch + (c << 4));
          }
#pragma unroll
          for (int u = 0; u < U; ++u)
            if (ok[u]) *reinterpret_cast<uint4*>(d + doff[u]) = v[u];
        }
      } else {
        const int64_t n = sg.row_bytes * sg.rows;
        for (int64_t i = tid; i < n; i += nthreads) {
          const int64_t r = i / sg.row_bytes, c = i - r * sg.row_bytes;
          d[r * sg.dst_pitch + c] = src[r * sg.src_pitch + c];
        }
      }
    }
  }
}

__global__ void signal_barrier_kernel(const CommDev c) {
  pdl_sync();
  cross_rank_barrier(c, 0);
  cross_rank_barrier(c, 1);
}

__global__ void bump_step_kernel2(int32_t* step) {
  pdl_sync(); *step += 1; }

static int fill_ctx(const b200_comm_ctx* ctx, CommDev* d) {
  B200_REQUIRE(ctx->world >= 1 && ctx->world <= MAX_WORLD, "comm: world size %d not in [1,%d]", ctx->world, MAX_WORLD);
  B200_REQUIRE(ctx->rank >= 0 && ctx->rank < ctx->world, "comm: bad rank %d", ctx->rank);
  memset(d, 0, sizeof(*d));
  d->rank = ctx->rank;
  d->world = ctx->world;
  for (int i = 0; i < ctx->world; ++i) {
    d->grad[i] = ctx->peer_grad ? static_cast<float*>(ctx->peer_grad[i]) : nullptr;
    d->param[i] = ctx->peer_param ? static_cast<float*>(ctx->peer_param[i]) : nullptr;
    d->bf16[i] = ctx->peer_bf16 ? static_cast<__nv_bfloat16*>(ctx->peer_bf16[i]) : nullptr;
    d->signal[i] = ctx->peer_signal ? static_cast<uint32_t*>(ctx->peer_signal[i]) : nullptr;
  }
  d->mc_grad = static_cast<float*>(ctx->mc_grad);
  d->mc_param = static_cast<float*>(ctx->mc_param);
  d->mc_bf16 = static_cast<__nv_bfloat16*>(ctx->mc_bf16);
  if (ctx->world > 1)
    for (int i = 0; i < ctx->world; ++i)
      B200_REQUIRE(d->signal[i] != nullptr, "comm: missing signal buffer of rank %d", i);
  return 0;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_fused_reduce_adamw(const b200_comm_ctx* ctx, const b200_seg* segs_dev, int nseg,
                                       int broadcast, float* p_local, void* bf16_local, float* m, float* v,
                                       const b200_adamw_hparams* hp, int32_t* step_count, int bump_step,
                                       const float* step_metrics, float* metrics_acc, int nmetrics,
                                       int blocks, void* stream) {
  return b200_fused_reduce_adamw_end(ctx, segs_dev, nseg, broadcast, p_local, bf16_local, m, v, hp, step_count, bump_step,
                                     step_metrics, metrics_acc, nmetrics, blocks, nullptr, stream);
}

extern "C" int b200_fused_reduce_adamw_end(const b200_comm_ctx* ctx, const b200_seg* segs_dev, int nseg,
                                           int broadcast, float* p_local, void* bf16_local, float* m, float* v,
                                           const b200_adamw_hparams* hp, int32_t* step_count, int bump_step,
                                           const float* step_metrics, float* metrics_acc, int nmetrics,
                                           int blocks, const b200_step_end* end, void* stream) {
  b200_step_end e;
  memset(&e, 0, sizeof(e));
  if (end != nullptr) {
    e = *end;
    B200_REQUIRE(e.rng_key != nullptr && e.step_count != nullptr && e.done_word != nullptr,
                 "comm: step_end needs rng_key, step_count and a zero-initialised done_word");
    B200_REQUIRE(!bump_step, "comm: bump_step and step_end both advance the step counter");
  }
  CommDev d;
  int rc = fill_ctx(ctx, &d);
  if (rc) return rc;
  B200_REQUIRE(nseg >= 0 && segs_dev != nullptr, "comm: no segments");
  B200_REQUIRE(nmetrics >= 0 && nmetrics <= 8, "comm: at most 8 metrics");
  for (int i = 0; i < ctx->world; ++i) {
    B200_REQUIRE(d.grad[i] != nullptr, "comm: missing gradient buffer of rank %d", i);
    if (broadcast) B200_REQUIRE(d.param[i] != nullptr, "comm: missing parameter buffer of rank %d", i);
  }
  if (!broadcast) B200_REQUIRE(p_local != nullptr, "comm: local mode needs p_local");
  if (blocks <= 0) blocks = 64;
  if (ctx->world == 1) {
    blocks = 2 * num_sms();     // no cross-GPU barrier: fill the machine (two resident blocks per SM)
  } else {
    if (blocks > B200_COMM_MAX_BLOCKS) blocks = B200_COMM_MAX_BLOCKS;
    if (blocks > num_sms()) blocks = num_sms();   // every block must be resident: they spin on each other across GPUs
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  B200_CUDA(launch_k(fused_reduce_adamw_kernel, dim3(blocks), dim3(256), 0, s, d, segs_dev, nseg, broadcast, p_local,
                                                   static_cast<__nv_bfloat16*>(bf16_local), m, v, *hp,
                                                   step_count, step_metrics, metrics_acc, nmetrics, e));
  if (bump_step) {
    B200_CUDA(launch_k(bump_step_kernel2, dim3(1), dim3(1), 0, s, step_count));
  }
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_signal_barrier(const b200_comm_ctx* ctx, void* stream) {
  CommDev d;
  int rc = fill_ctx(ctx, &d);
  if (rc) return rc;
  B200_CUDA(launch_k(signal_barrier_kernel, dim3(1), dim3(32), 0, static_cast<cudaStream_t>(stream), d));
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_peer_gather(void* const* peer_base, int world, const b200_gather_seg* segs_dev, int nseg,
                                void* dst, int blocks, void* stream) {
  B200_REQUIRE(world >= 1 && world <= MAX_WORLD, "peer_gather: world size %d not in [1,%d]", world, MAX_WORLD);
  PeerPtrs pp;
  memset(&pp, 0, sizeof(pp));
  for (int i = 0; i < world; ++i) {
    B200_REQUIRE(peer_base[i] != nullptr, "peer_gather: null peer pointer %d", i);
    pp.p[i] = static_cast<const uint8_t*>(peer_base[i]);
  }
  if (blocks <= 0) blocks = 4;
  if (blocks * world > 4 * num_sms()) blocks = 4 * num_sms() / world;
  B200_CUDA(launch_k(peer_gather_kernel, dim3(blocks, world), dim3(256), 0, static_cast<cudaStream_t>(stream), 
      pp, world, segs_dev, nseg, static_cast<uint8_t*>(dst)));
  B200_CUDA(cudaGetLastError());
  return 0;
}
