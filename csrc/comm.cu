// Fused collective + optimizer kernel over NVLink 5 / NVSwitch, and the optimizer-free collectives
// (all-reduce, reduce-scatter, all-gather) built from the same pieces.
//
// Every rank launches the kernel on symmetric-memory buffers. Data moves with loads/stores issued
// from inside the kernel: `multimem.ld_reduce` lets the switch sum the copies of a gradient vector
// held by all GPUs (NVLS) and `multimem.st` stores the updated parameters to all GPUs at once;
// without a multicast mapping the same kernel uses plain loads/stores on the peer-mapped pointers.
// The optimizer runs between the reduce and the publish on the values in registers, so gradients
// cross NVLink exactly once and reduced gradients never touch HBM.
//
// Reference semantics: pmean + replicated AdamW (data_paral.py:210-217) and
// psum_scatter/axis_size + sharded AdamW (param_sharding.py:134-138,353).
//
// Cross-GPU synchronisation (comm.cuh): monotonically increasing epoch words, one-way
// `st.relaxed.sys` by the sender and an `ld.acquire.sys` poll by the receiver. One launch is
//
//   exchange "gradients ready"   every block tells block b of every peer its epoch and waits for theirs.
//                                The gradients were written by earlier kernels of the stream, so they
//                                are already performed at the device's L2 (the point every peer and the
//                                switch read from) and the flag needs no fence in front of it.
//   reduce -> AdamW -> publish
//   fence.sys, then "published"  a fire-and-forget red.add into every peer's PUB counter. Nobody waits
//                                here: the matching wait sits in the first kernel of the next step
//                                (b200_cast_zero_wait / b200_comm_wait_published), off this step's
//                                critical path.
//
// All blocks of all ranks must be co-resident: the grid is capped at B200_COMM_MAX_BLOCKS CTAs.
#include "comm.cuh"

#include <stdlib.h>

namespace b200 {

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

// Everything one rank's launch needs. The real kernel takes it by value; the single-GPU emulation
// (several virtual ranks in one cooperative launch) reads one per blockIdx.y from device memory.
struct RankArgs {
  CommDev c;
  const b200_seg* segs;
  int nseg, broadcast, publish_fp32, nmetrics;
  float* p_local;
  __nv_bfloat16* bf16_local;
  float* m;
  float* v;
  b200_adamw_hparams hp;
  const int32_t* step_count;
  const float* step_metrics;
  float* metrics_acc;
  b200_step_end end;
  uint64_t timeout_ns;
  unsigned long long* stamps;   // optional [blocks][8] globaltimer stamps (profiling aid)
  int defer_publish;
};

static uint64_t g_timeout_ns = 0;
static int g_defer_publish = -1;
static unsigned long long* g_stamps = nullptr;

uint64_t comm_timeout_ns() {
  if (g_timeout_ns == 0) {
    const char* e = getenv("B200DIST_COMM_TIMEOUT_MS");
    long long ms = e ? atoll(e) : 0;
    if (ms <= 0) ms = 30000;
    g_timeout_ns = static_cast<uint64_t>(ms) * 1000000ull;
  }
  return g_timeout_ns;
}

bool comm_defer_publish() {
  if (g_defer_publish < 0) {
    const char* e = getenv("B200DIST_DEFER_PUBLISH");
    // 2 = fence by one thread per block behind a block barrier (default), 0 = fence by every thread
    // (the first version: ~5 us slower on a 70 us step), 1 = deferred to the next step's first kernel
    g_defer_publish = (e && e[0] == '1') ? 1 : (e && e[0] == '0') ? 0 : 2;
  }
  return g_defer_publish == 1;
}

// One all-to-all flag exchange for this block index: returns the epoch it used. Doubles as the
// block-wide barrier in front of and behind it.
__device__ __forceinline__ uint32_t exchange(const CommDev& c, int bid, uint64_t timeout_ns) {
  __shared__ uint32_t s_epoch;
  if (c.world == 1) {   // nobody to talk to (and no signal buffer needed)
    __syncthreads();
    return 0u;
  }
  uint32_t* mine = c.signal[c.rank];
  if (threadIdx.x == 0) {
    const uint32_t e = mine[SIG_EPOCH + bid] + 1u;
    mine[SIG_EPOCH + bid] = e;
    s_epoch = e;
  }
  __syncthreads();
  const uint32_t e = s_epoch;
  if (static_cast<int>(threadIdx.x) < c.world) {
    const int peer = threadIdx.x;
    st_relaxed_sys_u32(c.signal[peer] + SIG_READY + bid * MAX_WORLD + c.rank, e);
    spin_until_reached(mine + SIG_READY + bid * MAX_WORLD + peer, e, timeout_ns, c.rank, peer, 0);
  }
  __syncthreads();
  return e;
}

// Orders everything this block stored (locally, to peers, through the multicast address) before the
// "published" increments that follow it.
__device__ __forceinline__ void publish(const CommDev& c, int one_thread_fence) {
  if (one_thread_fence) {
    // bar.sync orders the block's stores before thread 0's fence; the fence is cumulative
    __syncthreads();
    if (threadIdx.x == 0) __threadfence_system();
  } else {
    __threadfence_system();
  }
  __syncthreads();
  if (c.world > 1 && static_cast<int>(threadIdx.x) < c.world)
    red_add_relaxed_sys_u32(c.signal[threadIdx.x] + SIG_PUB + c.rank, 1u);
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
  p = p - hp.lr * (__fdividef(m * c1, sqrtf(v * c2) + hp.eps) + hp.weight_decay * p);   // (2 ulp division)
}

__device__ __forceinline__ void stamp(const RankArgs& a, int bid, int k) {
  if (a.stamps != nullptr && threadIdx.x == 0) a.stamps[bid * 8 + k] = globaltimer_ns();
}

// Reduce + AdamW + publish of one segment by the blocks [0, vnblocks) of a (sub-)grid.
__device__ __forceinline__ void update_segment(const RankArgs& a, const b200_seg& sg, const int vbid, const int vnblocks,
                                               const float c1, const float c2) {
  const CommDev& c = a.c;
  const b200_adamw_hparams& hp = a.hp;
  const int broadcast = a.broadcast;
  float* const m = a.m;
  float* const v = a.v;
  const float* my_grad = c.grad[c.rank];
  // DP: the fp32 master leaves the owner only when asked to (publish_fp32); the bf16 working copy
  // always goes to every rank.
  const bool fp32_everywhere = broadcast && c.world > 1 && a.publish_fp32;
  const bool bf16_everywhere = broadcast && c.world > 1;
    const bool vec = ((sg.row_len | sg.src_pitch | sg.src_off | sg.dst_off | sg.loc_off) & 3) == 0;
  // Segments that several ranks reduce redundantly (replicated leaves of an FSDP run) are summed
  // with peer loads in a fixed rank order, so that every replica computes bit-identical values;
  // the in-switch reduction is not documented to return the same sum to different requesters.
  const bool use_mc = c.mc_grad != nullptr && !(sg.flags & B200_SEG_ORDERED);
  if (vec) {
    const int rl4 = sg.row_len >> 2;
    const int64_t total4 = static_cast<int64_t>(sg.rows) * rl4;   // multi-row segments: < 2^31 (checked by the engine)
    const int64_t nthreads = static_cast<int64_t>(vnblocks) * blockDim.x;
    constexpr int U = 4;   // independent 16-byte vectors per thread per iteration (memory-level parallelism)
    for (int64_t i0 = static_cast<int64_t>(vbid) * blockDim.x + threadIdx.x; i0 < total4; i0 += nthreads * U) {
      // Per vector two 32-bit element offsets relative to the segment: `rel` into the contiguous local /
      // destination rows (always 4 * index) and `srel` into the possibly strided source rows. The engines
      // check that a segment spans fewer than 2^31 elements, pitch included.
      uint32_t rel[U], srel[U];
      bool ok[U];
      float4 g[U], pp[U], mm[U], vv[U];
      float* p_base = broadcast ? c.param[c.rank] : a.p_local;
      const bool contiguous = sg.rows == 1 || sg.src_pitch == sg.row_len;
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t i = i0 + u * nthreads;
        ok[u] = i < total4;
        const uint32_t ii = ok[u] ? static_cast<uint32_t>(i) : 0u;
        rel[u] = ii << 2;
        srel[u] = rel[u];
        if (!contiguous) {       // column shards: rows of row_len elements, src_pitch apart
          const uint32_t row = ii / static_cast<uint32_t>(rl4);
          srel[u] = row * static_cast<uint32_t>(sg.src_pitch) + ((ii - row * static_cast<uint32_t>(rl4)) << 2);
        }
      }
      auto src_of = [&](int u) { return sg.src_off + srel[u]; };
      auto loc_of = [&](int u) { return sg.loc_off + rel[u]; };
      auto dst_of = [&](int u) { return (broadcast ? sg.dst_off : sg.loc_off) + rel[u]; };
      // ---- loads (all issued before any use; the peer-load path fetches m / v / p after its sums to
      //      keep the registers for the peers' vectors) ----
      const bool p2p = c.world > 1 && !use_mc;
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!ok[u]) continue;
        if (c.world == 1) g[u] = *reinterpret_cast<const float4*>(my_grad + src_of(u));
        else if (use_mc) g[u] = mc_ld_reduce_f32x4(c.mc_grad + src_of(u));
        else g[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (!p2p) {
          mm[u] = *reinterpret_cast<float4*>(m + loc_of(u));
          vv[u] = *reinterpret_cast<float4*>(v + loc_of(u));
          pp[u] = *reinterpret_cast<float4*>(p_base + dst_of(u));
        }
      }
      if (p2p) {
        // Peer loads, PF peers in flight at a time (each load is an NVLink round trip), summed in a
        // fixed order. DP slices are disjoint per rank: start at myself to spread the traffic.
        // Segments reduced by every rank use the same order everywhere (see above).
#pragma unroll 1
        constexpr int PF = 2;
        for (int k0 = 0; k0 < c.world; k0 += PF) {
          float4 t4[PF][U];
#pragma unroll
          for (int kk = 0; kk < PF; ++kk) {
            if (k0 + kk >= c.world) continue;
            int peer = broadcast ? c.rank + k0 + kk : k0 + kk;
            if (peer >= c.world) peer -= c.world;
#pragma unroll
            for (int u = 0; u < U; ++u)
              if (ok[u]) t4[kk][u] = ld_sys_f32x4(c.grad[peer] + src_of(u));
          }
#pragma unroll
          for (int kk = 0; kk < PF; ++kk) {
            if (k0 + kk >= c.world) continue;
#pragma unroll
            for (int u = 0; u < U; ++u)
              if (ok[u]) { g[u].x += t4[kk][u].x; g[u].y += t4[kk][u].y; g[u].z += t4[kk][u].z; g[u].w += t4[kk][u].w; }
          }
        }
      }
      if (p2p) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (!ok[u]) continue;
          mm[u] = *reinterpret_cast<float4*>(m + loc_of(u));
          vv[u] = *reinterpret_cast<float4*>(v + loc_of(u));
          pp[u] = *reinterpret_cast<float4*>(p_base + dst_of(u));
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
        // fp32 master
        if (!fp32_everywhere) {
          *reinterpret_cast<float4*>(p_base + dst_of(u)) = pp[u];
        } else if (c.mc_param != nullptr) {
          mc_st_f32x4(c.mc_param + dst_of(u), pp[u]);
        } else {
#pragma unroll 1
          for (int k = 0; k < c.world; ++k) {
            int peer = c.rank + k;
            if (peer >= c.world) peer -= c.world;
            *reinterpret_cast<float4*>(c.param[peer] + dst_of(u)) = pp[u];
          }
        }
        // bf16 working copy
        if (!bf16_everywhere) {
          __nv_bfloat16* b = broadcast ? c.bf16[c.rank] : a.bf16_local;
          if (b != nullptr) *reinterpret_cast<uint2*>(b + dst_of(u)) = make_uint2(w0, w1);
        } else if (c.mc_bf16 != nullptr) {
          mc_st_b32x2(c.mc_bf16 + dst_of(u), w0, w1);
        } else {
#pragma unroll 1
          for (int k = 0; k < c.world; ++k) {
            int peer = c.rank + k;
            if (peer >= c.world) peer -= c.world;
            if (c.bf16[peer] != nullptr) *reinterpret_cast<uint2*>(c.bf16[peer] + dst_of(u)) = make_uint2(w0, w1);
          }
        }
      }
    }
  } else {
    // scalar path for odd shapes (never hit by the padded layouts the engines build)
    const int64_t total = static_cast<int64_t>(sg.rows) * sg.row_len;
    for (int64_t i = static_cast<int64_t>(vbid) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<int64_t>(vnblocks) * blockDim.x) {
      const int64_t row = i / sg.row_len, col = i - row * sg.row_len;
      const int64_t src = sg.src_off + row * sg.src_pitch + col;
      const int64_t loc = sg.loc_off + i;
      float g = 0.f;
      for (int k = 0; k < c.world; ++k) g += *reinterpret_cast<const volatile float*>(c.grad[k] + src);
      const int64_t dst = broadcast ? sg.dst_off + i : loc;
      float* p_base = broadcast ? c.param[c.rank] : a.p_local;
      float pp = p_base[dst], mm = m[loc], vv = v[loc];
      adamw1(pp, mm, vv, g * hp.grad_scale, hp, c1, c2);
      m[loc] = mm; v[loc] = vv;
      if (broadcast) {
        for (int k = 0; k < c.world; ++k) {
          if (k == c.rank || a.publish_fp32) c.param[k][dst] = pp;
          if (c.bf16[k] != nullptr) c.bf16[k][dst] = __float2bfloat16_rn(pp);
        }
      } else {
        p_base[dst] = pp;
        if (a.bf16_local != nullptr) a.bf16_local[dst] = __float2bfloat16_rn(pp);
      }
    }
  }
}

// Blocks one pass over a segment needs (every thread: 4 vectors of 4 floats, or one scalar).
__device__ __forceinline__ int segment_blocks(const b200_seg& sg) {
  const bool vec = ((sg.row_len | sg.src_pitch | sg.src_off | sg.dst_off | sg.loc_off) & 3) == 0;
  const int64_t items = static_cast<int64_t>(sg.rows) * sg.row_len;
  const int64_t per_block = vec ? 256 * 4 * 4 : 256;
  const int64_t need = (items + per_block - 1) / per_block;
  return need < 1 ? 1 : need > (1 << 20) ? (1 << 20) : static_cast<int>(need);
}

__device__ __forceinline__ void fused_reduce_adamw_body(const RankArgs& a, const int bid, const int nblocks) {
  const CommDev& c = a.c;
  const b200_adamw_hparams& hp = a.hp;
  stamp(a, bid, 0);
  const uint32_t epoch = exchange(c, bid, a.timeout_ns);
  stamp(a, bid, 1);
  // metrics: every rank drops its partial sums, tagged with the epoch, into every peer's mailbox
  if (bid == 0 && a.step_metrics != nullptr && c.world > 1) {
    for (int i = threadIdx.x; i < c.world * a.nmetrics; i += blockDim.x) {
      const int peer = i / a.nmetrics, j = i - peer * a.nmetrics;
      uint32_t* dst = c.signal[peer] + SIG_METRICS + (c.rank * 8 + j) * 2;
      asm volatile("st.relaxed.sys.global.v2.b32 [%0], {%1, %2};" ::"l"(dst), "r"(epoch),
                   "r"(__float_as_uint(a.step_metrics[j])) : "memory");
    }
  }

  const float t = static_cast<float>(*a.step_count + 1);
  const float c1 = 1.0f / (1.0f - __powf(hp.beta1, t));
  const float c2 = 1.0f / (1.0f - __powf(hp.beta2, t));

  // Segments larger than a quarter of the grid take the whole grid, one after the other. Runs of
  // smaller segments (an MLP's leaves, the bias / norm vectors of a transformer) are spread over
  // disjoint block ranges so that a whole run is in flight at once: one NVLink round trip per run
  // instead of one per segment.
  const int small_limit = nblocks / 4 > 0 ? nblocks / 4 : 1;
  for (int s = 0; s < a.nseg;) {
    // one pass: either segment s alone on the whole grid, or the run [s, e) of small segments side by side
    int my_seg = s, my_first = 0, my_need = nblocks, e = s + 1;
    const int need = segment_blocks(a.segs[s]);
    if (need <= small_limit && s + 1 < a.nseg) {
      int first = need;
      my_seg = bid < need ? s : -1;
      my_need = need;
      while (e < a.nseg) {
        const int n2 = segment_blocks(a.segs[e]);
        if (n2 > small_limit || first + n2 > nblocks) break;
        if (bid >= first && bid < first + n2) { my_seg = e; my_first = first; my_need = n2; }
        first += n2;
        ++e;
      }
      if (e == s + 1) { my_seg = s; my_first = 0; my_need = nblocks; }   // a run of one: give it the whole grid
    }
    if (my_seg >= 0) {
      const b200_seg sg = a.segs[my_seg];
      update_segment(a, sg, bid - my_first, my_need, c1, c2);
    }
    s = e;
  }
  stamp(a, bid, 2);

  if (bid == 0 && a.step_metrics != nullptr && a.metrics_acc != nullptr && static_cast<int>(threadIdx.x) < a.nmetrics) {
    float acc = 0.f;
    if (c.world == 1) {
      acc = a.step_metrics[threadIdx.x];
    } else {
      const uint32_t* box = c.signal[c.rank] + SIG_METRICS;
      for (int k = 0; k < c.world; ++k) {
        const uint32_t* slot = box + (k * 8 + threadIdx.x) * 2;
        spin_until_reached(slot, epoch, a.timeout_ns, c.rank, k, 2);          // the tag: this launch's value is there
        acc += __uint_as_float(ld_relaxed_sys_u32(slot + 1));
      }
    }
    a.metrics_acc[threadIdx.x] += acc;
  }
  if (a.defer_publish != 1) publish(c, a.defer_publish == 2);
  stamp(a, bid, 3);
  // End-of-step bookkeeping folded into the last update of the step: the block that finishes last
  // (self-resetting counter) advances the RNG key and the step counter. Every block read the step
  // counter long before it got here, and nothing in this kernel reads the key.
  if (a.end.done_word != nullptr && threadIdx.x == 0) {
    __threadfence();
    if (atomicInc(a.end.done_word, nblocks - 1) == static_cast<uint32_t>(nblocks - 1)) {
      const uint64_t k = *a.end.rng_key;
      *a.end.rng_key = mix64(k ^ a.end.salt_next);
      *a.end.step_count += 1;
    }
  }
  stamp(a, bid, 4);
}

__global__ void __launch_bounds__(256, 2) fused_reduce_adamw_kernel(const __grid_constant__ RankArgs a) {
  pdl_sync();
  fused_reduce_adamw_body(a, blockIdx.x, gridDim.x);
}

// Several ranks emulated on one GPU: blockIdx.y selects the rank, every "peer" pointer is local.
// Launched cooperatively so that all blocks of all virtual ranks are resident while they wait for
// each other. Used by the single-GPU tests of the cross-rank protocol.
__global__ void __launch_bounds__(256, 2) fused_reduce_adamw_emul_kernel(const RankArgs* __restrict__ ranks) {
  __shared__ RankArgs a;
  if (threadIdx.x == 0) a = ranks[blockIdx.y];
  __syncthreads();
  fused_reduce_adamw_body(a, blockIdx.x, gridDim.x);
}

// ---------------------------------------------------------------------------------------------
// Collectives without an optimizer, for code that calls psum / pmean / psum_scatter / all_gather by
// axis name on tensors that live in symmetric memory (comm/collectives.py, backend "fused").
//   mode 0  all-reduce in place:     buf[i] = scale * sum_p buf_p[i]        on every rank
//   mode 1  reduce-scatter:          out[i] = scale * sum_p buf_p[rank*n/W + i],  i < n/W   (out local)
//   mode 2  all-gather:              out[p*n + i] = buf_p[i],  i < n                        (out local)
// fp32 elements for modes 0/1 (n multiple of 4*W), 16-byte units for mode 2. Each launch is bracketed
// by two flag exchanges, so the call is complete (and the buffers reusable) when the kernel ends.
// ---------------------------------------------------------------------------------------------
struct CollArgs {
  int rank, world, mode;
  float* buf[MAX_WORLD];
  uint32_t* signal[MAX_WORLD];
  float* mc_buf;
  float* out;
  int64_t n;
  float scale;
  uint64_t timeout_ns;
};

__device__ __forceinline__ void collective_body(const CollArgs& a, const int bid, const int nblocks) {
  CommDev c;
  c.rank = a.rank;
  c.world = a.world;
  for (int i = 0; i < MAX_WORLD; ++i) c.signal[i] = a.signal[i];
  exchange(c, bid, a.timeout_ns);
  const int64_t nthreads = static_cast<int64_t>(nblocks) * blockDim.x;
  const int64_t tid = static_cast<int64_t>(bid) * blockDim.x + threadIdx.x;
  constexpr int U = 4;
  if (a.mode == 0 || a.mode == 1) {
    const int64_t per4 = (a.n / a.world) >> 2;           // this rank's slice, in float4
    const int64_t base = static_cast<int64_t>(a.rank) * (a.n / a.world);
    for (int64_t i0 = tid; i0 < per4; i0 += nthreads * U) {
      float4 g[U];
      bool ok[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t i = i0 + u * nthreads;
        ok[u] = i < per4;
        if (!ok[u]) continue;
        if (a.mc_buf != nullptr) g[u] = mc_ld_reduce_f32x4(a.mc_buf + base + (i << 2));
        else g[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      if (a.mc_buf == nullptr) {
#pragma unroll 1
        for (int k = 0; k < a.world; ++k) {   // fixed order: every rank would compute the same bits
          float4 t4[U];
#pragma unroll
          for (int u = 0; u < U; ++u)
            if (ok[u]) t4[u] = ld_sys_f32x4(a.buf[k] + base + ((i0 + u * nthreads) << 2));
#pragma unroll
          for (int u = 0; u < U; ++u)
            if (ok[u]) { g[u].x += t4[u].x; g[u].y += t4[u].y; g[u].z += t4[u].z; g[u].w += t4[u].w; }
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!ok[u]) continue;
        const int64_t i = i0 + u * nthreads;
        g[u].x *= a.scale; g[u].y *= a.scale; g[u].z *= a.scale; g[u].w *= a.scale;
        if (a.mode == 1) {
          *reinterpret_cast<float4*>(a.out + (i << 2)) = g[u];
        } else if (a.mc_buf != nullptr) {
          mc_st_f32x4(a.mc_buf + base + (i << 2), g[u]);
        } else {
#pragma unroll 1
          for (int k = 0; k < a.world; ++k) {
            int peer = a.rank + k;
            if (peer >= a.world) peer -= a.world;
            *reinterpret_cast<float4*>(a.buf[peer] + base + (i << 2)) = g[u];
          }
        }
      }
    }
  } else {
    const int64_t n16 = a.n;   // 16-byte units per rank
    for (int k = 0; k < a.world; ++k) {
      int peer = a.rank + k;
      if (peer >= a.world) peer -= a.world;
      const uint4* src = reinterpret_cast<const uint4*>(a.buf[peer]);
      uint4* dst = reinterpret_cast<uint4*>(a.out) + static_cast<int64_t>(peer) * n16;
      for (int64_t i0 = tid; i0 < n16; i0 += nthreads * U) {
        uint4 t4[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int64_t i = i0 + u * nthreads;
          if (i < n16) t4[u] = src[i];
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int64_t i = i0 + u * nthreads;
          if (i < n16) dst[i] = t4[u];
        }
      }
    }
  }
  __threadfence_system();
  exchange(c, bid, a.timeout_ns);
}

__global__ void __launch_bounds__(256, 2) collective_kernel(const __grid_constant__ CollArgs a) {
  pdl_sync();
  collective_body(a, blockIdx.x, gridDim.x);
}
__global__ void __launch_bounds__(256, 2) collective_emul_kernel(const CollArgs* __restrict__ ranks) {
  __shared__ CollArgs a;
  if (threadIdx.x == 0) a = ranks[blockIdx.y];
  __syncthreads();
  collective_body(a, blockIdx.x, gridDim.x);
}

// All-gather by peer loads: for every segment and every rank p, `rows` rows of `row_bytes` bytes are
// copied from peer p (contiguous rows at src_off, pitch src_pitch) to
// dst + dst_off + p*peer_stride (+ r*dst_pitch per row). Row shards use rows = 1; column shards
// interleave the peers' column blocks into the full rows. No collective launch, no staging copy.
// The shards must already be published (wait_published in an earlier kernel of the stream).
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

__global__ void signal_barrier_kernel(const CommDev c, uint64_t timeout_ns) {
  pdl_sync();
  exchange(c, 0, timeout_ns);
}

__global__ void wait_published_kernel(const uint32_t* sig, int rank, int world, uint64_t timeout_ns) {
  pdl_sync();
  wait_published(sig, rank, world, timeout_ns);
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

static int fill_rank_args(RankArgs* ra, const b200_fused_update_args* u) {
  memset(ra, 0, sizeof(*ra));
  int rc = fill_ctx(u->ctx, &ra->c);
  if (rc) return rc;
  const int world = u->ctx->world;
  B200_REQUIRE(u->nseg >= 0 && u->segs_dev != nullptr, "comm: no segments");
  B200_REQUIRE(u->nmetrics >= 0 && u->nmetrics <= 8, "comm: at most 8 metrics");
  for (int i = 0; i < world; ++i) {
    B200_REQUIRE(ra->c.grad[i] != nullptr, "comm: missing gradient buffer of rank %d", i);
    if (u->broadcast) B200_REQUIRE(ra->c.param[i] != nullptr, "comm: missing parameter buffer of rank %d", i);
  }
  if (!u->broadcast) B200_REQUIRE(u->p_local != nullptr, "comm: local mode needs p_local");
  B200_REQUIRE(u->hp != nullptr && u->step_count != nullptr, "comm: missing hyper-parameters or step counter");
  ra->segs = u->segs_dev;
  ra->nseg = u->nseg;
  ra->broadcast = u->broadcast;
  ra->publish_fp32 = u->publish_fp32;
  ra->nmetrics = u->nmetrics;
  ra->p_local = u->p_local;
  ra->bf16_local = static_cast<__nv_bfloat16*>(u->bf16_local);
  ra->m = u->m;
  ra->v = u->v;
  ra->hp = *u->hp;
  ra->step_count = u->step_count;
  ra->step_metrics = u->step_metrics;
  ra->metrics_acc = u->metrics_acc;
  if (u->end != nullptr) {
    ra->end = *u->end;
    B200_REQUIRE(ra->end.rng_key != nullptr && ra->end.step_count != nullptr && ra->end.done_word != nullptr,
                 "comm: step_end needs rng_key, step_count and a zero-initialised done_word");
  }
  ra->timeout_ns = comm_timeout_ns();
  ra->stamps = g_stamps;
  ra->defer_publish = comm_defer_publish() ? 1 : (g_defer_publish == 2 ? 2 : 0);
  return 0;
}

static int clamp_blocks(int blocks, int world, int vranks) {
  if (blocks <= 0) blocks = 64;
  if (world == 1 && vranks == 1) return 2 * num_sms();   // nobody to wait for: fill the machine (two resident blocks per SM)
  if (blocks > B200_COMM_MAX_BLOCKS) blocks = B200_COMM_MAX_BLOCKS;
  const int cap = (vranks > 1 ? 2 * num_sms() : num_sms()) / vranks;   // every block must be resident: they wait for each other
  if (blocks > cap) blocks = cap;
  return blocks;
}

}  // namespace b200

using namespace b200;

extern "C" void b200_comm_set_timeout_ms(long long ms) {
  g_timeout_ns = ms > 0 ? static_cast<uint64_t>(ms) * 1000000ull : 0;
}

extern "C" void b200_comm_set_defer_publish(int on) { g_defer_publish = on < 0 ? -1 : (on > 2 ? 0 : on); }
extern "C" int b200_comm_get_defer_publish(void) { return comm_defer_publish() ? 1 : 0; }

extern "C" void b200_comm_set_stamps(void* dev_u64) { g_stamps = static_cast<unsigned long long*>(dev_u64); }

extern "C" int b200_fused_update(const b200_fused_update_args* u, void* stream) {
  RankArgs ra;
  int rc = fill_rank_args(&ra, u);
  if (rc) return rc;
  const int blocks = clamp_blocks(u->blocks, u->ctx->world, 1);
  B200_CUDA(launch_k(fused_reduce_adamw_kernel, dim3(blocks), dim3(256), 0, static_cast<cudaStream_t>(stream), ra));
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_fused_update_emulated(const b200_fused_update_args* ranks, int nranks, void* scratch_dev,
                                          void* stream) {
  B200_REQUIRE(nranks >= 1 && nranks <= MAX_WORLD, "comm: %d emulated ranks not in [1,%d]", nranks, MAX_WORLD);
  B200_REQUIRE(scratch_dev != nullptr, "comm: emulation needs a device scratch buffer of b200_fused_update_emulated_scratch_bytes()");
  RankArgs ra[MAX_WORLD];
  for (int r = 0; r < nranks; ++r) {
    B200_REQUIRE(ranks[r].ctx->world == nranks && ranks[r].ctx->rank == r, "comm: emulated rank %d has rank %d / world %d",
                 r, ranks[r].ctx->rank, ranks[r].ctx->world);
    B200_REQUIRE(ranks[r].ctx->mc_grad == nullptr && ranks[r].ctx->mc_param == nullptr && ranks[r].ctx->mc_bf16 == nullptr,
                 "comm: the emulation has no multicast addresses");
    int rc = fill_rank_args(&ra[r], &ranks[r]);
    if (rc) return rc;
  }
  const int blocks = clamp_blocks(ranks[0].blocks, nranks, nranks);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  B200_CUDA(cudaMemcpyAsync(scratch_dev, ra, sizeof(RankArgs) * nranks, cudaMemcpyHostToDevice, s));
  const RankArgs* dev = static_cast<const RankArgs*>(scratch_dev);
  void* kargs[] = {const_cast<RankArgs**>(&dev)};
  count_launch();
  B200_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(fused_reduce_adamw_emul_kernel), dim3(blocks, nranks), dim3(256),
                                        kargs, 0, s));
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" long long b200_fused_update_emulated_scratch_bytes(void) {
  const size_t a = sizeof(RankArgs), b = sizeof(CollArgs);
  return static_cast<long long>((a > b ? a : b) * MAX_WORLD);
}

// ---- the earlier, positional entry points (kept: they forward to b200_fused_update) ----
extern "C" int b200_fused_reduce_adamw_end(const b200_comm_ctx* ctx, const b200_seg* segs_dev, int nseg,
                                           int broadcast, float* p_local, void* bf16_local, float* m, float* v,
                                           const b200_adamw_hparams* hp, int32_t* step_count, int bump_step,
                                           const float* step_metrics, float* metrics_acc, int nmetrics,
                                           int blocks, const b200_step_end* end, void* stream) {
  B200_REQUIRE(!(end != nullptr && bump_step), "comm: bump_step and step_end both advance the step counter");
  b200_fused_update_args u;
  memset(&u, 0, sizeof(u));
  u.ctx = ctx; u.segs_dev = segs_dev; u.nseg = nseg; u.broadcast = broadcast; u.publish_fp32 = 1;
  u.p_local = p_local; u.bf16_local = bf16_local; u.m = m; u.v = v; u.hp = hp; u.step_count = step_count;
  u.step_metrics = step_metrics; u.metrics_acc = metrics_acc; u.nmetrics = nmetrics; u.blocks = blocks; u.end = end;
  int rc = b200_fused_update(&u, stream);
  if (rc) return rc;
  if (bump_step) {
    B200_CUDA(launch_k(bump_step_kernel2, dim3(1), dim3(1), 0, static_cast<cudaStream_t>(stream), step_count));
    B200_CUDA(cudaGetLastError());
  }
  return 0;
}

extern "C" int b200_fused_reduce_adamw(const b200_comm_ctx* ctx, const b200_seg* segs_dev, int nseg,
                                       int broadcast, float* p_local, void* bf16_local, float* m, float* v,
                                       const b200_adamw_hparams* hp, int32_t* step_count, int bump_step,
                                       const float* step_metrics, float* metrics_acc, int nmetrics,
                                       int blocks, void* stream) {
  return b200_fused_reduce_adamw_end(ctx, segs_dev, nseg, broadcast, p_local, bf16_local, m, v, hp, step_count, bump_step,
                                     step_metrics, metrics_acc, nmetrics, blocks, nullptr, stream);
}

static int fill_coll(CollArgs* a, const b200_collective_args* u) {
  memset(a, 0, sizeof(*a));
  B200_REQUIRE(u->world >= 1 && u->world <= MAX_WORLD && u->rank >= 0 && u->rank < u->world, "collective: bad rank %d / world %d",
               u->rank, u->world);
  B200_REQUIRE(u->mode >= 0 && u->mode <= 2, "collective: unknown mode %d", u->mode);
  B200_REQUIRE(u->n > 0, "collective: empty buffer");
  if (u->mode != 2) B200_REQUIRE(u->n % (4 * u->world) == 0, "collective: n = %lld must be a multiple of 4 x world", static_cast<long long>(u->n));
  if (u->mode != 0) B200_REQUIRE(u->out != nullptr, "collective: mode %d needs an output buffer", u->mode);
  a->rank = u->rank; a->world = u->world; a->mode = u->mode;
  for (int i = 0; i < u->world; ++i) {
    B200_REQUIRE(u->peer_buf[i] != nullptr && u->peer_signal[i] != nullptr, "collective: missing buffer of rank %d", i);
    a->buf[i] = static_cast<float*>(u->peer_buf[i]);
    a->signal[i] = static_cast<uint32_t*>(u->peer_signal[i]);
  }
  a->mc_buf = static_cast<float*>(u->mc_buf);
  a->out = static_cast<float*>(u->out);
  a->n = u->n;
  a->scale = u->scale;
  a->timeout_ns = comm_timeout_ns();
  return 0;
}

extern "C" int b200_fused_collective(const b200_collective_args* u, void* stream) {
  CollArgs a;
  int rc = fill_coll(&a, u);
  if (rc) return rc;
  const int blocks = clamp_blocks(u->blocks, u->world, 1);
  B200_CUDA(launch_k(collective_kernel, dim3(blocks), dim3(256), 0, static_cast<cudaStream_t>(stream), a));
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_fused_collective_emulated(const b200_collective_args* ranks, int nranks, void* scratch_dev, void* stream) {
  B200_REQUIRE(nranks >= 1 && nranks <= MAX_WORLD, "collective: %d emulated ranks not in [1,%d]", nranks, MAX_WORLD);
  B200_REQUIRE(scratch_dev != nullptr, "collective: emulation needs a device scratch buffer");
  CollArgs ca[MAX_WORLD];
  for (int r = 0; r < nranks; ++r) {
    B200_REQUIRE(ranks[r].world == nranks && ranks[r].rank == r && ranks[r].mc_buf == nullptr, "collective: bad emulated rank %d", r);
    int rc = fill_coll(&ca[r], &ranks[r]);
    if (rc) return rc;
  }
  const int blocks = clamp_blocks(ranks[0].blocks, nranks, nranks);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  B200_CUDA(cudaMemcpyAsync(scratch_dev, ca, sizeof(CollArgs) * nranks, cudaMemcpyHostToDevice, s));
  const CollArgs* dev = static_cast<const CollArgs*>(scratch_dev);
  void* kargs[] = {const_cast<CollArgs**>(&dev)};
  count_launch();
  B200_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(collective_emul_kernel), dim3(blocks, nranks), dim3(256), kargs, 0, s));
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_signal_barrier(const b200_comm_ctx* ctx, void* stream) {
  CommDev d;
  int rc = fill_ctx(ctx, &d);
  if (rc) return rc;
  B200_CUDA(launch_k(signal_barrier_kernel, dim3(1), dim3(32), 0, static_cast<cudaStream_t>(stream), d, comm_timeout_ns()));
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_comm_wait_published(const void* signal_local, int rank, int world, void* stream) {
  B200_REQUIRE(world >= 1 && world <= MAX_WORLD && rank >= 0 && rank < world, "wait_published: bad rank %d / world %d", rank, world);
  if (world == 1) return 0;
  B200_REQUIRE(signal_local != nullptr, "wait_published: no signal buffer");
  B200_CUDA(launch_k(wait_published_kernel, dim3(1), dim3(32), 0, static_cast<cudaStream_t>(stream),
                     static_cast<const uint32_t*>(signal_local), rank, world, comm_timeout_ns()));
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
