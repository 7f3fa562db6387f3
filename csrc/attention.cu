// Causal self-attention on tcgen05 tensor cores (head_dim 64), forward and backward.
// Not in the reference (it has no attention op); required by the GPT-2 configurations of
// BASELINE.json. Flash-attention dataflow re-cut for Blackwell: both GEMMs of every tile run on
// tcgen05.mma with accumulators in TMEM, K/V/Q/dO tiles arrive by TMA into 128B-swizzled shared
// memory, softmax runs in registers on rows read back with tcgen05.ld (one thread per row, no
// shuffles), and P / dS go back to the tensor cores through shared memory written in the UMMA
// canonical swizzled layout.
//
// Layout: qkv bf16 [batch][seq][3][heads][64] (the output of the fused QKV projection, read in
// place through one 2-D tensor map), out bf16 [batch][seq][heads][64], lse fp32
// [batch][heads][seq] in log2 units of the scaled scores.
//
// Warp roles: warp 0 TMA producer, warp 1 MMA issuer (+ TMEM owner), the rest math warps.
// Forward: 4 math warps (a thread owns one query row); two CTAs share an SM and fill each other's
// MMA <-> softmax dependency bubbles. Backward: 8 math warps, two per TMEM lane quadrant, a thread
// owns one query row and half of the tile's columns.
#include "../include/b200dist.h"
#include "common.cuh"

namespace b200 {

constexpr int HD = 64;        // head dim
constexpr int BQ = 128;       // query rows per tile
constexpr int BKV = 128;      // key/value rows per tile
constexpr int TILE_BYTES = 128 * 64 * 2;   // one [128][64] bf16 swizzled tile = 16 KB
constexpr int ATT_FWD_THREADS = 192;   // TMA warp + MMA warp + 4 math warps (one query row per thread); two CTAs per SM
constexpr int ATT_BWD_THREADS = 320;   // TMA warp + MMA warp + 8 math warps (two per TMEM lane quadrant, half the columns each)

__device__ __forceinline__ uint64_t kdesc(uint32_t addr) { return umma_smem_desc(addr, 16, 1024); }
__device__ __forceinline__ uint64_t mndesc(uint32_t addr, uint32_t lbo) { return umma_smem_desc(addr, lbo, 1024); }

// Store 8 consecutive bf16 (16 bytes) of row `r`, columns [c, c+8) of a [128][128] operand kept
// as two [128][64] K-major 128B-swizzled tiles.
__device__ __forceinline__ void st_swizzled_16B(uint8_t* base, int r, int c, uint4 v) {
  const int tile = c >> 6;
  const int chunk = (c & 63) >> 3;
  uint8_t* p = base + tile * TILE_BYTES + r * 128 + ((chunk ^ (r & 7)) << 4);
  *reinterpret_cast<uint4*>(p) = v;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}

__device__ __forceinline__ void named_barrier_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// =============================================================================================
// forward
// =============================================================================================
struct AttnFwdSmem {
  // offsets in bytes from the 1024-aligned base
  static constexpr int Q = 0;
  static constexpr int K0 = Q + TILE_BYTES;            // 2 stages
  static constexpr int V0 = K0 + 2 * TILE_BYTES;       // 2 stages
  static constexpr int P = V0 + 2 * TILE_BYTES;        // 2 tiles
  static constexpr int BAR = P + 2 * TILE_BYTES;
  // No alignment slack: two CTAs must fit in one SM's 228 KB (2 x (BYTES + 1 KB reserved) <= 233472),
  // so the kernel requires (and checks) a 1024-byte aligned dynamic shared memory base.
  static constexpr int BYTES = BAR + 128;
};
static_assert(2 * (AttnFwdSmem::BYTES + 1024) <= 233472, "forward attention: two CTAs per SM must fit");

// Work item = one (q tile, head, sequence). Items are numbered heavy-first (late q tiles see the
// most kv tiles) and dealt to the persistent CTAs in snake order, so every CTA gets nearly the
// same number of kv steps.
struct FwdItem {
  int qt, h, b;
};
__device__ __forceinline__ int fwd_item_index(int round, int n_cta, int cta) {
  return round * n_cta + ((round & 1) ? n_cta - 1 - cta : cta);
}
__device__ __forceinline__ FwdItem fwd_item_decode(int item, int nq, int heads, int hb) {
  FwdItem w;
  const int q = item / hb;
  const int rem = item - q * hb;
  w.qt = nq - 1 - q;
  w.b = rem / heads;
  w.h = rem - w.b * heads;
  return w;
}

constexpr float kRescaleThreshold = 8.0f;   // log2 units: P stays below 2^8, far inside bf16/fp32 range

// One kv step of the softmax for row `r`: S (TMEM) -> registers, row maximum, lazily rescaled running
// state, probabilities -> bf16 -> swizzled shared memory (A operand of the P V MMA).
// The output accumulator O stays in TMEM: the P V MMAs accumulate into it, and it is only
// multiplied by 2^(m_old - m_new) when the row maximum grew by more than kRescaleThreshold
// (otherwise the stale maximum keeps being used for P and l, which is exact after the final 1/l).
template <bool DIAG>
__device__ __forceinline__ void fwd_softmax_step(uint32_t tmem_s_lane, uint32_t tmem_o_lane, uint8_t* sp, int r, int lane,
                                                 float scale_log2, bool first, float& m_used, float& l_run,
                                                 uint64_t* s_free, uint64_t* o_full, uint32_t o_prev_parity,
                                                 uint64_t* p_full, bool store_issuer) {
  // pass 1: row maximum, 32 columns at a time (few live registers: the O rescale below needs room)
  float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    uint32_t v[32];
    tmem_ld_32x32b_x32(tmem_s_lane + c * 32, v);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      float x = __uint_as_float(v[i]);
      if (DIAG && (c * 32 + i) > r) x = -INFINITY;
      mx[i & 3] = fmaxf(mx[i & 3], x);
    }
  }
  const float m_tile = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])) * scale_log2;
  bool o_seen = first;                      // first step of an item: no earlier P V to wait for
  if (first) {
    m_used = m_tile;
  } else {
    const bool need = m_tile > m_used + kRescaleThreshold;
    if (__any_sync(0xffffffffu, need)) {
      mbar_wait(o_full, o_prev_parity);     // every P V issued so far has landed in O
      tc_fence_after();
      o_seen = true;
      const float alpha = need ? fast_exp2(m_used - m_tile) : 1.0f;
      if (need) m_used = m_tile;
      l_run *= alpha;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t o[32];
        tmem_ld_32x32b_x32(tmem_o_lane + c * 32, o);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
        tmem_st_32x32b_x32(tmem_o_lane + c * 32, o);
      }
      tmem_st_wait();
    }
  }
  // pass 2: the whole row into registers, then S is free for the next tile's Q K^T
  uint32_t sv[4][32];
#pragma unroll
  for (int c = 0; c < 4; ++c) tmem_ld_32x32b_x32(tmem_s_lane + c * 32, sv[c]);
  tmem_ld_wait();
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(s_free);
  float ls[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int c = 0; c < 4; ++c)
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      float e = fast_exp2(fmaf(__uint_as_float(sv[c][i]), scale_log2, -m_used));
      if (DIAG && (c * 32 + i) > r) e = 0.f;
      ls[i & 3] += e;
      sv[c][i] = __float_as_uint(e);
    }
  l_run += (ls[0] + ls[1]) + (ls[2] + ls[3]);
  if (!o_seen) {                            // P buffer: the previous step's P V must have read it
    mbar_wait(o_full, o_prev_parity);
    tc_fence_after();
  }
  if (first) {                              // ... and so must the previous item's output tile (TMA store)
    if (store_issuer) tma_store_wait_read<0>();
    named_barrier_sync(2, 128);
  }
#pragma unroll
  for (int c = 0; c < 4; ++c)
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      uint4 w;
      w.x = pack_bf16x2(__uint_as_float(sv[c][g * 8 + 0]), __uint_as_float(sv[c][g * 8 + 1]));
      w.y = pack_bf16x2(__uint_as_float(sv[c][g * 8 + 2]), __uint_as_float(sv[c][g * 8 + 3]));
      w.z = pack_bf16x2(__uint_as_float(sv[c][g * 8 + 4]), __uint_as_float(sv[c][g * 8 + 5]));
      w.w = pack_bf16x2(__uint_as_float(sv[c][g * 8 + 6]), __uint_as_float(sv[c][g * 8 + 7]));
      st_swizzled_16B(sp, r, c * 32 + g * 8, w);
    }
  tc_fence_before();
  fence_proxy_async_smem();     // generic-proxy smem writes -> visible to the tensor core
  __syncwarp();
  if (lane == 0) mbar_arrive(p_full);
}

// Persistent: gridDim.x CTAs (two per SM) walk the work items; all pipelines (K and V stages, S, P,
// O) keep running across item boundaries, so the next item's Q/K loads and first S = Q K^T are
// in flight while the math warps finish the current item.
__global__ void __launch_bounds__(ATT_FWD_THREADS, 2)
attention_fwd_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_out,
                     float* __restrict__ lse, int seq, int heads, int batch, float scale_log2) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem) & 1023u) != 0) {      // swizzled TMA / UMMA tiles need a 1024-byte aligned base
    if (threadIdx.x == 0) printf("b200: attention_fwd shared memory base is not 1024-byte aligned\n");
    __trap();
  }
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + AttnFwdSmem::BAR);
  uint64_t* q_full = bars + 0;
  uint64_t* q_empty = bars + 1;
  uint64_t* k_full = bars + 2;    // [2]
  uint64_t* k_empty = bars + 4;   // [2]
  uint64_t* v_full = bars + 6;    // [2]
  uint64_t* v_empty = bars + 8;   // [2]
  uint64_t* s_full = bars + 10;
  uint64_t* s_free = bars + 11;
  uint64_t* p_full = bars + 12;
  uint64_t* o_full = bars + 13;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 14);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = seq / BQ;
  const int hb = heads * batch;
  const int n_items = nq * hb;
  const int n_cta = gridDim.x, cta = blockIdx.x;
  const int C = heads * HD;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_qkv);
    tma_prefetch_desc(&tm_out);
    mbar_init(q_full, 1);
    mbar_init(q_empty, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&k_full[s], 1);
      mbar_init(&k_empty[s], 1);
      mbar_init(&v_full[s], 1);
      mbar_init(&v_empty[s], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(s_free, 4);
    mbar_init(p_full, 4);
    mbar_init(o_full, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<256>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_sync();
  const uint32_t tmem_s = tmem_base;          // 128 columns
  const uint32_t tmem_o = tmem_base + 128;    // 64 columns

  if (warp == 0) {
    if (elect_one()) {
      int g = 0, n = 0;       // g: kv steps issued so far (all items), n: items so far
      for (int round = 0;; ++round, ++n) {
        const int item = fwd_item_index(round, n_cta, cta);
        if (item >= n_items) break;
        const FwdItem w = fwd_item_decode(item, nq, heads, hb);
        const int row0 = w.b * seq;
        mbar_wait(q_empty, (n & 1) ^ 1);
        mbar_expect_tx(q_full, TILE_BYTES);
        tma_load_2d(smem + AttnFwdSmem::Q, &tm_qkv, q_full, w.h * HD, row0 + w.qt * BQ);
        for (int j = 0; j <= w.qt; ++j, ++g) {
          const int s = g & 1;
          const uint32_t ph = ((g >> 1) & 1) ^ 1;
          mbar_wait(&k_empty[s], ph);
          mbar_expect_tx(&k_full[s], TILE_BYTES);
          tma_load_2d(smem + AttnFwdSmem::K0 + s * TILE_BYTES, &tm_qkv, &k_full[s], C + w.h * HD, row0 + j * BKV);
          mbar_wait(&v_empty[s], ph);
          mbar_expect_tx(&v_full[s], TILE_BYTES);
          tma_load_2d(smem + AttnFwdSmem::V0 + s * TILE_BYTES, &tm_qkv, &v_full[s], 2 * C + w.h * HD, row0 + j * BKV);
        }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, 0, 0);   // S = Q K^T
    constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, 1);    // O += P V (V is MN-major)
    const uint32_t sq = smem_u32(smem + AttnFwdSmem::Q);
    const uint32_t sp = smem_u32(smem + AttnFwdSmem::P);
    // S(g) = Q K(g)^T into the single S accumulator; `last` also releases the Q tile
    auto issue_s = [&](int g, bool last) {
      const int s = g & 1;
      const uint32_t sk = smem_u32(smem + AttnFwdSmem::K0 + s * TILE_BYTES);
      mbar_wait(&k_full[s], (g >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < HD / 16; ++k) umma_bf16(tmem_s, kdesc(sq + k * 32), kdesc(sk + k * 32), idesc_s, k ? 1u : 0u);
        umma_commit(s_full);
        umma_commit(&k_empty[s]);
        if (last) umma_commit(q_empty);
      }
      __syncwarp();
    };
    int g = 0, n = 0;
    for (int round = 0;; ++round, ++n) {
      const int item = fwd_item_index(round, n_cta, cta);
      if (item >= n_items) break;
      const int nkv = (nq - 1 - item / hb) + 1;
      mbar_wait(q_full, n & 1);
      if (g > 0) mbar_wait(s_free, (g - 1) & 1);     // the previous item's last scores have been read
      issue_s(g, nkv == 1);
      for (int j = 0; j < nkv; ++j, ++g) {
        if (j + 1 < nkv) {
          mbar_wait(s_free, g & 1);                  // S(g) is in registers: overwrite it under the exponentials
          issue_s(g + 1, j + 2 == nkv);
        }
        const int s = g & 1;
        const uint32_t sv = smem_u32(smem + AttnFwdSmem::V0 + s * TILE_BYTES);
        mbar_wait(p_full, g & 1);
        mbar_wait(&v_full[s], (g >> 1) & 1);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < BKV / 16; ++k)
            umma_bf16(tmem_o, kdesc(sp + (k >> 2) * TILE_BYTES + (k & 3) * 32), mndesc(sv + k * 2048, TILE_BYTES),
                      idesc_o, (j | k) ? 1u : 0u);
          umma_commit(o_full);
          umma_commit(&v_empty[s]);
        }
        __syncwarp();
      }
    }
  } else {
    const int quad = warp & 3;
    const int r = quad * 32 + lane;                 // row within the q tile
    const uint32_t lane_off = static_cast<uint32_t>(quad * 32) << 16;
    const bool store_issuer = (threadIdx.x == 64);  // first math thread issues the output TMA stores
    uint8_t* sp = smem + AttnFwdSmem::P;
    int g = 0;
    for (int round = 0;; ++round) {
      const int item = fwd_item_index(round, n_cta, cta);
      if (item >= n_items) break;
      const FwdItem w = fwd_item_decode(item, nq, heads, hb);
      float m_used = -INFINITY, l_run = 0.f;
      for (int j = 0; j < w.qt; ++j, ++g) {
        mbar_wait(s_full, g & 1);
        tc_fence_after();
        fwd_softmax_step<false>(tmem_s + lane_off, tmem_o + lane_off, sp, r, lane, scale_log2, j == 0, m_used, l_run,
                                s_free, o_full, (g - 1) & 1, p_full, store_issuer);
      }
      mbar_wait(s_full, g & 1);     // the diagonal tile: the only one that needs the causal mask
      tc_fence_after();
      fwd_softmax_step<true>(tmem_s + lane_off, tmem_o + lane_off, sp, r, lane, scale_log2, w.qt == 0, m_used, l_run,
                             s_free, o_full, (g - 1) & 1, p_full, store_issuer);
      mbar_wait(o_full, g & 1);     // last P V of the item: O complete, P buffer free
      tc_fence_after();
      ++g;
      // O / l -> bf16 -> swizzled tile in the (idle) P buffer -> one TMA store
      const float inv = 1.0f / l_run;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t o[32];
        tmem_ld_32x32b_x32(tmem_o + lane_off + c * 32, o);
        tmem_ld_wait();
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint4 wv;
          wv.x = pack_bf16x2(__uint_as_float(o[q * 8 + 0]) * inv, __uint_as_float(o[q * 8 + 1]) * inv);
          wv.y = pack_bf16x2(__uint_as_float(o[q * 8 + 2]) * inv, __uint_as_float(o[q * 8 + 3]) * inv);
          wv.z = pack_bf16x2(__uint_as_float(o[q * 8 + 4]) * inv, __uint_as_float(o[q * 8 + 5]) * inv);
          wv.w = pack_bf16x2(__uint_as_float(o[q * 8 + 6]) * inv, __uint_as_float(o[q * 8 + 7]) * inv);
          st_swizzled_16B(sp, r, c * 32 + q * 8, wv);
        }
      }
      tc_fence_before();            // O reads done before the next item's first P V (ordered by p_full)
      lse[(static_cast<size_t>(w.b) * heads + w.h) * seq + w.qt * BQ + r] = m_used + __log2f(l_run);
      fence_proxy_async_smem();
      named_barrier_sync(1, 128);
      if (store_issuer) {
        tma_store_2d(&tm_out, sp, w.h * HD, w.b * seq + w.qt * BQ);
        tma_store_commit();
      }
    }
    if (store_issuer) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<256>(tmem_base);
  }
}

// =============================================================================================
// backward
// =============================================================================================
// delta[b,h,t] = sum_d dO*O ; dq_accum zeroed. 8 lanes per (b,t,h) row of 64, 16-byte loads/stores.
__global__ void __launch_bounds__(256)
attention_bwd_prep_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ d_o,
                          float* __restrict__ delta, float* __restrict__ dq_accum, int batch, int seq, int heads) {
  pdl_sync();
  const uint32_t rows = static_cast<uint32_t>(batch) * seq * heads;
  const uint32_t total = rows * 8u;                      // uint4 chunks (host checks rows < 2^28)
  const uint32_t useq = seq, uheads = heads;
  // warp-uniform trip count: the shuffles below need all 32 lanes
  for (uint32_t base = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; base < total; base += gridDim.x * blockDim.x) {
    const uint32_t i = base + (threadIdx.x & 31u);
    float s = 0.f;
    const bool live = i < total;
    if (live) {
      const uint4 a = reinterpret_cast<const uint4*>(o)[i];
      const uint4 g = reinterpret_cast<const uint4*>(d_o)[i];
      const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, gw[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 af = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&aw[k]));
        const float2 gf = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&gw[k]));
        s = fmaf(af.x, gf.x, fmaf(af.y, gf.y, s));
      }
      float4* z = reinterpret_cast<float4*>(dq_accum) + static_cast<size_t>(i) * 2;
      z[0] = make_float4(0.f, 0.f, 0.f, 0.f);
      z[1] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    if (live && (i & 7u) == 0) {
      const uint32_t row = i >> 3;
      const uint32_t bt = row / uheads, hh = row - bt * uheads;
      const uint32_t bb = bt / useq, t = bt - bb * useq;
      delta[(static_cast<size_t>(bb) * uheads + hh) * useq + t] = s;
    }
  }
}

// dq bf16 <- dq_accum fp32, written into dqkv[..., 0, h, :]; 8 values per thread.
// dq_accum is tile-major: [(b*heads + h)*nq + q tile][128 rows][64] fp32, and within a row the
// sixteen 16-byte chunks are stored at position chunk ^ (row & 15) (the layout the backward kernel
// stages in shared memory so that its bulk reduce-add is one contiguous 32 KB transfer).
__global__ void __launch_bounds__(256)
attention_bwd_post_kernel(const float* __restrict__ dq_accum, __nv_bfloat16* __restrict__ dqkv, int batch, int seq,
                          int heads) {
  pdl_sync();
  const uint32_t uheads = heads, useq = seq, nq = useq / BQ;
  const uint32_t total = static_cast<uint32_t>(batch) * useq * uheads * 8u;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const uint32_t c8 = i & 7u;
    const uint32_t rh = i >> 3;
    const uint32_t bt = rh / uheads, hh = rh - bt * uheads;
    const uint32_t bb = bt / useq, t = bt - bb * useq;
    const uint32_t qi = t / BQ, r = t - qi * BQ;
    const float* src = dq_accum + (static_cast<size_t>((bb * uheads + hh) * nq + qi) * BQ + r) * HD;
    const float4 v0 = *reinterpret_cast<const float4*>(src + (((2 * c8) ^ (r & 15u)) << 2));
    const float4 v1 = *reinterpret_cast<const float4*>(src + (((2 * c8 + 1) ^ (r & 15u)) << 2));
    uint4 w;
    w.x = pack_bf16x2(v0.x, v0.y);
    w.y = pack_bf16x2(v0.z, v0.w);
    w.z = pack_bf16x2(v1.x, v1.y);
    w.w = pack_bf16x2(v1.z, v1.w);
    *reinterpret_cast<uint4*>(dqkv + static_cast<size_t>(bt) * 3 * uheads * HD + hh * HD + c8 * 8) = w;
  }
}

struct AttnBwdSmem {
  static constexpr int K0 = 0;                           // 2 stages (items alternate)
  static constexpr int V0 = K0 + 2 * TILE_BYTES;         // 2 stages
  static constexpr int Q0 = V0 + 2 * TILE_BYTES;         // 2 stages
  static constexpr int DO0 = Q0 + 2 * TILE_BYTES;        // 2 stages
  static constexpr int P = DO0 + 2 * TILE_BYTES;         // 2 tiles
  static constexpr int DS = P + 2 * TILE_BYTES;          // 2 tiles
  static constexpr int DQ = DS + 2 * TILE_BYTES;         // fp32 [128][64] staging for the bulk reduce-add
  static constexpr int BAR = DQ + 2 * TILE_BYTES;
  static constexpr int BYTES = BAR + 128;                // 1024-aligned base required (checked in the kernel)
};
static_assert(AttnBwdSmem::BYTES + 1024 <= 232448, "backward attention: shared memory budget");

// Work item = one (kv tile, head, sequence); kv tile jt meets q tiles jt..nq-1. Heavy-first
// numbering and snake dealing as in the forward.
struct BwdItem {
  int jt, h, b;
};
__device__ __forceinline__ BwdItem bwd_item_decode(int item, int heads, int hb) {
  BwdItem w;
  w.jt = item / hb;
  const int rem = item - w.jt * hb;
  w.b = rem / heads;
  w.h = rem - w.b * heads;
  return w;
}

// One q tile of the backward math for this warp's 64 columns (`half`) of row r:
//   P = exp2(S * scale_log2 - lse), dS = P * (dP - delta) * scale
// S and dP are pulled out of TMEM in one go (then released to the MMA warp through `sdp_free`), the
// results are kept packed in registers until the previous tile's MMAs have stopped reading the
// P / dS shared-memory operands (`prev_mma_done`), and only then stored.
template <bool DIAG>
__device__ __forceinline__ void bwd_math_tile(uint32_t tm_s, uint32_t tm_dp, int half, int r, int lane, float lse_r,
                                              float dls, float scale, float scale_log2, uint8_t* sp, uint8_t* sds,
                                              uint64_t* sdp_free, uint64_t* prev_mma_done, uint32_t prev_parity,
                                              bool wait_prev) {
  uint32_t sv_[2][32], dv_[2][32];
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    tmem_ld_32x32b_x32(tm_s + half * 64 + c * 32, sv_[c]);
    tmem_ld_32x32b_x32(tm_dp + half * 64 + c * 32, dv_[c]);
  }
  tmem_ld_wait();
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(sdp_free);     // S / dP accumulators may be overwritten by the next q tile
  uint32_t pp[2][16], dd[2][16];
#pragma unroll
  for (int c = 0; c < 2; ++c)
#pragma unroll
    for (int i = 0; i < 32; i += 2) {
      float e0 = fast_exp2(fmaf(__uint_as_float(sv_[c][i]), scale_log2, -lse_r));
      float e1 = fast_exp2(fmaf(__uint_as_float(sv_[c][i + 1]), scale_log2, -lse_r));
      if (DIAG && (half * 64 + c * 32 + i) > r) e0 = 0.f;
      if (DIAG && (half * 64 + c * 32 + i + 1) > r) e1 = 0.f;
      const float d0 = e0 * fmaf(__uint_as_float(dv_[c][i]), scale, -dls);
      const float d1 = e1 * fmaf(__uint_as_float(dv_[c][i + 1]), scale, -dls);
      pp[c][i >> 1] = pack_bf16x2(e0, e1);
      dd[c][i >> 1] = pack_bf16x2(d0, d1);
    }
  if (wait_prev) {
    mbar_wait(prev_mma_done, prev_parity);
    tc_fence_after();
  }
#pragma unroll
  for (int c = 0; c < 2; ++c)
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      st_swizzled_16B(sp, r, half * 64 + c * 32 + g * 8,
                      make_uint4(pp[c][g * 4 + 0], pp[c][g * 4 + 1], pp[c][g * 4 + 2], pp[c][g * 4 + 3]));
      st_swizzled_16B(sds, r, half * 64 + c * 32 + g * 8,
                      make_uint4(dd[c][g * 4 + 0], dd[c][g * 4 + 1], dd[c][g * 4 + 2], dd[c][g * 4 + 3]));
    }
}

// Persistent, one CTA per SM. Per item the q tiles jt..nq-1 are pipelined as
//   tensor pipe :  S,dP(0) | S,dP(1)  dV,dK,dQ(0) | S,dP(2)  dV,dK,dQ(1) | ...
//   math warps  :          | P,dS(0)              | P,dS(1)  dQ(0) out   | P,dS(2)  dQ(1) out ...
// S/dP(i+1) are issued as soon as the math warps hold S/dP(i) in registers, so they run under the
// exponentials; dV/dK/dQ(i) run under the math of tile i+1. dQ tiles leave through shared memory
// with one bulk reduce-add (cp.reduce.async.bulk ... add.f32) per tile into the fp32 accumulator
// that the CTAs of the other kv tiles add to as well. All pipelines run across item boundaries
// (K/V double-buffered), so the next item's loads and first S,dP overlap the current item's tail.
__global__ void __launch_bounds__(ATT_BWD_THREADS, 1)
attention_bwd_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_do,
                     const __grid_constant__ CUtensorMap tm_dqkv, const float* __restrict__ lse,
                     const float* __restrict__ delta, float* __restrict__ dq_accum, int seq, int heads, int batch,
                     float scale, float scale_log2) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem) & 1023u) != 0) {
    if (threadIdx.x == 0) printf("b200: attention_bwd shared memory base is not 1024-byte aligned\n");
    __trap();
  }
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + AttnBwdSmem::BAR);
  uint64_t* kv_full = bars + 0;    // [2]
  uint64_t* kv_empty = bars + 2;   // [2]
  uint64_t* qdo_full = bars + 4;   // [2]
  uint64_t* qdo_empty = bars + 6;  // [2]
  uint64_t* sdp_full = bars + 8;
  uint64_t* sdp_free = bars + 9;
  uint64_t* pds_full = bars + 10;
  uint64_t* dq_full = bars + 11;   // [2]
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 13);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = seq / BQ;
  const int hb = heads * batch;
  const int n_items = nq * hb;
  const int n_cta = gridDim.x, cta = blockIdx.x;
  const int C = heads * HD;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_qkv);
    tma_prefetch_desc(&tm_do);
    tma_prefetch_desc(&tm_dqkv);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&kv_full[s], 1);
      mbar_init(&kv_empty[s], 1);
      mbar_init(&qdo_full[s], 1);
      mbar_init(&qdo_empty[s], 1);
      mbar_init(&dq_full[s], 1);
    }
    mbar_init(sdp_full, 1);
    mbar_init(sdp_free, 8);
    mbar_init(pds_full, 8);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_sync();
  const uint32_t tm_s = tmem_base, tm_dp = tmem_base + 128, tm_dv = tmem_base + 256, tm_dk = tmem_base + 320,
                 tm_dq = tmem_base + 384;     // two 64-column dQ buffers: 384, 448

  if (warp == 0) {
    if (elect_one()) {
      int g = 0, n = 0;       // g: q-tile steps so far (all items), n: items so far
      for (int round = 0;; ++round, ++n) {
        const int item = fwd_item_index(round, n_cta, cta);
        if (item >= n_items) break;
        const BwdItem w = bwd_item_decode(item, heads, hb);
        const int row0 = w.b * seq;
        const int ks = n & 1;
        mbar_wait(&kv_empty[ks], ((n >> 1) & 1) ^ 1);
        mbar_expect_tx(&kv_full[ks], 2 * TILE_BYTES);
        tma_load_2d(smem + AttnBwdSmem::K0 + ks * TILE_BYTES, &tm_qkv, &kv_full[ks], C + w.h * HD, row0 + w.jt * BKV);
        tma_load_2d(smem + AttnBwdSmem::V0 + ks * TILE_BYTES, &tm_qkv, &kv_full[ks], 2 * C + w.h * HD, row0 + w.jt * BKV);
        for (int qi = w.jt; qi < nq; ++qi, ++g) {
          const int s = g & 1;
          mbar_wait(&qdo_empty[s], ((g >> 1) & 1) ^ 1);
          mbar_expect_tx(&qdo_full[s], 2 * TILE_BYTES);
          tma_load_2d(smem + AttnBwdSmem::Q0 + s * TILE_BYTES, &tm_qkv, &qdo_full[s], w.h * HD, row0 + qi * BQ);
          tma_load_2d(smem + AttnBwdSmem::DO0 + s * TILE_BYTES, &tm_do, &qdo_full[s], w.h * HD, row0 + qi * BQ);
        }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t id_s = umma_idesc_bf16(128, 128, 0, 0);    // S = Q K^T, dP = dO V^T
    constexpr uint32_t id_t = umma_idesc_bf16(128, 64, 1, 1);     // dV = P^T dO, dK = dS^T Q
    constexpr uint32_t id_q = umma_idesc_bf16(128, 64, 0, 1);     // dQ = dS K
    const uint32_t sp = smem_u32(smem + AttnBwdSmem::P), sds = smem_u32(smem + AttnBwdSmem::DS);
    auto issue_sdp = [&](int g, int ks) {      // S and dP of step g against the K/V stage ks
      const uint32_t sk = smem_u32(smem + AttnBwdSmem::K0 + ks * TILE_BYTES);
      const uint32_t sv = smem_u32(smem + AttnBwdSmem::V0 + ks * TILE_BYTES);
      const uint32_t sq = smem_u32(smem + AttnBwdSmem::Q0 + (g & 1) * TILE_BYTES);
      const uint32_t sdo = smem_u32(smem + AttnBwdSmem::DO0 + (g & 1) * TILE_BYTES);
      mbar_wait(&qdo_full[g & 1], (g >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < HD / 16; ++k) umma_bf16(tm_s, kdesc(sq + k * 32), kdesc(sk + k * 32), id_s, k ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k) umma_bf16(tm_dp, kdesc(sdo + k * 32), kdesc(sv + k * 32), id_s, k ? 1u : 0u);
        umma_commit(sdp_full);
      }
      __syncwarp();
    };
    int g = 0, n = 0;
    for (int round = 0;; ++round, ++n) {
      const int item = fwd_item_index(round, n_cta, cta);
      if (item >= n_items) break;
      const int niter = nq - item / hb;
      const int ks = n & 1;
      const uint32_t sk = smem_u32(smem + AttnBwdSmem::K0 + ks * TILE_BYTES);
      mbar_wait(&kv_full[ks], (n >> 1) & 1);
      if (g > 0) mbar_wait(sdp_free, (g - 1) & 1);     // the previous item's last S/dP have been read
      issue_sdp(g, ks);
      for (int it = 0; it < niter; ++it, ++g) {
        const int s = g & 1;
        const uint32_t sq = smem_u32(smem + AttnBwdSmem::Q0 + s * TILE_BYTES);
        const uint32_t sdo = smem_u32(smem + AttnBwdSmem::DO0 + s * TILE_BYTES);
        if (it + 1 < niter) {
          mbar_wait(sdp_free, g & 1);       // S/dP(g) are in the math warps' registers
          issue_sdp(g + 1, ks);
        }
        mbar_wait(pds_full, g & 1);         // P/dS(g) in shared memory
        tc_fence_after();
        if (elect_one()) {
          // contraction over the 128 query rows of this tile, 16 at a time
#pragma unroll
          for (int k = 0; k < BQ / 16; ++k) {
            const uint32_t acc = (it | k) ? 1u : 0u;
            umma_bf16(tm_dv, mndesc(sp + k * 2048, TILE_BYTES), mndesc(sdo + k * 2048, TILE_BYTES), id_t, acc);
            umma_bf16(tm_dk, mndesc(sds + k * 2048, TILE_BYTES), mndesc(sq + k * 2048, TILE_BYTES), id_t, acc);
          }
          // dQ tile: contraction over the 128 kv rows
#pragma unroll
          for (int k = 0; k < BKV / 16; ++k)
            umma_bf16(tm_dq + s * 64, kdesc(sds + (k >> 2) * TILE_BYTES + (k & 3) * 32), mndesc(sk + k * 2048, TILE_BYTES),
                      id_q, k ? 1u : 0u);
          umma_commit(&dq_full[s]);
          umma_commit(&qdo_empty[s]);
          if (it + 1 == niter) umma_commit(&kv_empty[ks]);
        }
        __syncwarp();
      }
    }
  } else {
    const int quad = warp & 3;
    const int half = (warp - 2) >> 2;     // columns [64*half, +64) of S / dP, [32*half, +32) of dQ / dK / dV
    const int r = quad * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(quad * 32) << 16;
    const bool issuer = (threadIdx.x == 64);      // first math thread issues the bulk reduce-adds
    uint8_t* sp = smem + AttnBwdSmem::P;
    uint8_t* sds = smem + AttnBwdSmem::DS;
    float* sdq = reinterpret_cast<float*>(smem + AttnBwdSmem::DQ);
    // dQ tile of step g (complete in TMEM once dq_full[g&1] has fired): TMEM -> swizzled fp32 staging
    // -> one 32 KB bulk reduce-add into the accumulator tile `tile`
    auto flush_dq = [&](int g, size_t tile) {
      uint32_t v[32];
      tmem_ld_32x32b_x32(tm_dq + (g & 1) * 64 + lane_off + half * 32, v);
      tmem_ld_wait();
      if (issuer) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // staging no longer read
      named_barrier_sync(1, 256);
#pragma unroll
      for (int q = 0; q < 8; ++q)
        *reinterpret_cast<uint4*>(sdq + r * HD + (((half * 8 + q) ^ (r & 15)) << 2)) =
            make_uint4(v[q * 4 + 0], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
      fence_proxy_async_smem();
      named_barrier_sync(2, 256);
      if (issuer) {
        float* dst = dq_accum + tile * (BQ * HD);
        asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(dst),
                     "r"(smem_u32(sdq)), "r"(BQ * HD * 4)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    };
    int g = 0;
    BwdItem w = bwd_item_decode(cta, heads, hb);      // round 0 item = this CTA's index (the grid never exceeds n_items)
    size_t stat0 = (static_cast<size_t>(w.b) * heads + w.h) * seq + r;
    float lse_r = lse[stat0 + w.jt * BQ];
    float dls = delta[stat0 + w.jt * BQ] * scale;
    for (int round = 0;; ++round) {
      const int niter = nq - w.jt;
      const size_t tile0 = static_cast<size_t>(w.b * heads + w.h) * nq + w.jt;
      for (int it = 0; it < niter; ++it, ++g) {
        // row statistics of the next q tile: in flight while this one is processed
        float lse_n = 0.f, dls_n = 0.f;
        if (it + 1 < niter) {
          lse_n = lse[stat0 + (w.jt + it + 1) * BQ];
          dls_n = delta[stat0 + (w.jt + it + 1) * BQ] * scale;
        }
        mbar_wait(sdp_full, g & 1);
        tc_fence_after();
        // The diagonal tile (q tile == kv tile, it == 0) is the only one that needs the causal mask.
        // P/dS stores wait for the previous step's MMAs (dq_full of step g-1); on the first step of
        // an item that barrier was already consumed by the previous item's tail.
        if (it == 0)
          bwd_math_tile<true>(tm_s + lane_off, tm_dp + lane_off, half, r, lane, lse_r, dls, scale, scale_log2, sp, sds,
                              sdp_free, nullptr, 0, false);
        else
          bwd_math_tile<false>(tm_s + lane_off, tm_dp + lane_off, half, r, lane, lse_r, dls, scale, scale_log2, sp, sds,
                               sdp_free, &dq_full[(g - 1) & 1], ((g - 1) >> 1) & 1, true);
        tc_fence_before();
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(pds_full);
        if (it > 0) flush_dq(g - 1, tile0 + it - 1);     // dq_full of step g-1 was observed inside bwd_math_tile
        lse_r = lse_n;
        dls = dls_n;
      }
      // next item's identity and first row statistics: loads in flight under this item's tail
      const int next_item = fwd_item_index(round + 1, n_cta, cta);
      const bool more = next_item < n_items;
      BwdItem wn = w;
      size_t stat0n = stat0;
      if (more) {
        wn = bwd_item_decode(next_item, heads, hb);
        stat0n = (static_cast<size_t>(wn.b) * heads + wn.h) * seq + r;
        lse_r = lse[stat0n + wn.jt * BQ];
        dls = delta[stat0n + wn.jt * BQ] * scale;
      }
      mbar_wait(&dq_full[(g - 1) & 1], ((g - 1) >> 1) & 1);
      tc_fence_after();
      flush_dq(g - 1, tile0 + niter - 1);
      // dK, dV of this kv tile (the last dq_full commit covers every MMA issued before it):
      // TMEM -> bf16 -> swizzled tiles in the dQ staging buffer -> two TMA stores
      uint32_t a[32], bb[32];
      tmem_ld_32x32b_x32(tm_dk + lane_off + half * 32, a);
      tmem_ld_32x32b_x32(tm_dv + lane_off + half * 32, bb);
      tmem_ld_wait();
      tc_fence_before();      // dK/dV reads precede the next item's first accumulate=0 MMA (ordered by pds_full)
      if (issuer) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // last dQ reduce has read the staging
      named_barrier_sync(1, 256);
      uint8_t* sdk = smem + AttnBwdSmem::DQ;
      uint8_t* sdv = sdk + TILE_BYTES;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 wv, u;
        wv.x = pack_bf16x2(__uint_as_float(a[q * 8 + 0]), __uint_as_float(a[q * 8 + 1]));
        wv.y = pack_bf16x2(__uint_as_float(a[q * 8 + 2]), __uint_as_float(a[q * 8 + 3]));
        wv.z = pack_bf16x2(__uint_as_float(a[q * 8 + 4]), __uint_as_float(a[q * 8 + 5]));
        wv.w = pack_bf16x2(__uint_as_float(a[q * 8 + 6]), __uint_as_float(a[q * 8 + 7]));
        u.x = pack_bf16x2(__uint_as_float(bb[q * 8 + 0]), __uint_as_float(bb[q * 8 + 1]));
        u.y = pack_bf16x2(__uint_as_float(bb[q * 8 + 2]), __uint_as_float(bb[q * 8 + 3]));
        u.z = pack_bf16x2(__uint_as_float(bb[q * 8 + 4]), __uint_as_float(bb[q * 8 + 5]));
        u.w = pack_bf16x2(__uint_as_float(bb[q * 8 + 6]), __uint_as_float(bb[q * 8 + 7]));
        st_swizzled_16B(sdk, r, half * 32 + q * 8, wv);
        st_swizzled_16B(sdv, r, half * 32 + q * 8, u);
      }
      fence_proxy_async_smem();
      named_barrier_sync(2, 256);
      if (issuer) {
        tma_store_2d(&tm_dqkv, sdk, C + w.h * HD, w.b * seq + w.jt * BKV);
        tma_store_2d(&tm_dqkv, sdv, 2 * C + w.h * HD, w.b * seq + w.jt * BKV);
        tma_store_commit();
      }
      if (!more) break;
      w = wn;
      stat0 = stat0n;
    }
    if (issuer) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");    // staging must outlive the last reduce
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

static int check_attn_args(int batch, int seq, int heads) {
  B200_REQUIRE(batch > 0 && heads > 0 && seq > 0 && seq % 128 == 0, "attention: seq=%d must be a positive multiple of 128", seq);
  B200_REQUIRE(batch <= 65535 && heads <= 65535, "attention: batch/heads too large for the grid");
  B200_REQUIRE((long long)batch * seq * heads < (1ll << 28), "attention: batch*seq*heads=%lld exceeds the 2^28 row limit",
               (long long)batch * seq * heads);
  return 0;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_attention_fwd(const void* qkv, void* out, float* lse, int batch, int seq, int heads,
                                  void* stream) {
  int rc = check_attn_args(batch, seq, heads);
  if (rc) return rc;
  const int C = heads * HD;
  CUtensorMap tm;
  rc = make_tmap_2d(&tm, qkv, 2, (uint64_t)3 * C, (uint64_t)batch * seq, (uint64_t)3 * C * 2, 64, 128, 128);
  if (rc) return rc;
  static unsigned long long attr = 0;
  if (first_use_on_this_device(attr)) {
    B200_CUDA(cudaFuncSetAttribute(attention_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnFwdSmem::BYTES));
  }
  const float scale_log2 = 1.4426950408889634f / 8.0f;   // log2(e) / sqrt(64)
  CUtensorMap tm_out;
  rc = make_tmap_2d(&tm_out, out, 2, (uint64_t)C, (uint64_t)batch * seq, (uint64_t)C * 2, 64, 128, 128);
  if (rc) return rc;
  const int n_items = (seq / BQ) * heads * batch;
  const int grid = n_items < 2 * num_sms() ? n_items : 2 * num_sms();      // persistent, two CTAs per SM
  B200_CUDA(launch_k(attention_fwd_kernel, dim3(grid), dim3(ATT_FWD_THREADS), AttnFwdSmem::BYTES, static_cast<cudaStream_t>(stream),
      tm, tm_out, lse, seq, heads, batch, scale_log2));
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" long long b200_attention_bwd_scratch_bytes(int batch, int seq, int heads) {
  // delta fp32 [batch][heads][seq] + dq accumulator fp32 [batch][seq][heads][64]
  return (long long)batch * seq * heads * (HD + 1) * (long long)sizeof(float);
}

extern "C" int b200_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                                  float* scratch, int batch, int seq, int heads, void* stream) {
  int rc = check_attn_args(batch, seq, heads);
  if (rc) return rc;
  B200_REQUIRE(scratch != nullptr && (reinterpret_cast<uintptr_t>(scratch) & 15) == 0, "attention_bwd: scratch must be a 16-byte aligned buffer of b200_attention_bwd_scratch_bytes()");
  const int C = heads * HD;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int64_t rows = (int64_t)batch * seq * heads;
  float* dq_accum = scratch;                 // [batch*seq*heads][64], first so that it stays 16-byte aligned
  float* delta = scratch + rows * HD;        // [batch][heads][seq]
  CUtensorMap tm_qkv, tm_do;
  rc = make_tmap_2d(&tm_qkv, qkv, 2, (uint64_t)3 * C, (uint64_t)batch * seq, (uint64_t)3 * C * 2, 64, 128, 128);
  if (rc) return rc;
  rc = make_tmap_2d(&tm_do, dout, 2, (uint64_t)C, (uint64_t)batch * seq, (uint64_t)C * 2, 64, 128, 128);
  if (rc) return rc;
  CUtensorMap tm_dqkv;
  rc = make_tmap_2d(&tm_dqkv, dqkv, 2, (uint64_t)3 * C, (uint64_t)batch * seq, (uint64_t)3 * C * 2, 64, 128, 128);
  if (rc) return rc;
  static unsigned long long attr = 0;
  if (first_use_on_this_device(attr)) {
    B200_CUDA(cudaFuncSetAttribute(attention_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnBwdSmem::BYTES));
  }
  int pblocks = (int)((rows * 8 + 255) / 256);
  if (pblocks > num_sms() * 8) pblocks = num_sms() * 8;
  B200_CUDA(launch_k(attention_bwd_prep_kernel, dim3(pblocks), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(out),
                                                    static_cast<const __nv_bfloat16*>(dout), delta, dq_accum, batch, seq,
                                                    heads));
  const float scale = 0.125f, scale_log2 = 1.4426950408889634f / 8.0f;
  const int n_items = (seq / BKV) * heads * batch;
  const int grid = n_items < num_sms() ? n_items : num_sms();       // persistent, one CTA per SM
  B200_CUDA(launch_k(attention_bwd_kernel, dim3(grid), dim3(ATT_BWD_THREADS), AttnBwdSmem::BYTES, s, tm_qkv, tm_do, tm_dqkv,
                     lse, static_cast<const float*>(delta), dq_accum, seq, heads, batch, scale, scale_log2));
  int64_t total = (int64_t)batch * seq * C / 8;
  int qblocks = (int)((total + 255) / 256);
  if (qblocks > num_sms() * 8) qblocks = num_sms() * 8;
  B200_CUDA(launch_k(attention_bwd_post_kernel, dim3(qblocks), dim3(256), 0, s, static_cast<const float*>(dq_accum),
                     static_cast<__nv_bfloat16*>(dqkv), batch, seq, heads));
  B200_CUDA(cudaGetLastError());
  return 0;
}

// =============================================================================================
// Allocation-free variant of the attention backward (no counterpart in the reference).
//
// Who needs it: a host that cannot hand us scratch memory. Examples are a plain C/C++ program
// or plugin that links this library without any device allocator of its own, and a runtime that
// lays out all of its device memory once at start-up and then forbids cudaMalloc, because the
// call synchronises the whole device and can fail late when the memory budget is already
// spent. For them the scratch (fp32 dQ accumulator + the row dot products delta) lives in the
// fixed array below instead of in a buffer the caller passes in.
//
// Costs and rules that follow from that choice:
//   * every GPU that loads the library carries the array (about 4.1 MiB) until the process exits;
//   * the problem must fit: batch * seq * heads <= B200_ATTN_BWD_NOALLOC_MAX_ROWS, larger
//     problems are refused with an error, never truncated;
//   * all calls on one GPU use the same array. It still holds the previous call's numbers; the
//     part that has to start from zero (the dQ accumulator) is cleared by the first kernel of
//     every call, the rest is fully overwritten before it is read. Two calls that run on the GPU
//     at the same time (different streams) would trample each other: issue them on one stream or
//     order them with events;
//   * the first call on each device looks the array's device address up (cudaGetSymbolAddress);
//     make that first call before starting any stream capture. Later calls only launch kernels.
// It calls no cudaMalloc/cudaFree and no library routine that allocates.
// =============================================================================================
static __device__ __align__(256) float g_attn_bwd_scratch[B200_ATTN_BWD_NOALLOC_MAX_ROWS * (64 + 1)];

extern "C" int b200_attention_bwd_noalloc(const void* qkv, const void* out, const void* dout, const float* lse,
                                          void* dqkv, int batch, int seq, int heads, void* stream) {
  int rc = check_attn_args(batch, seq, heads);
  if (rc) return rc;
  const long long rows = (long long)batch * seq * heads;
  B200_REQUIRE(rows <= B200_ATTN_BWD_NOALLOC_MAX_ROWS,
               "attention_bwd_noalloc: batch*seq*heads = %lld exceeds the built-in scratch (%d rows); "
               "use b200_attention_bwd with caller-provided scratch",
               rows, B200_ATTN_BWD_NOALLOC_MAX_ROWS);
  // In host code `g_attn_bwd_scratch` is only a handle for the symbol, not a device address:
  // resolve it once per device, with that device current, and keep the result.
  constexpr int kMaxDevices = 64;
  static float* dev_ptr[kMaxDevices] = {nullptr};
  int dev = -1;
  B200_CUDA(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < kMaxDevices, "attention_bwd_noalloc: unsupported device ordinal %d", dev);
  float* scratch = dev_ptr[dev];
  if (scratch == nullptr) {
    void* p = nullptr;
    B200_CUDA(cudaGetSymbolAddress(&p, g_attn_bwd_scratch));
    B200_REQUIRE(p != nullptr, "attention_bwd_noalloc: could not resolve the scratch symbol on device %d", dev);
    scratch = static_cast<float*>(p);
    dev_ptr[dev] = scratch;
  }
  return b200_attention_bwd(qkv, out, dout, lse, dqkv, scratch, batch, seq, heads, stream);
}
