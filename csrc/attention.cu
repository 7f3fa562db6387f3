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
// Warp roles (192 threads): warp 0 TMA producer, warp 1 MMA issuer (+ TMEM owner),
// warps 2..5 one thread per query row (softmax / gradient math / epilogue).
#include "../include/b200dist.h"
#include "common.cuh"

namespace b200 {

constexpr int HD = 64;        // head dim
constexpr int BQ = 128;       // query rows per tile
constexpr int BKV = 128;      // key/value rows per tile
constexpr int TILE_BYTES = 128 * 64 * 2;   // one [128][64] bf16 swizzled tile = 16 KB
constexpr int ATT_THREADS = 192;

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
  static constexpr int BYTES = BAR + 128 + 1024;
};

__global__ void __launch_bounds__(ATT_THREADS, 2)
attention_fwd_kernel(const __grid_constant__ CUtensorMap tm_qkv, __nv_bfloat16* __restrict__ out,
                     float* __restrict__ lse, int seq, int heads, float scale_log2) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + AttnFwdSmem::BAR);
  uint64_t* q_full = bars + 0;
  uint64_t* kv_full = bars + 1;   // [2]
  uint64_t* kv_empty = bars + 3;  // [2]
  uint64_t* s_full = bars + 5;
  uint64_t* p_full = bars + 6;
  uint64_t* o_full = bars + 7;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = seq / BQ;
  const int qt = nq - 1 - static_cast<int>(blockIdx.x);   // heavy (late) tiles first
  const int h = blockIdx.y, b = blockIdx.z;
  const int C = heads * HD;
  const int row0 = b * seq;
  const int nkv = qt + 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_qkv);
    mbar_init(q_full, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(&kv_full[s], 1); mbar_init(&kv_empty[s], 1); }
    mbar_init(s_full, 1);
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
    if (lane == 0) {
      mbar_expect_tx(q_full, TILE_BYTES);
      tma_load_2d(smem + AttnFwdSmem::Q, &tm_qkv, q_full, h * HD, row0 + qt * BQ);
      for (int j = 0; j < nkv; ++j) {
        const int s = j & 1;
        mbar_wait(&kv_empty[s], ((j >> 1) & 1) ^ 1);
        mbar_expect_tx(&kv_full[s], 2 * TILE_BYTES);
        tma_load_2d(smem + AttnFwdSmem::K0 + s * TILE_BYTES, &tm_qkv, &kv_full[s], C + h * HD, row0 + j * BKV);
        tma_load_2d(smem + AttnFwdSmem::V0 + s * TILE_BYTES, &tm_qkv, &kv_full[s], 2 * C + h * HD, row0 + j * BKV);
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, 0, 0);   // S = Q K^T
    constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, 1);    // O = P V (V is MN-major)
    const uint32_t sq = smem_u32(smem + AttnFwdSmem::Q);
    const uint32_t sp = smem_u32(smem + AttnFwdSmem::P);
    mbar_wait(q_full, 0);
    for (int j = 0; j < nkv; ++j) {
      const int s = j & 1;
      const uint32_t sk = smem_u32(smem + AttnFwdSmem::K0 + s * TILE_BYTES);
      const uint32_t sv = smem_u32(smem + AttnFwdSmem::V0 + s * TILE_BYTES);
      mbar_wait(&kv_full[s], (j >> 1) & 1);
      tc_fence_after();
      if (lane == 0) {
#pragma unroll
        for (int k = 0; k < HD / 16; ++k) umma_bf16(tmem_s, kdesc(sq + k * 32), kdesc(sk + k * 32), idesc_s, k ? 1u : 0u);
        umma_commit(s_full);
      }
      __syncwarp();
      mbar_wait(p_full, j & 1);
      tc_fence_after();
      if (lane == 0) {
#pragma unroll
        for (int k = 0; k < BKV / 16; ++k)
          umma_bf16(tmem_o, kdesc(sp + (k >> 2) * TILE_BYTES + (k & 3) * 32), mndesc(sv + k * 2048, TILE_BYTES),
                    idesc_o, k ? 1u : 0u);
        umma_commit(o_full);
        umma_commit(&kv_empty[s]);
      }
      __syncwarp();
    }
  } else {
    const int quad = warp & 3;
    const int r = quad * 32 + lane;                 // row within the q tile
    const uint32_t lane_off = static_cast<uint32_t>(quad * 32) << 16;
    float o_acc[HD];
#pragma unroll
    for (int i = 0; i < HD; ++i) o_acc[i] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;
    uint8_t* sp = smem + AttnFwdSmem::P;
    for (int j = 0; j < nkv; ++j) {
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      const bool diag = (j == qt);
      // pass 1: row maximum
      float m_tile = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_s + lane_off + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float x = __uint_as_float(v[i]);
          if (diag && (c * 32 + i) > r) x = -INFINITY;
          m_tile = fmaxf(m_tile, x);
        }
      }
      const float m_new = fmaxf(m_run, m_tile * scale_log2);
      const float alpha = fast_exp2(m_run - m_new);     // m_run = -inf on the first tile -> 0
      float l_tile = 0.f;
      // pass 2: probabilities -> bf16 -> swizzled shared memory
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_s + lane_off + c * 32, v);
        tmem_ld_wait();
        float p[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float x = __uint_as_float(v[i]) * scale_log2 - m_new;
          float e = fast_exp2(x);
          if (diag && (c * 32 + i) > r) e = 0.f;
          p[i] = e;
          l_tile += e;
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 w;
          w.x = pack_bf16x2(p[g * 8 + 0], p[g * 8 + 1]);
          w.y = pack_bf16x2(p[g * 8 + 2], p[g * 8 + 3]);
          w.z = pack_bf16x2(p[g * 8 + 4], p[g * 8 + 5]);
          w.w = pack_bf16x2(p[g * 8 + 6], p[g * 8 + 7]);
          st_swizzled_16B(sp, r, c * 32 + g * 8, w);
        }
      }
      l_run = l_run * alpha + l_tile;
      m_run = m_new;
#pragma unroll
      for (int i = 0; i < HD; ++i) o_acc[i] *= alpha;
      tc_fence_before();
      fence_proxy_async_smem();     // generic-proxy smem writes -> visible to the tensor core
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
      // O += P V
      mbar_wait(o_full, j & 1);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_o + lane_off + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o_acc[c * 32 + i] += __uint_as_float(v[i]);
      }
    }
    const float inv = 1.0f / l_run;
    const size_t grow = static_cast<size_t>(row0 + qt * BQ + r);
    __nv_bfloat16* op = out + (grow * heads + h) * HD;
#pragma unroll
    for (int g = 0; g < HD / 8; ++g) {
      uint4 w;
      w.x = pack_bf16x2(o_acc[g * 8 + 0] * inv, o_acc[g * 8 + 1] * inv);
      w.y = pack_bf16x2(o_acc[g * 8 + 2] * inv, o_acc[g * 8 + 3] * inv);
      w.z = pack_bf16x2(o_acc[g * 8 + 4] * inv, o_acc[g * 8 + 5] * inv);
      w.w = pack_bf16x2(o_acc[g * 8 + 6] * inv, o_acc[g * 8 + 7] * inv);
      reinterpret_cast<uint4*>(op)[g] = w;
    }
    lse[(static_cast<size_t>(b) * heads + h) * seq + qt * BQ + r] = m_run + __log2f(l_run);
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
// delta[b,h,t] = sum_d dO*O ; dq_accum zeroed. One warp per (b,t,h) row of 64.
__global__ void attention_bwd_prep_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ d_o,
                                          float* __restrict__ delta, float* __restrict__ dq_accum, int batch, int seq,
                                          int heads) {
  pdl_sync();
  const int lane = threadIdx.x & 31;
  const int64_t rows = static_cast<int64_t>(batch) * seq * heads;
  for (int64_t row = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows;
       row += static_cast<int64_t>(gridDim.x) * (blockDim.x >> 5)) {
    const __nv_bfloat162 a = reinterpret_cast<const __nv_bfloat162*>(o + row * HD)[lane];
    const __nv_bfloat162 g = reinterpret_cast<const __nv_bfloat162*>(d_o + row * HD)[lane];
    float s = __bfloat162float(a.x) * __bfloat162float(g.x) + __bfloat162float(a.y) * __bfloat162float(g.y);
    s = warp_sum(s);
    reinterpret_cast<float2*>(dq_accum + row * HD)[lane] = make_float2(0.f, 0.f);
    if (lane == 0) {
      const int64_t bt = row / heads;
      const int hh = static_cast<int>(row - bt * heads);
      const int64_t bb = bt / seq;
      const int t = static_cast<int>(bt - bb * seq);
      delta[(bb * heads + hh) * seq + t] = s;
    }
  }
}

// dq bf16 <- dq_accum fp32, written into dqkv[..., 0, h, :]
__global__ void attention_bwd_post_kernel(const float* __restrict__ dq_accum, __nv_bfloat16* __restrict__ dqkv,
                                          int64_t rows_bt, int heads) {
  pdl_sync();
  const int64_t total = rows_bt * heads * (HD / 2);
  const int C = heads * HD;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t bt = i / (C / 2);
    const int c2 = static_cast<int>(i - bt * (C / 2));
    const float2 v = reinterpret_cast<const float2*>(dq_accum)[i];
    reinterpret_cast<__nv_bfloat162*>(dqkv + bt * 3 * C)[c2] = __floats2bfloat162_rn(v.x, v.y);
  }
}

struct AttnBwdSmem {
  static constexpr int K = 0;
  static constexpr int V = K + TILE_BYTES;
  static constexpr int Q0 = V + TILE_BYTES;              // 2 stages
  static constexpr int DO0 = Q0 + 2 * TILE_BYTES;        // 2 stages
  static constexpr int P = DO0 + 2 * TILE_BYTES;         // 2 tiles
  static constexpr int DS = P + 2 * TILE_BYTES;          // 2 tiles
  static constexpr int BAR = DS + 2 * TILE_BYTES;
  static constexpr int BYTES = BAR + 128 + 1024;
};

__global__ void __launch_bounds__(ATT_THREADS, 1)
attention_bwd_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_do,
                     const float* __restrict__ lse, const float* __restrict__ delta, float* __restrict__ dq_accum,
                     __nv_bfloat16* __restrict__ dqkv, int seq, int heads, float scale, float scale_log2) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + AttnBwdSmem::BAR);
  uint64_t* kv_full = bars + 0;
  uint64_t* qdo_full = bars + 1;   // [2]
  uint64_t* qdo_empty = bars + 3;  // [2]
  uint64_t* sdp_full = bars + 5;
  uint64_t* pds_full = bars + 6;
  uint64_t* dq_full = bars + 7;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = seq / BQ;
  const int jt = blockIdx.x;            // kv tile owned by this CTA (tile 0 has the most work)
  const int h = blockIdx.y, b = blockIdx.z;
  const int C = heads * HD;
  const int row0 = b * seq;
  const int niter = nq - jt;            // q tiles jt .. nq-1

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_qkv);
    tma_prefetch_desc(&tm_do);
    mbar_init(kv_full, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(&qdo_full[s], 1); mbar_init(&qdo_empty[s], 1); }
    mbar_init(sdp_full, 1);
    mbar_init(pds_full, 4);
    mbar_init(dq_full, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_sync();
  const uint32_t tm_s = tmem_base, tm_dp = tmem_base + 128, tm_dv = tmem_base + 256, tm_dk = tmem_base + 320,
                 tm_dq = tmem_base + 384;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(kv_full, 2 * TILE_BYTES);
      tma_load_2d(smem + AttnBwdSmem::K, &tm_qkv, kv_full, C + h * HD, row0 + jt * BKV);
      tma_load_2d(smem + AttnBwdSmem::V, &tm_qkv, kv_full, 2 * C + h * HD, row0 + jt * BKV);
      for (int it = 0; it < niter; ++it) {
        const int s = it & 1;
        const int qi = jt + it;
        mbar_wait(&qdo_empty[s], ((it >> 1) & 1) ^ 1);
        mbar_expect_tx(&qdo_full[s], 2 * TILE_BYTES);
        tma_load_2d(smem + AttnBwdSmem::Q0 + s * TILE_BYTES, &tm_qkv, &qdo_full[s], h * HD, row0 + qi * BQ);
        tma_load_2d(smem + AttnBwdSmem::DO0 + s * TILE_BYTES, &tm_do, &qdo_full[s], h * HD, row0 + qi * BQ);
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t id_s = umma_idesc_bf16(128, 128, 0, 0);    // S = Q K^T, dP = dO V^T
    constexpr uint32_t id_t = umma_idesc_bf16(128, 64, 1, 1);     // dV = P^T dO, dK = dS^T Q
    constexpr uint32_t id_q = umma_idesc_bf16(128, 64, 0, 1);     // dQ = dS K
    const uint32_t sk = smem_u32(smem + AttnBwdSmem::K), sv = smem_u32(smem + AttnBwdSmem::V);
    const uint32_t sp = smem_u32(smem + AttnBwdSmem::P), sds = smem_u32(smem + AttnBwdSmem::DS);
    mbar_wait(kv_full, 0);
    for (int it = 0; it < niter; ++it) {
      const int s = it & 1;
      const uint32_t sq = smem_u32(smem + AttnBwdSmem::Q0 + s * TILE_BYTES);
      const uint32_t sdo = smem_u32(smem + AttnBwdSmem::DO0 + s * TILE_BYTES);
      mbar_wait(&qdo_full[s], (it >> 1) & 1);
      tc_fence_after();
      if (lane == 0) {
#pragma unroll
        for (int k = 0; k < HD / 16; ++k) umma_bf16(tm_s, kdesc(sq + k * 32), kdesc(sk + k * 32), id_s, k ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k) umma_bf16(tm_dp, kdesc(sdo + k * 32), kdesc(sv + k * 32), id_s, k ? 1u : 0u);
        umma_commit(sdp_full);
      }
      __syncwarp();
      mbar_wait(pds_full, it & 1);
      tc_fence_after();
      if (lane == 0) {
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
          umma_bf16(tm_dq, kdesc(sds + (k >> 2) * TILE_BYTES + (k & 3) * 32), mndesc(sk + k * 2048, TILE_BYTES), id_q,
                    k ? 1u : 0u);
        umma_commit(dq_full);
        umma_commit(&qdo_empty[s]);
      }
      __syncwarp();
    }
  } else {
    const int quad = warp & 3;
    const int r = quad * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(quad * 32) << 16;
    uint8_t* sp = smem + AttnBwdSmem::P;
    uint8_t* sds = smem + AttnBwdSmem::DS;
    for (int it = 0; it < niter; ++it) {
      const int qi = jt + it;
      const bool diag = (qi == jt);
      const size_t stat = (static_cast<size_t>(b) * heads + h) * seq + qi * BQ + r;
      const float lse_r = lse[stat];
      const float dl_r = delta[stat];
      mbar_wait(sdp_full, it & 1);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t sv_[32], dv_[32];
        tmem_ld_32x32b_x32(tm_s + lane_off + c * 32, sv_);
        tmem_ld_32x32b_x32(tm_dp + lane_off + c * 32, dv_);
        tmem_ld_wait();
        float p[32], ds[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float e = fast_exp2(__uint_as_float(sv_[i]) * scale_log2 - lse_r);
          if (diag && (c * 32 + i) > r) e = 0.f;
          p[i] = e;
          ds[i] = e * (__uint_as_float(dv_[i]) - dl_r) * scale;
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 w, u;
          w.x = pack_bf16x2(p[g * 8 + 0], p[g * 8 + 1]);   w.y = pack_bf16x2(p[g * 8 + 2], p[g * 8 + 3]);
          w.z = pack_bf16x2(p[g * 8 + 4], p[g * 8 + 5]);   w.w = pack_bf16x2(p[g * 8 + 6], p[g * 8 + 7]);
          u.x = pack_bf16x2(ds[g * 8 + 0], ds[g * 8 + 1]); u.y = pack_bf16x2(ds[g * 8 + 2], ds[g * 8 + 3]);
          u.z = pack_bf16x2(ds[g * 8 + 4], ds[g * 8 + 5]); u.w = pack_bf16x2(ds[g * 8 + 6], ds[g * 8 + 7]);
          st_swizzled_16B(sp, r, c * 32 + g * 8, w);
          st_swizzled_16B(sds, r, c * 32 + g * 8, u);
        }
      }
      tc_fence_before();
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(pds_full);
      // dQ tile -> fp32 accumulation buffer (other kv tiles add into the same rows)
      mbar_wait(dq_full, it & 1);
      tc_fence_after();
      float* dqp = dq_accum + ((static_cast<size_t>(row0 + qi * BQ + r)) * heads + h) * HD;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tm_dq + lane_off + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; i += 4)
          asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dqp + c * 32 + i), "f"(__uint_as_float(v[i])),
                       "f"(__uint_as_float(v[i + 1])), "f"(__uint_as_float(v[i + 2])), "f"(__uint_as_float(v[i + 3]))
                       : "memory");
      }
    }
    // dK, dV of this kv tile (the last dq_full commit covers every MMA issued before it)
    const size_t grow = static_cast<size_t>(row0 + jt * BKV + r);
    __nv_bfloat16* dkp = dqkv + grow * 3 * C + C + h * HD;
    __nv_bfloat16* dvp = dqkv + grow * 3 * C + 2 * C + h * HD;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      uint32_t a[32], bb[32];
      tmem_ld_32x32b_x32(tm_dk + lane_off + c * 32, a);
      tmem_ld_32x32b_x32(tm_dv + lane_off + c * 32, bb);
      tmem_ld_wait();
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        uint4 w, u;
        w.x = pack_bf16x2(__uint_as_float(a[g * 8 + 0]), __uint_as_float(a[g * 8 + 1]));
        w.y = pack_bf16x2(__uint_as_float(a[g * 8 + 2]), __uint_as_float(a[g * 8 + 3]));
        w.z = pack_bf16x2(__uint_as_float(a[g * 8 + 4]), __uint_as_float(a[g * 8 + 5]));
        w.w = pack_bf16x2(__uint_as_float(a[g * 8 + 6]), __uint_as_float(a[g * 8 + 7]));
        u.x = pack_bf16x2(__uint_as_float(bb[g * 8 + 0]), __uint_as_float(bb[g * 8 + 1]));
        u.y = pack_bf16x2(__uint_as_float(bb[g * 8 + 2]), __uint_as_float(bb[g * 8 + 3]));
        u.z = pack_bf16x2(__uint_as_float(bb[g * 8 + 4]), __uint_as_float(bb[g * 8 + 5]));
        u.w = pack_bf16x2(__uint_as_float(bb[g * 8 + 6]), __uint_as_float(bb[g * 8 + 7]));
        reinterpret_cast<uint4*>(dkp)[c * 4 + g] = w;
        reinterpret_cast<uint4*>(dvp)[c * 4 + g] = u;
      }
    }
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
  static bool attr = false;
  if (!attr) {
    B200_CUDA(cudaFuncSetAttribute(attention_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnFwdSmem::BYTES));
    attr = true;
  }
  const float scale_log2 = 1.4426950408889634f / 8.0f;   // log2(e) / sqrt(64)
  dim3 grid(seq / BQ, heads, batch);
  B200_CUDA(launch_k(attention_fwd_kernel, dim3(grid), dim3(ATT_THREADS), AttnFwdSmem::BYTES, static_cast<cudaStream_t>(stream), 
      tm, static_cast<__nv_bfloat16*>(out), lse, seq, heads, scale_log2));
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
  static bool attr = false;
  if (!attr) {
    B200_CUDA(cudaFuncSetAttribute(attention_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnBwdSmem::BYTES));
    attr = true;
  }
  int pblocks = (int)((rows + 7) / 8);
  if (pblocks > num_sms() * 16) pblocks = num_sms() * 16;
  B200_CUDA(launch_k(attention_bwd_prep_kernel, dim3(pblocks), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(out),
                                                    static_cast<const __nv_bfloat16*>(dout), delta, dq_accum, batch, seq,
                                                    heads));
  const float scale = 0.125f, scale_log2 = 1.4426950408889634f / 8.0f;
  dim3 grid(seq / BKV, heads, batch);
  B200_CUDA(launch_k(attention_bwd_kernel, dim3(grid), dim3(ATT_THREADS), AttnBwdSmem::BYTES, s, tm_qkv, tm_do, lse, delta, dq_accum,
                                                                    static_cast<__nv_bfloat16*>(dqkv), seq, heads, scale,
                                                                    scale_log2));
  int64_t total = (int64_t)batch * seq * C / 2;
  int qblocks = (int)((total + 255) / 256);
  if (qblocks > num_sms() * 16) qblocks = num_sms() * 16;
  B200_CUDA(launch_k(attention_bwd_post_kernel, dim3(qblocks), dim3(256), 0, s, dq_accum, static_cast<__nv_bfloat16*>(dqkv), (int64_t)batch * seq, heads));
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
