// Persistent, warp-specialised bf16 GEMM for sm_100a.
//
//   warp 0      TMA producer   cp.async.bulk.tensor tiles -> 128B-swizzled shared memory ring
//   warp 1      MMA issuer     one lane issues tcgen05.mma (M=128, N=BN, K=16), accumulators in TMEM
//   warps 2..5  epilogue       tcgen05.ld -> registers -> bias/activation/dropout/residual/cast -> global
//
// TMEM holds two accumulator stages so the epilogue of tile i overlaps the main loop of tile
// i+1. The B operand can come from several equally sized shards (peer GPUs' memory) selected
// per K block or per N tile: that is the FSDP all-gather fused into the consuming GEMM.
//
// What the reference does here: flax nn.Dense lowered by XLA to a library GEMM followed by
// separate elementwise fusions (data_paral.py:86-100); FSDP gathers the weight with a
// stand-alone all_gather before the matmul (param_sharding.py:140,153-158).
#include "../include/b200dist.h"
#include "common.cuh"

namespace b200 {

constexpr int BM = 128;
constexpr int BK = 64;       // 64 bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int GEMM_THREADS = 320;   // TMA warp + MMA warp + 8 epilogue warps

struct TmapSet {
  CUtensorMap m[8];
};

struct GemmDev {
  int M, N, K;
  int shard_mode, nsrc, shard_len, rot;
  // epilogue
  void* d;
  int ldd, d_fp32, accumulate;
  float alpha;
  const __nv_bfloat16* bias;
  __nv_bfloat16* pre;
  int ldpre;
  int act;
  const __nv_bfloat16* resid;
  int ldres;
  int bwd;
  const __nv_bfloat16* aux;
  int ldaux;
  const uint64_t* drop_key;
  uint64_t drop_salt;
  int drop_nsalt, drop_rows_per_salt;
  uint64_t drop_salts[16];
  uint32_t drop_thresh16;
  float drop_scale;
  int vec_d, vec_pre, vec_aux, vec_res;  // 16-byte vector access allowed
  int epi_kind;                          // EPI_*: compile-time specialised epilogue, or EPI_GENERIC
  int k_splits, atomic_out;              // split-K: tile index also selects a K range; partials are red.add-ed
  float* dbias;                          // column sums of the stored output, added atomically (or null)
  const int64_t* xent_labels;            // EPI_XENT: integer labels, metrics {loss sum, #correct}, gradient scale
  float* xent_metrics;
  float xent_scale;
};

// Specialised epilogues (fully vectorised, no run-time feature tests in the inner loop). Anything
// else - dropout, SiLU, odd alignments, partial 4-column groups - takes EPI_GENERIC.
enum { EPI_GENERIC = 0, EPI_PLAIN = 1, EPI_PLAIN_F32 = 2, EPI_BIAS = 3, EPI_BIAS_RESID = 4, EPI_BIAS_GELU_PRE = 5,
       EPI_BWD_GELU = 6, EPI_XENT = 7 };

template <int BN>
struct GemmCfg {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (BN == 256) ? 4 : (BN == 128) ? 5 : 6;
  static constexpr int EPI_STAGE_BYTES = 8 * 32 * 32 * 4 + 8 * 16 * 8;   // per-warp 32x32 fp32 transpose buffers + dropout seed tables
  static constexpr uint32_t TMEM_COLS = (2 * BN <= 32) ? 32 : (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128 : (2 * BN <= 256) ? 256 : 512;
  static constexpr int BAR_BYTES = 256;   // (2*STAGES + 4) mbarriers + the TMEM base address, 16-byte multiple
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + BAR_BYTES + EPI_STAGE_BYTES + 1024;  // + alignment slack
};

struct TileCoord {
  int m0, n0, n_limit, b_n, b_src;
};

__device__ __forceinline__ TileCoord tile_coord(const GemmDev& p, int tile, int m_tiles, int BN) {
  TileCoord t;
  int mt = tile % m_tiles;
  int nt = tile / m_tiles;
  t.m0 = mt * BM;
  if (p.shard_mode == 2) {
    int tps = (p.shard_len + BN - 1) / BN;
    int n_tiles = tps * p.nsrc;
    nt = (nt + p.rot * tps) % n_tiles;
    int src = nt / tps, loc = nt % tps;
    t.b_src = src;
    t.b_n = loc * BN;
    t.n0 = src * p.shard_len + loc * BN;
    int lim = (src + 1) * p.shard_len;
    t.n_limit = lim < p.N ? lim : p.N;
  } else {
    t.b_src = 0;
    t.b_n = nt * BN;
    t.n0 = nt * BN;
    t.n_limit = p.N;
  }
  return t;
}

__device__ __forceinline__ float bf16_bits_to_f(uint32_t v) { return __uint_as_float(v << 16); }

// Explicit shared-space accesses for the epilogue staging buffer: through a plain pointer the
// compiler emits generic LD.E/ST.E (64-bit addresses, generic-path latency) because it cannot
// prove the address space.
__device__ __forceinline__ void sts_f32x4(uint32_t addr, float4 v) {
  asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 lds_f32x4(uint32_t addr) {
  float4 v;
  // volatile keeps it ordered after the staging stores / __syncwarp; no "memory" clobber, so kernel
  // parameters are not re-read from the constant bank after every load
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}

template <int BN, int KIND, bool DBIAS>
__device__ __forceinline__ void epilogue_tile_fast(const GemmDev& p, int row_base, int col_base, int n_limit, uint32_t taddr,
                                                   float* stage, int lane) {
  constexpr int CHUNKS = BN / 64;
  const uint32_t stage_s = smem_u32(stage);
  constexpr bool HAS_BIAS = KIND == EPI_BIAS || KIND == EPI_BIAS_RESID || KIND == EPI_BIAS_GELU_PRE;
  const int lrow = lane >> 3, lchunk = lane & 7;
  if (row_base >= p.M || col_base >= n_limit) return;   // warp-uniform
  const int rows_left = p.M - row_base;
  const float alpha = p.alpha;
  // kinds that read global memory in the epilogue keep the loads of the *next* 32-column chunk in
  // flight while the current chunk is processed (and issue them as a batch: the output stores
  // could alias them as far as the compiler knows, so it would not hoist them itself)
  constexpr bool LOADS8 = KIND == EPI_BWD_GELU || KIND == EPI_BIAS_RESID;
  const __nv_bfloat16* lsrc = KIND == EPI_BWD_GELU ? p.aux : p.resid;
  const int lds = KIND == EPI_BWD_GELU ? p.ldaux : p.ldres;
  uint2 nxt8[LOADS8 ? 8 : 1];
  auto issue_loads = [&](int cc) {
    if (!LOADS8) return;
    const int colx = col_base + cc * 32 + lchunk * 4;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      nxt8[LOADS8 ? i : 0] = make_uint2(0u, 0u);
      if (i * 4 + lrow < rows_left && colx < n_limit)
        nxt8[LOADS8 ? i : 0] = *reinterpret_cast<const uint2*>(lsrc + static_cast<size_t>(row_base + lrow + i * 4) * lds + colx);
    }
  };
  issue_loads(0);
  uint32_t r[32];
  tmem_ld_32x32b_x32(taddr, r);
#pragma unroll 1
  for (int c = 0; c < CHUNKS; ++c) {
    const int col0 = col_base + c * 32;
    if (col0 >= n_limit) break;
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float4 v4 = make_float4(__uint_as_float(r[j * 4 + 0]), __uint_as_float(r[j * 4 + 1]),
                              __uint_as_float(r[j * 4 + 2]), __uint_as_float(r[j * 4 + 3]));
      sts_f32x4(stage_s + ((lane * 32 + ((j ^ (lane & 7)) << 2)) << 2), v4);
    }
    __syncwarp();
    if (c + 1 < CHUNKS && col0 + 32 < n_limit) tmem_ld_32x32b_x32(taddr + (c + 1) * 32, r);
    const int col = col0 + lchunk * 4;
    const bool col_ok = col < n_limit;      // n_limit is a multiple of 4 on this path
    float b0 = 0.f, b1 = 0.f, b2 = 0.f, b3 = 0.f;
    if (HAS_BIAS && col_ok) {
      const uint2 bw = *reinterpret_cast<const uint2*>(p.bias + col);
      b0 = bf16_bits_to_f(bw.x & 0xFFFF); b1 = bf16_bits_to_f(bw.x >> 16);
      b2 = bf16_bits_to_f(bw.y & 0xFFFF); b3 = bf16_bits_to_f(bw.y >> 16);
    }
    const size_t row0 = static_cast<size_t>(row_base + lrow);
    float cs0 = 0.f, cs1 = 0.f, cs2 = 0.f, cs3 = 0.f;   // column sums of what this lane stores (dbias)
    uint2 ld8[LOADS8 ? 8 : 1];
    float4 acc8[KIND == EPI_PLAIN_F32 ? 8 : 1];
    if (LOADS8) {
#pragma unroll
      for (int i = 0; i < 8; ++i) ld8[i] = nxt8[i];
      if (c + 1 < CHUNKS && col0 + 32 < n_limit) issue_loads(c + 1);
    }
    if (KIND == EPI_PLAIN_F32 && p.accumulate && !p.atomic_out) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        acc8[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i * 4 + lrow < rows_left && col_ok)
          acc8[i] = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.d) + (row0 + i * 4) * p.ldd + col);
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int lr = i * 4 + lrow;
      if (i * 4 >= rows_left) break;        // warp-uniform
      if (lr < rows_left && col_ok) {
        const size_t row = row0 + i * 4;
        const float4 a4 = lds_f32x4(stage_s + ((lr * 32 + ((lchunk ^ (lr & 7)) << 2)) << 2));
        float v0 = a4.x * alpha + b0, v1 = a4.y * alpha + b1, v2 = a4.z * alpha + b2, v3 = a4.w * alpha + b3;
        if (KIND == EPI_BIAS_GELU_PRE) {
          __nv_bfloat162 t0 = __floats2bfloat162_rn(v0, v1), t1 = __floats2bfloat162_rn(v2, v3);
          *reinterpret_cast<uint2*>(p.pre + row * p.ldpre + col) =
              make_uint2(*reinterpret_cast<uint32_t*>(&t0), *reinterpret_cast<uint32_t*>(&t1));
          v0 = gelu_f(__bfloat162float(t0.x)); v1 = gelu_f(__bfloat162float(t0.y));
          v2 = gelu_f(__bfloat162float(t1.x)); v3 = gelu_f(__bfloat162float(t1.y));
        }
        if (KIND == EPI_BWD_GELU) {
          const uint2 w = ld8[LOADS8 ? i : 0];
          v0 *= gelu_grad_f(bf16_bits_to_f(w.x & 0xFFFF)); v1 *= gelu_grad_f(bf16_bits_to_f(w.x >> 16));
          v2 *= gelu_grad_f(bf16_bits_to_f(w.y & 0xFFFF)); v3 *= gelu_grad_f(bf16_bits_to_f(w.y >> 16));
        }
        if (KIND == EPI_BIAS_RESID) {
          const uint2 w = ld8[LOADS8 ? i : 0];
          v0 += bf16_bits_to_f(w.x & 0xFFFF); v1 += bf16_bits_to_f(w.x >> 16);
          v2 += bf16_bits_to_f(w.y & 0xFFFF); v3 += bf16_bits_to_f(w.y >> 16);
        }
        if (KIND == EPI_PLAIN_F32) {
          float* dp = reinterpret_cast<float*>(p.d) + row * p.ldd + col;
          if (p.atomic_out) {   // split-K partial sums meet in memory
            asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dp), "f"(v0), "f"(v1), "f"(v2), "f"(v3) : "memory");
          } else {
            float4 o = make_float4(v0, v1, v2, v3);
            if (p.accumulate) {
              const float4 c4 = acc8[KIND == EPI_PLAIN_F32 ? i : 0];
              o.x += c4.x; o.y += c4.y; o.z += c4.z; o.w += c4.w;
            }
            *reinterpret_cast<float4*>(dp) = o;
          }
        } else {
          __nv_bfloat162 t0 = __floats2bfloat162_rn(v0, v1), t1 = __floats2bfloat162_rn(v2, v3);
          *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(p.d) + row * p.ldd + col) =
              make_uint2(*reinterpret_cast<uint32_t*>(&t0), *reinterpret_cast<uint32_t*>(&t1));
          if (DBIAS) {
            cs0 += __bfloat162float(t0.x); cs1 += __bfloat162float(t0.y);
            cs2 += __bfloat162float(t1.x); cs3 += __bfloat162float(t1.y);
          }
        }
      }
    }
    if (KIND != EPI_PLAIN_F32 && DBIAS) {
      // lanes l, l^8, l^16, l^24 hold the same four columns for different rows
      cs0 += __shfl_xor_sync(0xffffffffu, cs0, 8);  cs1 += __shfl_xor_sync(0xffffffffu, cs1, 8);
      cs2 += __shfl_xor_sync(0xffffffffu, cs2, 8);  cs3 += __shfl_xor_sync(0xffffffffu, cs3, 8);
      cs0 += __shfl_xor_sync(0xffffffffu, cs0, 16); cs1 += __shfl_xor_sync(0xffffffffu, cs1, 16);
      cs2 += __shfl_xor_sync(0xffffffffu, cs2, 16); cs3 += __shfl_xor_sync(0xffffffffu, cs3, 16);
      if (lrow == 0 && col_ok)
        asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p.dbias + col), "f"(cs0), "f"(cs1), "f"(cs2), "f"(cs3) : "memory");
    }
    __syncwarp();
  }
  tmem_ld_wait();
}

// Classifier head with at most 32 classes: the accumulator row of a sample sits in one thread
// (TMEM lane = row), so softmax, loss, accuracy and dlogits are computed in registers right here and
// the logits never go to memory. Only the warps that hold columns [0, 32) take part.
__device__ __noinline__ void epilogue_tile_xent(const GemmDev& p, int row_base, int col_base, uint32_t taddr, float* stage,
                                                int lane, int label_pre, float bias_pre) {
  if (col_base != 0 || row_base >= p.M) return;   // warp-uniform
  // Deliberately compact code (rolled loops over the classes, row parked in the warp's staging
  // buffer): this runs once per step in a launch-latency-bound chain, where instruction fetch of a
  // fully unrolled 32-class version from a cold cache costs more than the arithmetic.
  uint32_t r[32];
  tmem_ld_32x32b_x32(taddr, r);
  tmem_ld_wait();
  float* mine = stage + lane * 32;                // element j of this lane's row lives at (j ^ lane): conflict-free
#pragma unroll
  for (int j = 0; j < 32; ++j) mine[j ^ lane] = __uint_as_float(r[j]);
  __syncwarp();
  const int row = row_base + lane;
  const bool live = row < p.M;
  const int nc = p.N;
  const int label = live ? label_pre : 0;         // label of this lane's row and bias[lane]: loaded before the
  float m = -INFINITY, l_label = 0.f;             // accumulator wait, i.e. under the main loop
  int amax = 0;
#pragma unroll 1
  for (int j = 0; j < nc; ++j) {
    const float v = mine[j ^ lane] * p.alpha + __shfl_sync(0xffffffffu, bias_pre, j);
    mine[j ^ lane] = v;
    if (v > m) { m = v; amax = j; }               // strict: the first maximum wins
    if (j == label) l_label = v;
  }
  float ssum = 0.f;
#pragma unroll 1
  for (int j = 0; j < nc; ++j) ssum += __expf(mine[j ^ lane] - m);
  const float inv = 1.0f / ssum;
  __nv_bfloat16* drow = reinterpret_cast<__nv_bfloat16*>(p.d) + static_cast<size_t>(row) * p.ldd;
#pragma unroll 1
  for (int j = 0; j < nc; ++j) {
    const __nv_bfloat16 g = __float2bfloat16_rn((__expf(mine[j ^ lane] - m) * inv - (j == label ? 1.0f : 0.f)) * p.xent_scale);
    if (live) drow[j] = g;
    if (p.dbias != nullptr) {
      const float cs = warp_sum(live ? __bfloat162float(g) : 0.f);
      if (lane == 0) atomicAdd(&p.dbias[j], cs);
    }
  }
  const float loss = warp_sum(live ? (m + __logf(ssum) - l_label) : 0.f);
  const float corr = warp_sum((live && amax == label) ? 1.0f : 0.f);
  if (lane == 0 && p.xent_metrics != nullptr) {
    atomicAdd(&p.xent_metrics[0], loss);
    atomicAdd(&p.xent_metrics[1], corr);
  }
}

// Operands of the cross-entropy epilogue that do not depend on the accumulator: fetched before
// the wait on the accumulator so that their (cold-cache) latency hides under the main loop.
__device__ __forceinline__ void xent_prefetch(const GemmDev& p, int row_base, int col_base, int lane, int& label_pre,
                                              float& bias_pre) {
  label_pre = 0;
  bias_pre = 0.f;
  if (p.epi_kind != EPI_XENT || col_base != 0) return;
  const int row = row_base + lane;
  if (row < p.M) label_pre = static_cast<int>(p.xent_labels[row]);
  if (lane < p.N && p.bias != nullptr) bias_pre = __bfloat162float(p.bias[lane]);
}

// Dispatch once per tile on the epilogue kind (uniform across the kernel).
template <int BN>
__device__ __forceinline__ void epilogue_dispatch(const GemmDev& p, int row_base, int col_base, int n_limit, uint32_t taddr,
                                                  float* stage, const uint64_t* seed_tab, uint64_t seed, int lane,
                                                  int label_pre, float bias_pre);

// Epilogue of one warp for one tile: rows [row_base, row_base+32) x columns [col_base, col_base+BN/2)
// of the output, accumulator at TMEM address `taddr` (lane quadrant, stage and column half already
// applied). See the comment at the call site for the data path.
template <int BN>
__device__ __forceinline__ void epilogue_tile(const GemmDev& p, int row_base, int col_base, int n_limit, uint32_t taddr,
                                              float* stage, const uint64_t* seed_tab, uint64_t seed, int lane) {
  constexpr int CHUNKS = BN / 64;     // 32-column chunks per warp
  const int lrow = lane >> 3;         // row within a group of 4 rows
  const int lchunk = lane & 7;        // 16-byte chunk (4 columns) within the 32-column row
  const bool warp_active = row_base < p.M && col_base < n_limit;
  uint32_t r[32];
  if (warp_active) tmem_ld_32x32b_x32(taddr, r);
#pragma unroll 1
  for (int c = 0; c < CHUNKS; ++c) {
    const int col0 = col_base + c * 32;
    if (!warp_active || col0 >= n_limit) break;  // warp-uniform
    tmem_ld_wait();
    // registers (thread = row) -> staging, 16-byte chunks XOR-swizzled by row: conflict-free
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float4 v4 = make_float4(__uint_as_float(r[j * 4 + 0]), __uint_as_float(r[j * 4 + 1]),
                              __uint_as_float(r[j * 4 + 2]), __uint_as_float(r[j * 4 + 3]));
      *reinterpret_cast<float4*>(stage + lane * 32 + ((j ^ (lane & 7)) << 2)) = v4;
    }
    __syncwarp();
    // prefetch the next chunk from TMEM while this one is being processed
    if (c + 1 < CHUNKS && col0 + 32 < n_limit) tmem_ld_32x32b_x32(taddr + (c + 1) * 32, r);
    const int col = col0 + lchunk * 4;
    const bool full4 = col + 4 <= n_limit;
    const bool col_ok = col < n_limit;
    float bias4[4] = {0.f, 0.f, 0.f, 0.f};
    float cs[4] = {0.f, 0.f, 0.f, 0.f};   // column sums of what this lane stores (dbias)
    if (!p.bwd && p.bias != nullptr && col_ok) {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (col + j < n_limit) bias4[j] = __bfloat162float(p.bias[col + j]);
    }
#pragma unroll 2
    for (int i = 0; i < 8; ++i) {
      const int lr = i * 4 + lrow;          // row within the warp's 32 rows
      const int row = row_base + lr;
      if (row_base + i * 4 >= p.M) break;   // warp-uniform: nothing left in this tile
      if (row >= p.M || !col_ok) continue;
      const float4 a4 = *reinterpret_cast<const float4*>(stage + lr * 32 + ((lchunk ^ (lr & 7)) << 2));
      float v[4] = {a4.x * p.alpha, a4.y * p.alpha, a4.z * p.alpha, a4.w * p.alpha};
      if (!p.bwd) {
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] += bias4[j];
        if (p.pre != nullptr) {
          __nv_bfloat16* pp = p.pre + static_cast<size_t>(row) * p.ldpre + col;
          __nv_bfloat162 t0 = __floats2bfloat162_rn(v[0], v[1]), t1 = __floats2bfloat162_rn(v[2], v[3]);
          if (full4 && p.vec_pre) {
            *reinterpret_cast<uint2*>(pp) = make_uint2(*reinterpret_cast<uint32_t*>(&t0), *reinterpret_cast<uint32_t*>(&t1));
          } else {
            __nv_bfloat16 tv[4] = {t0.x, t0.y, t1.x, t1.y};
#pragma unroll
            for (int j = 0; j < 4; ++j)
              if (col + j < n_limit) pp[j] = tv[j];
          }
          // the activation sees the value as it was stored (bf16), like the backward will
          v[0] = __bfloat162float(t0.x); v[1] = __bfloat162float(t0.y);
          v[2] = __bfloat162float(t1.x); v[3] = __bfloat162float(t1.y);
        }
        if (p.act == B200_ACT_SILU) {
#pragma unroll
          for (int j = 0; j < 4; ++j) v[j] = silu_f(v[j]);
        } else if (p.act == B200_ACT_GELU) {
#pragma unroll
          for (int j = 0; j < 4; ++j) v[j] = gelu_f(v[j]);
        }
      } else if (p.aux != nullptr) {
        const __nv_bfloat16* ap = p.aux + static_cast<size_t>(row) * p.ldaux + col;
        float x[4];
        if (full4 && p.vec_aux) {
          const uint2 w = *reinterpret_cast<const uint2*>(ap);
          x[0] = bf16_bits_to_f(w.x & 0xFFFF); x[1] = bf16_bits_to_f(w.x >> 16);
          x[2] = bf16_bits_to_f(w.y & 0xFFFF); x[3] = bf16_bits_to_f(w.y >> 16);
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) x[j] = (col + j < n_limit) ? __bfloat162float(ap[j]) : 0.0f;
        }
        if (p.act == B200_ACT_SILU) {
#pragma unroll
          for (int j = 0; j < 4; ++j) v[j] *= silu_grad_f(x[j]);
        } else if (p.act == B200_ACT_GELU) {
#pragma unroll
          for (int j = 0; j < 4; ++j) v[j] *= gelu_grad_f(x[j]);
        }
      }
      if (p.drop_key != nullptr) {
        int drow = row;
        uint64_t sd = seed;
        if (p.drop_nsalt > 0) {
          int blk = row / p.drop_rows_per_salt;
          drow = row - blk * p.drop_rows_per_salt;
          sd = seed_tab[blk < p.drop_nsalt ? blk : p.drop_nsalt - 1];
        }
        const uint32_t keep = dropout_keep4(sd, static_cast<uint32_t>(drow), static_cast<uint32_t>(col) >> 2,
                                            p.drop_thresh16);
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = ((keep >> j) & 1u) ? v[j] * p.drop_scale : 0.0f;
      }
      if (!p.bwd && p.resid != nullptr) {
        const __nv_bfloat16* rp = p.resid + static_cast<size_t>(row) * p.ldres + col;
        if (full4 && p.vec_res) {
          const uint2 w = *reinterpret_cast<const uint2*>(rp);
          v[0] += bf16_bits_to_f(w.x & 0xFFFF); v[1] += bf16_bits_to_f(w.x >> 16);
          v[2] += bf16_bits_to_f(w.y & 0xFFFF); v[3] += bf16_bits_to_f(w.y >> 16);
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (col + j < n_limit) v[j] += __bfloat162float(rp[j]);
        }
      }
      // ---- store ----
      if (p.d_fp32) {
        float* dp = reinterpret_cast<float*>(p.d) + static_cast<size_t>(row) * p.ldd + col;
        if (full4 && p.vec_d) {
          float4 o = make_float4(v[0], v[1], v[2], v[3]);
          if (p.accumulate) {
            const float4 c4 = *reinterpret_cast<const float4*>(dp);
            o.x += c4.x; o.y += c4.y; o.z += c4.z; o.w += c4.w;
          }
          *reinterpret_cast<float4*>(dp) = o;
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (col + j < n_limit) dp[j] = p.accumulate ? dp[j] + v[j] : v[j];
        }
      } else {
        __nv_bfloat16* dp = reinterpret_cast<__nv_bfloat16*>(p.d) + static_cast<size_t>(row) * p.ldd + col;
        __nv_bfloat162 t0 = __floats2bfloat162_rn(v[0], v[1]), t1 = __floats2bfloat162_rn(v[2], v[3]);
        cs[0] += __bfloat162float(t0.x); cs[1] += __bfloat162float(t0.y);
        cs[2] += __bfloat162float(t1.x); cs[3] += __bfloat162float(t1.y);
        if (full4 && p.vec_d) {
          *reinterpret_cast<uint2*>(dp) = make_uint2(*reinterpret_cast<uint32_t*>(&t0), *reinterpret_cast<uint32_t*>(&t1));
        } else {
          __nv_bfloat16 tv[4] = {t0.x, t0.y, t1.x, t1.y};
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (col + j < n_limit) dp[j] = tv[j];
        }
      }
    }
    __syncwarp();   // staging is reused by the next chunk
    if (p.dbias != nullptr && !p.d_fp32) {   // warp-uniform
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        cs[j] += __shfl_xor_sync(0xffffffffu, cs[j], 8);
        cs[j] += __shfl_xor_sync(0xffffffffu, cs[j], 16);
      }
      if (lrow == 0) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (col + j < n_limit) atomicAdd(p.dbias + col + j, cs[j]);
      }
    }
  }
  if (warp_active) tmem_ld_wait();
}

template <int BN>
__device__ __forceinline__ void epilogue_dispatch(const GemmDev& p, int row_base, int col_base, int n_limit, uint32_t taddr,
                                                  float* stage, const uint64_t* seed_tab, uint64_t seed, int lane,
                                                  int label_pre, float bias_pre) {
  // the column-sum (bias gradient) by-product is only wanted by the backward GEMMs: keep it out of
  // the instruction stream of the forward ones (it is ~20% of a plain epilogue)
  const bool db = p.dbias != nullptr;
  switch (p.epi_kind) {
    case EPI_PLAIN:
      if (db) epilogue_tile_fast<BN, EPI_PLAIN, true>(p, row_base, col_base, n_limit, taddr, stage, lane);
      else epilogue_tile_fast<BN, EPI_PLAIN, false>(p, row_base, col_base, n_limit, taddr, stage, lane);
      break;
    case EPI_PLAIN_F32: epilogue_tile_fast<BN, EPI_PLAIN_F32, false>(p, row_base, col_base, n_limit, taddr, stage, lane); break;
    case EPI_BIAS: epilogue_tile_fast<BN, EPI_BIAS, false>(p, row_base, col_base, n_limit, taddr, stage, lane); break;
    case EPI_BIAS_RESID: epilogue_tile_fast<BN, EPI_BIAS_RESID, false>(p, row_base, col_base, n_limit, taddr, stage, lane); break;
    case EPI_BIAS_GELU_PRE: epilogue_tile_fast<BN, EPI_BIAS_GELU_PRE, false>(p, row_base, col_base, n_limit, taddr, stage, lane); break;
    case EPI_BWD_GELU:
      if (db) epilogue_tile_fast<BN, EPI_BWD_GELU, true>(p, row_base, col_base, n_limit, taddr, stage, lane);
      else epilogue_tile_fast<BN, EPI_BWD_GELU, false>(p, row_base, col_base, n_limit, taddr, stage, lane);
      break;
    case EPI_XENT: epilogue_tile_xent(p, row_base, col_base, taddr, stage, lane, label_pre, bias_pre); break;
    default: epilogue_tile<BN>(p, row_base, col_base, n_limit, taddr, stage, seed_tab, seed, lane); break;
  }
}

template <int BN, int A_MN, int B_MN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ TmapSet tma_b,
                    const GemmDev p) {
  using Cfg = GemmCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int m_tiles = (p.M + BM - 1) / BM;
  int n_tiles;
  if (p.shard_mode == 2)
    n_tiles = ((p.shard_len + BN - 1) / BN) * p.nsrc;
  else
    n_tiles = (p.N + BN - 1) / BN;
  const int num_tiles = m_tiles * n_tiles * p.k_splits;
  int kps = 0, nk;
  if (p.shard_mode == 1) {
    // TMA needs the innermost coordinate 16-byte aligned. A shard that starts at an A column that
    // is not a multiple of 8 is read from the aligned column below it, and B's row coordinate
    // (outer dimension, may go negative: zero-filled) moves down by the same shift, so up to 7
    // extra K rows per shard may be touched.
    kps = (p.shard_len + ((p.shard_len & 7) ? 7 : 0) + BK - 1) / BK;
    nk = kps * p.nsrc;
  } else {
    nk = (p.K + BK - 1) / BK;
  }
  // split-K (never combined with K-sharded B): split s of a tile covers k blocks [s*nk_split, ...)
  const int nk_total = nk;
  const int nk_split = (nk_total + p.k_splits - 1) / p.k_splits;
  const int mn_tiles = m_tiles * n_tiles;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    for (int s = 0; s < p.nsrc; ++s) tma_prefetch_desc(&tma_b.m[s]);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], 8);
    }
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<Cfg::TMEM_COLS>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_sync();   // everything above overlapped the previous kernel's tail; global memory is touched below

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {   // elect.sync keeps the issue code on the uniform datapath
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int split = tile / mn_tiles;
        TileCoord tc = tile_coord(p, tile - split * mn_tiles, m_tiles, BN);
        const int it_end = min(nk_total, (split + 1) * nk_split);
        for (int it = split * nk_split; it < it_end; ++it) {
          int a_k, b_k, b_src;
          if (p.shard_mode == 1) {
            int s = it / kps, kb = it - s * kps;
            b_src = p.rot + s;
            if (b_src >= p.nsrc) b_src -= p.nsrc;
            const int base = b_src * p.shard_len;
            const int shift = base & 7;
            a_k = base - shift + kb * BK;
            b_k = kb * BK - shift;
          } else {
            a_k = b_k = it * BK;
            b_src = tc.b_src;
          }
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * Cfg::STAGE_BYTES;
          uint8_t* sb = sa + Cfg::A_BYTES;
          mbar_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
          if (A_MN == 0) {
            tma_load_2d(sa, &tma_a, &full_bar[stage], a_k, tc.m0);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j)
              tma_load_2d(sa + j * (64 * BK * 2), &tma_a, &full_bar[stage], tc.m0 + j * 64, a_k);
          }
          const CUtensorMap* bm = &tma_b.m[b_src];
          if (B_MN == 0) {
            tma_load_2d(sb, bm, &full_bar[stage], b_k, tc.b_n);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_2d(sb + j * (64 * BK * 2), bm, &full_bar[stage], tc.b_n + j * 64, b_k);
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = umma_idesc_bf16(BM, BN, A_MN, B_MN);
    // K-major SW128: rows of 128 B, 8-row groups 1024 B apart; next K=16 slice is +32 B.
    // MN-major SW128: k rows of 128 B (64 MN elements), 8-k groups 1024 B apart (SBO),
    //                 64-element MN blocks BK*128 B apart (LBO); next K=16 slice is +2048 B.
    constexpr uint32_t A_LBO = A_MN ? BK * 128 : 16, A_SBO = 1024, A_KSTEP = A_MN ? 2048 : 32;
    constexpr uint32_t B_LBO = B_MN ? BK * 128 : 16, B_SBO = 1024, B_KSTEP = B_MN ? 2048 : 32;
    int stage = 0;
    uint32_t phase = 0;
    int local = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
      const int as = local & 1;
      const uint32_t aphase = (local >> 1) & 1;
      mbar_wait(&tmem_empty[as], aphase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * BN;
      const int split = tile / mn_tiles;
      const int it_begin = split * nk_split, it_end = min(nk_total, (split + 1) * nk_split);
      for (int it = it_begin; it < it_end; ++it) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {   // elect.sync keeps the issue code on the uniform datapath
          const uint32_t sa = smem_u32(smem + stage * Cfg::STAGE_BYTES);
          const uint32_t sb = sa + Cfg::A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            uint64_t da = umma_smem_desc(sa + k * A_KSTEP, A_LBO, A_SBO);
            uint64_t db = umma_smem_desc(sb + k * B_KSTEP, B_LBO, B_SBO);
            umma_bf16(d_tmem, da, db, idesc, ((it - it_begin) | k) ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);
          if (it == it_end - 1) umma_commit(&tmem_full[as]);
        }
        __syncwarp();
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else {
    // ===================== epilogue (warps 2..9) =====================
    // Two warps per TMEM lane quadrant; each takes half of the tile's columns. A 32-row x 32-col
    // fp32 chunk goes TMEM -> registers (thread = row) -> per-warp swizzled staging in shared
    // memory -> registers again with 8 lanes per row (lane = 4 consecutive columns), so that every
    // global access of the epilogue (bias, saved pre-activation, residual, output) is a
    // row-contiguous vector access and the element-wise math runs on all 32 lanes whatever M is.
    const int ew = warp - 2;
    const int quad = warp & 3;          // TMEM lane quadrant this warp may read
    const int half = ew >> 2;           // which half of the tile's columns
    float* stage = reinterpret_cast<float*>(smem + STAGES * Cfg::STAGE_BYTES + Cfg::BAR_BYTES) + ew * (32 * 32);
    uint64_t seed = 0;
    uint64_t* seed_tab = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES + Cfg::BAR_BYTES + 8 * 32 * 32 * 4) + ew * 16;
    if (p.drop_key != nullptr) {
      const uint64_t key = *p.drop_key;
      seed = mix64(key ^ p.drop_salt);
      if (lane < p.drop_nsalt) seed_tab[lane] = mix64(key ^ p.drop_salts[lane]);   // one RNG stream per minibatch
      __syncwarp();
    }
    int local = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
      const int as = local & 1;
      const uint32_t aphase = (local >> 1) & 1;
      TileCoord tc = tile_coord(p, tile % mn_tiles, m_tiles, BN);
      int label_pre;
      float bias_pre;
      xent_prefetch(p, tc.m0 + quad * 32, tc.n0 + half * (BN / 2), lane, label_pre, bias_pre);
      mbar_wait(&tmem_full[as], aphase);
      tc_fence_after();
      epilogue_dispatch<BN>(p, tc.m0 + quad * 32, tc.n0 + half * (BN / 2), tc.n_limit,
                        tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + as * BN + half * (BN / 2), stage, seed_tab,
                        seed, lane, label_pre, bias_pre);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[as]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
  }
}

// =============================================================================================
// CTA-pair variant (cta_group::2): a cluster of two CTAs on the two SMs of a TPC computes one
// 256 x BN tile. Each CTA loads its own 128 rows of A and its own half (BN/2 columns) of B; the
// leader CTA issues tcgen05.mma with M = 256 and the hardware reads both halves of B from the two
// CTAs' shared memories. Per flop each SM pulls 1/3 fewer operand bytes from L2 than with a
// single-CTA 128 x BN tile, which is what the K ~ 768 GEMMs of the transformer are bound by.
// Barriers: TMA of both CTAs completes on the leader's full barrier; MMA completion is multicast
// to the empty / accumulator-full barriers of both CTAs; epilogue warps of both CTAs arrive on
// the leader's accumulator-empty barrier.
// =============================================================================================
template <int BN>
struct PairCfg {
  static constexpr int BNH = BN / 2;                 // B columns held by each CTA
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BNH * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = 6;
  static constexpr uint32_t TMEM_COLS = (2 * BN <= 256) ? 256 : 512;
  static constexpr int BAR_BYTES = 256;
  static constexpr int EPI_STAGE_BYTES = 8 * 32 * 32 * 4 + 8 * 16 * 8;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + BAR_BYTES + EPI_STAGE_BYTES + 1024;
};

template <int BN, int A_MN, int B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_pair_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ TmapSet tma_b,
                         const GemmDev p) {
  using Cfg = PairCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  constexpr int BNH = Cfg::BNH;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES);   // used in the leader
  uint64_t* empty_bar = full_bar + STAGES;                                              // both CTAs
  uint64_t* tmem_full = empty_bar + STAGES;                                             // both CTAs
  uint64_t* tmem_empty = tmem_full + 2;                                                 // used in the leader
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;

  constexpr int BM2 = 2 * BM;
  const int m_tiles = (p.M + BM2 - 1) / BM2;
  int n_tiles;
  if (p.shard_mode == 2)
    n_tiles = ((p.shard_len + BN - 1) / BN) * p.nsrc;
  else
    n_tiles = (p.N + BN - 1) / BN;
  const int mn_tiles = m_tiles * n_tiles;
  const int num_tiles = mn_tiles * p.k_splits;
  int kps = 0, nk;
  if (p.shard_mode == 1) {
    kps = (p.shard_len + ((p.shard_len & 7) ? 7 : 0) + BK - 1) / BK;
    nk = kps * p.nsrc;
  } else {
    nk = (p.K + BK - 1) / BK;
  }
  const int nk_total = nk;
  const int nk_split = (nk_total + p.k_splits - 1) / p.k_splits;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    for (int s = 0; s < p.nsrc; ++s) tma_prefetch_desc(&tma_b.m[s]);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], 16);   // 8 epilogue warps in each CTA of the pair
    }
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc_pair<Cfg::TMEM_COLS>(tmem_ptr);
  tc_fence_before();
  cluster_sync_all();                  // barriers of both CTAs initialised before any remote arrive
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_sync();

  // tile -> coordinates (same order as the single-CTA kernel, M tiles of 256 rows)
  auto coord = [&](int tile) {
    TileCoord t;
    int mt = tile % m_tiles;
    int nt = tile / m_tiles;
    t.m0 = mt * BM2;
    if (p.shard_mode == 2) {
      int tps = (p.shard_len + BN - 1) / BN;
      int ntl = tps * p.nsrc;
      nt = (nt + p.rot * tps) % ntl;
      int src = nt / tps, loc = nt % tps;
      t.b_src = src;
      t.b_n = loc * BN;
      t.n0 = src * p.shard_len + loc * BN;
      int lim = (src + 1) * p.shard_len;
      t.n_limit = lim < p.N ? lim : p.N;
    } else {
      t.b_src = 0;
      t.b_n = nt * BN;
      t.n0 = nt * BN;
      t.n_limit = p.N;
    }
    return t;
  };

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (elect_one()) {   // elect.sync keeps the issue code on the uniform datapath
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        const int split = tile / mn_tiles;
        TileCoord tc = coord(tile - split * mn_tiles);
        const int my_m0 = tc.m0 + static_cast<int>(rank) * BM;
        const int my_bn = tc.b_n + static_cast<int>(rank) * BNH;
        const int it_end = min(nk_total, (split + 1) * nk_split);
        for (int it = split * nk_split; it < it_end; ++it) {
          int a_k, b_k, b_src;
          if (p.shard_mode == 1) {
            int s = it / kps, kb = it - s * kps;
            b_src = p.rot + s;
            if (b_src >= p.nsrc) b_src -= p.nsrc;
            const int base = b_src * p.shard_len;
            const int shift = base & 7;
            a_k = base - shift + kb * BK;
            b_k = kb * BK - shift;
          } else {
            a_k = b_k = it * BK;
            b_src = tc.b_src;
          }
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * Cfg::STAGE_BYTES;
          uint8_t* sb = sa + Cfg::A_BYTES;
          const uint32_t full_leader = mapa_cluster(smem_u32(&full_bar[stage]), 0);
          if (leader) mbar_expect_tx(&full_bar[stage], 2 * Cfg::STAGE_BYTES);   // bytes of both CTAs
          if (A_MN == 0) {
            tma_load_2d_pair(sa, &tma_a, full_leader, a_k, my_m0);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j)
              tma_load_2d_pair(sa + j * (64 * BK * 2), &tma_a, full_leader, my_m0 + j * 64, a_k);
          }
          const CUtensorMap* bm = &tma_b.m[b_src];
          if (B_MN == 0) {
            tma_load_2d_pair(sb, bm, full_leader, b_k, my_bn);
          } else {
#pragma unroll
            for (int j = 0; j < BNH / 64; ++j)
              tma_load_2d_pair(sb + j * (64 * BK * 2), bm, full_leader, my_bn + j * 64, b_k);
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM2, BN, A_MN, B_MN);
      constexpr uint32_t A_LBO = A_MN ? BK * 128 : 16, A_SBO = 1024, A_KSTEP = A_MN ? 2048 : 32;
      constexpr uint32_t B_LBO = B_MN ? BK * 128 : 16, B_SBO = 1024, B_KSTEP = B_MN ? 2048 : 32;
      int stage = 0;
      uint32_t phase = 0;
      int local = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++local) {
        const int as = local & 1;
        const uint32_t aphase = (local >> 1) & 1;
        mbar_wait(&tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        const int split = tile / mn_tiles;
        const int it_begin = split * nk_split, it_end = min(nk_total, (split + 1) * nk_split);
        for (int it = it_begin; it < it_end; ++it) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {   // elect.sync keeps the issue code on the uniform datapath
            const uint32_t sa = smem_u32(smem + stage * Cfg::STAGE_BYTES);
            const uint32_t sb = sa + Cfg::A_BYTES;
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) {
              uint64_t da = umma_smem_desc(sa + k * A_KSTEP, A_LBO, A_SBO);
              uint64_t db = umma_smem_desc(sb + k * B_KSTEP, B_LBO, B_SBO);
              umma_bf16_pair(d_tmem, da, db, idesc, ((it - it_begin) | k) ? 1u : 0u);
            }
            umma_commit_pair(&empty_bar[stage]);
            if (it == it_end - 1) umma_commit_pair(&tmem_full[as]);
          }
          __syncwarp();
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else {
    // ===================== epilogue (warps 2..9 of both CTAs) =====================
    const int ew = warp - 2;
    const int quad = warp & 3;
    const int half = ew >> 2;
    float* stage = reinterpret_cast<float*>(smem + STAGES * Cfg::STAGE_BYTES + Cfg::BAR_BYTES) + ew * (32 * 32);
    uint64_t seed = 0;
    uint64_t* seed_tab = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES + Cfg::BAR_BYTES + 8 * 32 * 32 * 4) + ew * 16;
    if (p.drop_key != nullptr) {
      const uint64_t key = *p.drop_key;
      seed = mix64(key ^ p.drop_salt);
      if (lane < p.drop_nsalt) seed_tab[lane] = mix64(key ^ p.drop_salts[lane]);
      __syncwarp();
    }
    int local = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++local) {
      const int as = local & 1;
      const uint32_t aphase = (local >> 1) & 1;
      TileCoord tc = coord(tile % mn_tiles);
      int label_pre;
      float bias_pre;
      xent_prefetch(p, tc.m0 + static_cast<int>(rank) * BM + quad * 32, tc.n0 + half * (BN / 2), lane, label_pre, bias_pre);
      mbar_wait(&tmem_full[as], aphase);
      tc_fence_after();
      epilogue_dispatch<BN>(p, tc.m0 + static_cast<int>(rank) * BM + quad * 32, tc.n0 + half * (BN / 2), tc.n_limit,
                        tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + as * BN + half * (BN / 2), stage, seed_tab,
                        seed, lane, label_pre, bias_pre);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa_cluster(smem_u32(&tmem_empty[as]), 0));
    }
  }

  tc_fence_before();
  cluster_sync_all();                  // neither CTA may exit (or free TMEM) while the other still uses it
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair<Cfg::TMEM_COLS>(tmem_base);
  }
}

// ---------------------------------------------------------------------------------------------
// host
// ---------------------------------------------------------------------------------------------
template <int BN, int A_MN, int B_MN>
static int launch_gemm(const b200_gemm_args& a, const GemmDev& dev, cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  CUtensorMap ta;
  TmapSet tb;
  memset(&tb, 0, sizeof(tb));
  int rc;
  if (A_MN == 0)
    rc = make_tmap_2d(&ta, a.a, 2, (uint64_t)a.k, (uint64_t)a.m, (uint64_t)a.lda * 2, BK, BM, 128);
  else
    rc = make_tmap_2d(&ta, a.a, 2, (uint64_t)a.m, (uint64_t)a.k, (uint64_t)a.lda * 2, 64, BK, 128);
  if (rc) return rc;
  // logical extent of one B source
  uint64_t bn = a.n, bk = a.k;
  if (a.b_shard_mode == 1) bk = a.b_shard_len;
  if (a.b_shard_mode == 2) bn = a.b_shard_len;
  for (int s = 0; s < dev.nsrc; ++s) {
    if (B_MN == 0)
      rc = make_tmap_2d(&tb.m[s], a.b_src[s], 2, bk, bn, (uint64_t)a.ldb * 2, BK, BN, 128);
    else
      rc = make_tmap_2d(&tb.m[s], a.b_src[s], 2, bn, bk, (uint64_t)a.ldb * 2, 64, BK, 128);
    if (rc) return rc;
  }
  auto kern = gemm_tcgen05_kernel<BN, A_MN, B_MN>;
  static unsigned long long attr_done = 0;
  if (first_use_on_this_device(attr_done)) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   Cfg::SMEM_BYTES));
  }
  int m_tiles = (a.m + BM - 1) / BM;
  int n_tiles = (a.b_shard_mode == 2) ? ((a.b_shard_len + BN - 1) / BN) * dev.nsrc
                                      : (a.n + BN - 1) / BN;
  long long tiles = (long long)m_tiles * n_tiles * dev.k_splits;
  int grid = (int)(tiles < num_sms() ? tiles : num_sms());
  B200_CUDA(launch_k(kern, dim3(grid), dim3(GEMM_THREADS), Cfg::SMEM_BYTES, stream, ta, tb, dev));
  B200_CUDA(cudaGetLastError());
  return 0;
}

template <int BN, int A_MN, int B_MN>
static int launch_gemm_pair(const b200_gemm_args& a, const GemmDev& dev, cudaStream_t stream) {
  using Cfg = PairCfg<BN>;
  CUtensorMap ta;
  TmapSet tb;
  memset(&tb, 0, sizeof(tb));
  int rc;
  if (A_MN == 0)
    rc = make_tmap_2d(&ta, a.a, 2, (uint64_t)a.k, (uint64_t)a.m, (uint64_t)a.lda * 2, BK, BM, 128);
  else
    rc = make_tmap_2d(&ta, a.a, 2, (uint64_t)a.m, (uint64_t)a.k, (uint64_t)a.lda * 2, 64, BK, 128);
  if (rc) return rc;
  uint64_t bn = a.n, bk = a.k;
  if (a.b_shard_mode == 1) bk = a.b_shard_len;
  if (a.b_shard_mode == 2) bn = a.b_shard_len;
  for (int s = 0; s < dev.nsrc; ++s) {
    if (B_MN == 0)
      rc = make_tmap_2d(&tb.m[s], a.b_src[s], 2, bk, bn, (uint64_t)a.ldb * 2, BK, Cfg::BNH, 128);
    else
      rc = make_tmap_2d(&tb.m[s], a.b_src[s], 2, bn, bk, (uint64_t)a.ldb * 2, 64, BK, 128);
    if (rc) return rc;
  }
  auto kern = gemm_tcgen05_pair_kernel<BN, A_MN, B_MN>;
  static unsigned long long attr_done = 0;
  if (first_use_on_this_device(attr_done)) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
  }
  int m_tiles = (a.m + 2 * BM - 1) / (2 * BM);
  int n_tiles = (a.b_shard_mode == 2) ? ((a.b_shard_len + BN - 1) / BN) * dev.nsrc : (a.n + BN - 1) / BN;
  long long tiles = (long long)m_tiles * n_tiles * dev.k_splits;
  int clusters = (int)(tiles < num_sms() / 2 ? tiles : num_sms() / 2);
  B200_CUDA(launch_k(kern, dim3(clusters * 2), dim3(GEMM_THREADS), Cfg::SMEM_BYTES, stream, ta, tb, dev));
  B200_CUDA(cudaGetLastError());
  return 0;
}

template <int BN>
static int dispatch_major_pair(const b200_gemm_args& a, const GemmDev& dev, cudaStream_t s) {
  if (a.a_mn == 0 && a.b_mn == 0) return launch_gemm_pair<BN, 0, 0>(a, dev, s);
  if (a.a_mn == 0 && a.b_mn == 1) return launch_gemm_pair<BN, 0, 1>(a, dev, s);
  if (a.a_mn == 1 && a.b_mn == 0) return launch_gemm_pair<BN, 1, 0>(a, dev, s);
  return launch_gemm_pair<BN, 1, 1>(a, dev, s);
}

template <int BN>
static int dispatch_major(const b200_gemm_args& a, const GemmDev& dev, cudaStream_t s) {
  if (a.a_mn == 0 && a.b_mn == 0) return launch_gemm<BN, 0, 0>(a, dev, s);
  if (a.a_mn == 0 && a.b_mn == 1) return launch_gemm<BN, 0, 1>(a, dev, s);
  if (a.a_mn == 1 && a.b_mn == 0) return launch_gemm<BN, 1, 0>(a, dev, s);
  return launch_gemm<BN, 1, 1>(a, dev, s);
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace b200

extern "C" int b200_gemm_bf16(const b200_gemm_args* args, void* stream) {
  using namespace b200;
  const b200_gemm_args& a = *args;
  B200_REQUIRE(a.m > 0 && a.n > 0 && a.k > 0, "gemm: empty problem %dx%dx%d", a.m, a.n, a.k);
  B200_REQUIRE(a.a != nullptr && a.d != nullptr, "gemm: null operand");
  B200_REQUIRE(a.lda % 8 == 0 && a.ldb % 8 == 0, "gemm: lda/ldb must be multiples of 8 (got %d, %d)",
               a.lda, a.ldb);
  B200_REQUIRE(!(a.accumulate && !a.d_fp32), "gemm: accumulate needs an fp32 output");
  int nsrc = a.b_shard_mode == 0 ? 1 : a.b_nsrc;
  B200_REQUIRE(nsrc >= 1 && nsrc <= 8, "gemm: b_nsrc %d out of range", nsrc);
  for (int s = 0; s < nsrc; ++s) B200_REQUIRE(a.b_src[s] != nullptr, "gemm: null B source %d", s);
  if (a.b_shard_mode == 1)
    B200_REQUIRE((long long)a.b_shard_len * nsrc == a.k, "gemm: K shards (%d x %d) != K (%d)",
                 a.b_shard_len, nsrc, a.k);
  if (a.b_shard_mode == 1 && !a.b_mn)
    B200_REQUIRE(a.b_shard_len % 8 == 0,
                 "gemm: K-sharded K-major B needs shard_len %% 8 == 0 (got %d): gather it first",
                 a.b_shard_len);
  if (a.b_shard_mode == 2)
    B200_REQUIRE((long long)a.b_shard_len * nsrc == a.n, "gemm: N shards (%d x %d) != N (%d)",
                 a.b_shard_len, nsrc, a.n);
  GemmDev dev;
  memset(&dev, 0, sizeof(dev));
  dev.M = a.m; dev.N = a.n; dev.K = a.k;
  dev.shard_mode = a.b_shard_mode; dev.nsrc = nsrc; dev.shard_len = a.b_shard_len;
  dev.rot = nsrc > 1 ? ((a.b_rot % nsrc) + nsrc) % nsrc : 0;
  dev.d = a.d; dev.ldd = a.ldd; dev.d_fp32 = a.d_fp32; dev.accumulate = a.accumulate;
  dev.alpha = a.alpha == 0.0f ? 1.0f : a.alpha;
  dev.bias = static_cast<const __nv_bfloat16*>(a.bias);
  dev.pre = static_cast<__nv_bfloat16*>(a.pre); dev.ldpre = a.ldpre;
  dev.act = a.act;
  dev.resid = static_cast<const __nv_bfloat16*>(a.resid); dev.ldres = a.ldres;
  dev.bwd = a.bwd;
  dev.aux = static_cast<const __nv_bfloat16*>(a.aux); dev.ldaux = a.ldaux;
  dev.drop_key = (a.drop_rate > 0.0f) ? a.drop_key : nullptr;
  dev.drop_salt = a.drop_salt;
  B200_REQUIRE(a.drop_nsalt >= 0 && a.drop_nsalt <= 16, "gemm: at most 16 dropout row blocks (got %d)", a.drop_nsalt);
  dev.drop_nsalt = a.drop_nsalt;
  dev.drop_rows_per_salt = a.drop_rows_per_salt > 0 ? a.drop_rows_per_salt : 1;
  for (int i = 0; i < a.drop_nsalt; ++i) dev.drop_salts[i] = a.drop_salts[i];
  float keep = 1.0f - a.drop_rate;
  dev.drop_thresh16 = (uint32_t)(keep * 65536.0f + 0.5f);
  dev.drop_scale = keep > 0.f ? 1.0f / keep : 0.f;
  // the epilogue moves 4 elements per lane: 16 bytes of fp32 or 8 bytes of bf16; column offsets
  // are multiples of 4 unless an N-shard boundary is not
  const bool cols_ok = a.b_shard_mode != 2 || a.b_shard_len % 4 == 0;
  dev.vec_d = aligned16(a.d) && (a.ldd % 4 == 0) && cols_ok;
  dev.vec_pre = a.pre && aligned16(a.pre) && (a.ldpre % 4 == 0) && cols_ok;
  dev.vec_aux = a.aux && aligned16(a.aux) && (a.ldaux % 4 == 0) && cols_ok;
  dev.vec_res = a.resid && aligned16(a.resid) && (a.ldres % 4 == 0) && cols_ok;

  {
    const bool n4 = (a.n % 4 == 0) && cols_ok;
    const bool no_drop = dev.drop_key == nullptr;
    const bool bias_ok = a.bias == nullptr || (reinterpret_cast<uintptr_t>(a.bias) & 7) == 0;
    int kind = EPI_GENERIC;
    if (n4 && no_drop && dev.vec_d && bias_ok) {
      if (a.bwd) {
        if (a.aux == nullptr && a.act == B200_ACT_NONE) kind = a.d_fp32 ? EPI_PLAIN_F32 : EPI_PLAIN;
        else if (a.aux != nullptr && dev.vec_aux && a.act == B200_ACT_GELU && !a.d_fp32) kind = EPI_BWD_GELU;
      } else if (a.bias == nullptr && a.pre == nullptr && a.resid == nullptr && a.act == B200_ACT_NONE) {
        kind = a.d_fp32 ? EPI_PLAIN_F32 : EPI_PLAIN;
      } else if (!a.d_fp32 && a.bias != nullptr && a.act == B200_ACT_NONE && a.pre == nullptr) {
        if (a.resid == nullptr) kind = EPI_BIAS;
        else if (dev.vec_res) kind = EPI_BIAS_RESID;
      } else if (!a.d_fp32 && a.bias != nullptr && a.act == B200_ACT_GELU && a.pre != nullptr && dev.vec_pre &&
                 a.resid == nullptr) {
        kind = EPI_BIAS_GELU_PRE;
      }
    }
    // the specialised epilogues produce the bias-gradient column sums only where a backward pass asks
    // for them (plain and GELU' dgrad); any other combination goes through the generic epilogue
    if (a.dbias != nullptr && kind != EPI_PLAIN && kind != EPI_BWD_GELU) kind = EPI_GENERIC;
    if (a.xent_labels != nullptr) {
      B200_REQUIRE(a.n <= 32 && !a.bwd && !a.d_fp32 && a.act == B200_ACT_NONE && a.pre == nullptr && a.resid == nullptr &&
                       dev.drop_key == nullptr && a.b_shard_mode != 2 && !a.accumulate,
                   "gemm: the fused cross-entropy epilogue needs n <= 32, a bf16 output and a plain forward product");
      kind = EPI_XENT;
    }
    dev.epi_kind = kind;
  }
  int bn = a.block_n;
  if (bn == 0) {
    int n_eff = a.b_shard_mode == 2 ? a.b_shard_len : a.n;
    if (n_eff <= 64) bn = 64;
    else if (n_eff <= 128) bn = 128;
    else {
      // 128-wide tiles move twice the operand bytes per flop from L2 (measured: 30-50% slower on
      // every transformer shape), so they only win when 256-wide tiles would leave most of the
      // chip idle and the output cannot be split along K.
      long long m_tiles = (a.m + BM - 1) / BM;
      long long reps = a.b_shard_mode == 2 ? nsrc : 1;
      long long t256 = m_tiles * ((n_eff + 255) / 256) * reps;
      long long t128 = m_tiles * ((n_eff + 127) / 128) * reps;
      long long sms = num_sms();
      const bool splittable = dev.epi_kind == EPI_PLAIN_F32 && a.b_shard_mode != 1 && a.split_k != 1 &&
                              (a.k + BK - 1) / BK >= 8;   // enough K blocks to split
      double e256 = (double)t256 / (double)(((t256 + sms - 1) / sms) * sms);
      double e128 = (double)t128 / (double)(((t128 + sms - 1) / sms) * sms);
      // a problem that fits in one wave even with 128-wide tiles is latency-bound: more, smaller CTAs
      long long t64 = m_tiles * ((n_eff + 63) / 64) * reps;
      if (splittable) bn = 256;            // split-K fills the chip; wide tiles halve L2 traffic
      else if (t64 <= sms / 2) bn = 64;    // tiny problem: latency-bound, spread the weight stream over more CTAs
      else if (t128 <= sms) bn = 128;
      else bn = (e256 * 1.5 >= e128) ? 256 : 128;
    }
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // CTA pairs (256-row tiles) once the problem has at least two row tiles; pair_mode: 0 = auto,
  // 1 = never, 2 = always (when the tile shape allows it)
  const bool pair_ok = bn >= 128;
  // (auto: only for problems large enough to amortise the cluster launch and the pair barriers)
  const bool big = (long long)a.m * a.n * a.k >= (1LL << 27);
  const bool use_pair = pair_ok && (a.pair_mode == 2 || (a.pair_mode == 0 && a.m > BM && big));
  // split-K for fp32 outputs whose M x N grid cannot fill the chip (weight gradients: small M x N,
  // K = tokens). Partial products are added in memory with red.global.add, after a memset when the
  // call does not accumulate.
  dev.k_splits = 1;
  dev.atomic_out = 0;
  dev.dbias = a.dbias;
  dev.xent_labels = a.xent_labels;
  dev.xent_metrics = a.xent_metrics;
  dev.xent_scale = a.xent_scale;
  B200_REQUIRE(a.dbias == nullptr || (!a.d_fp32 && (reinterpret_cast<uintptr_t>(a.dbias) & 15) == 0),
               "gemm: dbias needs a bf16 output and a 16-byte aligned fp32 vector");
  if (dev.epi_kind == EPI_PLAIN_F32 && a.b_shard_mode != 1 && a.split_k != 1) {
    const long long row_tile = use_pair ? 2 * BM : BM;
    const long long mn = ((a.m + row_tile - 1) / row_tile) * (((a.b_shard_mode == 2 ? a.b_shard_len : a.n) + bn - 1) / bn) *
                         (a.b_shard_mode == 2 ? nsrc : 1);
    const long long slots = use_pair ? num_sms() / 2 : num_sms();
    const int nkb = (a.k + BK - 1) / BK;
    int splits = a.split_k > 1 ? a.split_k : (int)(slots / mn);
    if (splits > nkb / 4) splits = nkb / 4;          // keep at least 4 k-blocks per split
    if (splits > 16) splits = 16;
    if (splits >= 2) {
      int per = (nkb + splits - 1) / splits;
      splits = (nkb + per - 1) / per;                // no empty split
      if (splits >= 2) {
        dev.k_splits = splits;
        dev.atomic_out = 1;
        if (!a.accumulate) {
          if (a.ldd == a.n) {
            B200_CUDA(cudaMemsetAsync(a.d, 0, sizeof(float) * (size_t)a.m * a.n, s));
          } else {
            B200_CUDA(cudaMemset2DAsync(a.d, sizeof(float) * (size_t)a.ldd, 0, sizeof(float) * (size_t)a.n, a.m, s));
          }
        }
      }
    }
  }
  if (use_pair) {
    if (bn == 128) return dispatch_major_pair<128>(a, dev, s);
    return dispatch_major_pair<256>(a, dev, s);
  }
  switch (bn) {
    case 64: return dispatch_major<64>(a, dev, s);
    case 128: return dispatch_major<128>(a, dev, s);
    case 256: return dispatch_major<256>(a, dev, s);
    default: B200_REQUIRE(false, "gemm: unsupported block_n %d", bn);
  }
  return 0;
}
