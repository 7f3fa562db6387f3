// LayerNorm and embedding kernels for the GPT-2 configurations (BASELINE.json configs 2-5; the
// reference itself has no transformer). Memory-bound: one warp per row, 16-byte accesses, rows
// kept in registers between the statistics pass and the normalisation pass.
#include "../include/b200dist.h"
#include "common.cuh"

namespace b200 {

// LayerNorm rows are processed as 16-byte chunks, one per lane: NCH chunks per lane cover NCH * 256 columns
// (instantiated for 256 ... 2048 columns).

__device__ __forceinline__ void unpack8(const uint4& w, float (&f)[8]) {
  f[0] = __uint_as_float(w.x << 16); f[1] = __uint_as_float(w.x & 0xFFFF0000u);
  f[2] = __uint_as_float(w.y << 16); f[3] = __uint_as_float(w.y & 0xFFFF0000u);
  f[4] = __uint_as_float(w.z << 16); f[5] = __uint_as_float(w.z & 0xFFFF0000u);
  f[6] = __uint_as_float(w.w << 16); f[7] = __uint_as_float(w.w & 0xFFFF0000u);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  __nv_bfloat162 a = __floats2bfloat162_rn(f[0], f[1]), b = __floats2bfloat162_rn(f[2], f[3]);
  __nv_bfloat162 c = __floats2bfloat162_rn(f[4], f[5]), d = __floats2bfloat162_rn(f[6], f[7]);
  uint4 w;
  w.x = *reinterpret_cast<uint32_t*>(&a); w.y = *reinterpret_cast<uint32_t*>(&b);
  w.z = *reinterpret_cast<uint32_t*>(&c); w.w = *reinterpret_cast<uint32_t*>(&d);
  return w;
}

// cols % 8 == 0, cols <= NCH * 256. One warp per row, the row stays in registers; NCH is sized to the
// model width so that small rows leave registers for more resident warps (bytes in flight).
template <int NCH>
__global__ void __launch_bounds__(128)
layernorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ gamma,
                     const __nv_bfloat16* __restrict__ beta, __nv_bfloat16* __restrict__ y,
                     float* __restrict__ mean, float* __restrict__ rstd, int rows, int cols, float eps) {
  pdl_sync();
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  const int nchunk = cols >> 3;
  for (int row = blockIdx.x * wpb + (threadIdx.x >> 5); row < rows; row += gridDim.x * wpb) {
    const uint4* xr = reinterpret_cast<const uint4*>(x + static_cast<size_t>(row) * cols);
    float v[NCH][8];
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      int idx = c * 32 + lane;
      if (idx < nchunk) {
        unpack8(xr[idx], v[c]);
#pragma unroll
        for (int j = 0; j < 8; ++j) s += v[c][j];
      }
    }
    s = warp_sum(s);
    const float mu = s / cols;
    float q = 0.f;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      int idx = c * 32 + lane;
      if (idx < nchunk) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { float d = v[c][j] - mu; q += d * d; }
      }
    }
    q = warp_sum(q);
    const float rs = rsqrtf(q / cols + eps);
    uint4* yr = reinterpret_cast<uint4*>(y + static_cast<size_t>(row) * cols);
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      int idx = c * 32 + lane;
      if (idx < nchunk) {
        float g[8], b[8], o[8];
        unpack8(reinterpret_cast<const uint4*>(gamma)[idx], g);
        unpack8(reinterpret_cast<const uint4*>(beta)[idx], b);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (v[c][j] - mu) * rs * g[j] + b[j];
        yr[idx] = pack8(o);
      }
    }
    if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
  }
}

// dx = rstd * (dy*g - mean(dy*g) - xhat * mean(dy*g*xhat)) (+ dresid); dgamma += dy*xhat; dbeta += dy.
// A lane owns the same columns for every row it visits, so its dgamma/dbeta partial sums stay in
// registers across rows; they are combined across the block's warps in shared memory once at the
// end and flushed with one atomicAdd per column per block.
// Rows reach the block through a two-stage ring in shared memory filled by 1-D bulk copies
// (cp.async.bulk + mbarrier), one copy per operand per group of 8 consecutive rows (one row per
// warp): the next group of x / dy / dresid is always in flight while the current one is reduced,
// which a load -> reduce -> store loop over registers alone cannot do without doubling the
// register footprint. (Per-row copies of 1.5 KB were tried first: the TMA unit's per-copy cost
// made the kernel slower than plain loads.)
__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

constexpr int LN_BWD_WARPS = 8;

template <int NCH>   // 16-byte chunks per lane: cols <= NCH * 256
__global__ void __launch_bounds__(256, NCH <= 3 ? 2 : 1)    // wide rows: registers for the accumulators beat a second block
layernorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                     const __nv_bfloat16* __restrict__ gamma, const float* __restrict__ mean,
                     const float* __restrict__ rstd, const __nv_bfloat16* __restrict__ dresid,
                     __nv_bfloat16* __restrict__ dx, float* __restrict__ dgamma,
                     float* __restrict__ dbeta, float* __restrict__ dxsum, int rows, int cols) {
  extern __shared__ __align__(128) uint8_t sm_raw[];
  constexpr int STAGES = 2, W = LN_BWD_WARPS;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nchunk = cols >> 3;
  const uint32_t row_bytes = static_cast<uint32_t>(cols) * 2u;
  const uint32_t grp_bytes = row_bytes * W;          // one operand, one group of W rows
  const int narr = dresid != nullptr ? 3 : 2;
  uint8_t* ring = sm_raw;                                                       // [STAGES][3][W rows]
  uint64_t* full = reinterpret_cast<uint64_t*>(ring + static_cast<size_t>(STAGES) * 3 * grp_bytes);
  uint64_t* empty = full + STAGES;
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], W); }
    mbar_fence_init();
  }
  __syncthreads();
  pdl_sync();
  const int ngroups = (rows + W - 1) / W;
  auto issue = [&](int group, int stage) {     // one thread
    const int r0 = group * W;
    const uint32_t bytes = static_cast<uint32_t>(min(W, rows - r0)) * row_bytes;
    uint8_t* dst = ring + static_cast<size_t>(stage) * 3 * grp_bytes;
    mbar_expect_tx(&full[stage], bytes * narr);
    bulk_load_1d(dst, x + static_cast<size_t>(r0) * cols, bytes, &full[stage]);
    bulk_load_1d(dst + grp_bytes, dy + static_cast<size_t>(r0) * cols, bytes, &full[stage]);
    if (narr == 3) bulk_load_1d(dst + 2 * grp_bytes, dresid + static_cast<size_t>(r0) * cols, bytes, &full[stage]);
  };
  if (threadIdx.x == 0) {
    if (blockIdx.x < ngroups) issue(blockIdx.x, 0);
    if (blockIdx.x + gridDim.x < ngroups) issue(blockIdx.x + gridDim.x, 1);
  }
  float acc_g[NCH][8], acc_b[NCH][8], acc_x[NCH][8];   // acc_x: column sums of the emitted dx (optional)
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
#pragma unroll
    for (int j = 0; j < 8; ++j) { acc_g[c][j] = 0.f; acc_b[c][j] = 0.f; acc_x[c][j] = 0.f; }
  }
  int k = 0;
  for (int group = blockIdx.x; group < ngroups; group += gridDim.x, ++k) {
    const int stage = k & 1;
    const int row = group * W + warp;
    const bool live = row < rows;            // warp-uniform; only the last group can be short
    float mu = 0.f, rs = 0.f;
    if (live) { mu = mean[row]; rs = rstd[row]; }
    mbar_wait(&full[stage], (k >> 1) & 1);
    const uint8_t* base = ring + static_cast<size_t>(stage) * 3 * grp_bytes + static_cast<size_t>(warp) * row_bytes;
    const uint4* xs = reinterpret_cast<const uint4*>(base);
    const uint4* ds = reinterpret_cast<const uint4*>(base + grp_bytes);
    const uint4* rsd = reinterpret_cast<const uint4*>(base + 2 * grp_bytes);
    uint4 xw[NCH], dw[NCH], rw[NCH];       // the row stays in registers as packed bf16
    if (live) {
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        const int idx = c * 32 + lane;
        if (idx < nchunk) {
          xw[c] = xs[idx];
          dw[c] = ds[idx];
          if (narr == 3) rw[c] = rsd[idx];
        }
      }
    }
    __syncwarp();                          // every lane has its chunks
    if (lane == 0) mbar_arrive(&empty[stage]);
    if (threadIdx.x == 0 && group + 2 * gridDim.x < ngroups) {     // refill the stage once all warps have drained it
      mbar_wait(&empty[stage], (k >> 1) & 1);
      fence_proxy_async_smem();
      issue(group + 2 * gridDim.x, stage);
    }
    if (!live) continue;
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int idx = c * 32 + lane;
      if (idx < nchunk) {
        float xv[8], dv[8], g[8];
        unpack8(xw[c], xv);
        unpack8(dw[c], dv);
        unpack8(reinterpret_cast<const uint4*>(gamma)[idx], g);   // L1-resident
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float xh = (xv[j] - mu) * rs, dg = dv[j] * g[j];
          s1 += dg;
          s2 += dg * xh;
          acc_g[c][j] += dv[j] * xh;
          acc_b[c][j] += dv[j];
        }
      }
    }
    s1 = warp_sum(s1) / cols;
    s2 = warp_sum(s2) / cols;
    uint4* or_ = reinterpret_cast<uint4*>(dx + static_cast<size_t>(row) * cols);
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int idx = c * 32 + lane;
      if (idx < nchunk) {
        float xv[8], dv[8], g[8], o[8];
        unpack8(xw[c], xv);
        unpack8(dw[c], dv);
        unpack8(reinterpret_cast<const uint4*>(gamma)[idx], g);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rs * (dv[j] * g[j] - s1 - (xv[j] - mu) * rs * s2);
        if (narr == 3) {   // gradient arriving through the residual connection
          float rr[8];
          unpack8(rw[c], rr);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += rr[j];
        }
        const uint4 ow = pack8(o);
        or_[idx] = ow;
        if (dxsum != nullptr) {
          float q[8];
          unpack8(ow, q);      // sum what was stored (bf16), like a column sum over dx would
#pragma unroll
          for (int j = 0; j < 8; ++j) acc_x[c][j] += q[j];
        }
      }
    }
  }
  // Block reduction of the per-warp column sums: plain stores into the (now idle) ring laid out as
  // [warp][array][col] (exactly the ring's size), then one pass that adds the W partials of a
  // column and issues one reduction per column per block. (Shared-memory atomics here cost more
  // than the whole streaming part: the lanes' 32-byte stride makes them 8-way bank conflicted.)
  __syncthreads();
  float* part = reinterpret_cast<float*>(ring);
  const int narr_out = dxsum != nullptr ? 3 : 2;
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int idx = c * 32 + lane;
    if (idx < nchunk) {
      float* pg = part + (static_cast<size_t>(warp) * 3 + 0) * cols + idx * 8;
      float* pb = part + (static_cast<size_t>(warp) * 3 + 1) * cols + idx * 8;
      float* px = part + (static_cast<size_t>(warp) * 3 + 2) * cols + idx * 8;
      *reinterpret_cast<float4*>(pg) = make_float4(acc_g[c][0], acc_g[c][1], acc_g[c][2], acc_g[c][3]);
      *reinterpret_cast<float4*>(pg + 4) = make_float4(acc_g[c][4], acc_g[c][5], acc_g[c][6], acc_g[c][7]);
      *reinterpret_cast<float4*>(pb) = make_float4(acc_b[c][0], acc_b[c][1], acc_b[c][2], acc_b[c][3]);
      *reinterpret_cast<float4*>(pb + 4) = make_float4(acc_b[c][4], acc_b[c][5], acc_b[c][6], acc_b[c][7]);
      if (narr_out == 3) {
        *reinterpret_cast<float4*>(px) = make_float4(acc_x[c][0], acc_x[c][1], acc_x[c][2], acc_x[c][3]);
        *reinterpret_cast<float4*>(px + 4) = make_float4(acc_x[c][4], acc_x[c][5], acc_x[c][6], acc_x[c][7]);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < cols * narr_out; i += blockDim.x) {
    const int a = i / cols, col = i - a * cols;
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < W; ++w) t += part[(static_cast<size_t>(w) * 3 + a) * cols + col];
    float* dst = a == 0 ? dgamma : (a == 1 ? dbeta : dxsum);
    atomicAdd(&dst[col], t);
  }
}

__global__ void __launch_bounds__(128)
embedding_fwd_kernel(const int64_t* __restrict__ tokens, const __nv_bfloat16* __restrict__ wte,
                     const __nv_bfloat16* __restrict__ wpe, __nv_bfloat16* __restrict__ out, int rows, int seq,
                     int dim) {
  pdl_sync();
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  const int nchunk = dim >> 3;
  for (int row = blockIdx.x * wpb + (threadIdx.x >> 5); row < rows; row += gridDim.x * wpb) {
    const int64_t tok = tokens[row];
    const int t = row % seq;
    const uint4* a = reinterpret_cast<const uint4*>(wte + static_cast<size_t>(tok) * dim);
    const uint4* b = reinterpret_cast<const uint4*>(wpe + static_cast<size_t>(t) * dim);
    uint4* o = reinterpret_cast<uint4*>(out + static_cast<size_t>(row) * dim);
    for (int idx = lane; idx < nchunk; idx += 32) {
      float fa[8], fb[8];
      unpack8(a[idx], fa);
      unpack8(b[idx], fb);
#pragma unroll
      for (int j = 0; j < 8; ++j) fa[j] += fb[j];
      o[idx] = pack8(fa);
    }
  }
}

// Row-sharded tables in peer memory (FSDP): row `tok` lives on shard tok / rows_per_shard.
struct TablePeers { const __nv_bfloat16* p[8]; };
__global__ void __launch_bounds__(128)
embedding_fwd_sharded_kernel(const int64_t* __restrict__ tokens, const TablePeers wte, int64_t wte_rows,
                             const TablePeers wpe, int64_t wpe_rows, __nv_bfloat16* __restrict__ out, int rows,
                             int seq, int dim) {
  pdl_sync();
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  const int nchunk = dim >> 3;
  for (int row = blockIdx.x * wpb + (threadIdx.x >> 5); row < rows; row += gridDim.x * wpb) {
    const int64_t tok = tokens[row];
    const int64_t t = row % seq;
    const int64_t ws = tok / wte_rows, ps = t / wpe_rows;
    const uint4* a = reinterpret_cast<const uint4*>(wte.p[ws] + static_cast<size_t>(tok - ws * wte_rows) * dim);
    const uint4* b = reinterpret_cast<const uint4*>(wpe.p[ps] + static_cast<size_t>(t - ps * wpe_rows) * dim);
    uint4* o = reinterpret_cast<uint4*>(out + static_cast<size_t>(row) * dim);
    for (int idx = lane; idx < nchunk; idx += 32) {
      float fa[8], fb[8];
      unpack8(a[idx], fa);
      unpack8(b[idx], fb);
#pragma unroll
      for (int j = 0; j < 8; ++j) fa[j] += fb[j];
      o[idx] = pack8(fa);
    }
  }
}

__global__ void __launch_bounds__(128)
embedding_bwd_kernel(const int64_t* __restrict__ tokens, const __nv_bfloat16* __restrict__ dout,
                     float* __restrict__ dwte, float* __restrict__ dwpe, int rows, int seq, int dim) {
  pdl_sync();
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  const int nchunk = dim >> 3;
  for (int row = blockIdx.x * wpb + (threadIdx.x >> 5); row < rows; row += gridDim.x * wpb) {
    const int64_t tok = tokens[row];
    const int t = row % seq;
    const uint4* d = reinterpret_cast<const uint4*>(dout + static_cast<size_t>(row) * dim);
    float* gt = dwte + static_cast<size_t>(tok) * dim;
    float* gp = dwpe + static_cast<size_t>(t) * dim;
    for (int idx = lane; idx < nchunk; idx += 32) {
      float f[8];
      unpack8(d[idx], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (dwte != nullptr) atomicAdd(&gt[idx * 8 + j], f[j]);
        if (dwpe != nullptr) atomicAdd(&gp[idx * 8 + j], f[j]);
      }
    }
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean,
                                  float* rstd, int rows, int cols, float eps, void* stream) {
  B200_REQUIRE(rows > 0 && cols > 0 && cols % 8 == 0 && cols <= 2048, "layernorm: cols=%d must be a multiple of 8, <= 2048", cols);
  int blocks = (rows + 3) / 4;
  int cap = num_sms() * 16;
  if (blocks > cap) blocks = cap;
  auto launch = [&](auto kern) -> int {
    B200_CUDA(launch_k(kern, dim3(blocks), dim3(128), 0, static_cast<cudaStream_t>(stream),
        static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(gamma),
        static_cast<const __nv_bfloat16*>(beta), static_cast<__nv_bfloat16*>(y), mean, rstd, rows, cols, eps));
    return 0;
  };
  int rc;
  if (cols <= 256) rc = launch(layernorm_fwd_kernel<1>);
  else if (cols <= 512) rc = launch(layernorm_fwd_kernel<2>);
  else if (cols <= 768) rc = launch(layernorm_fwd_kernel<3>);
  else if (cols <= 1024) rc = launch(layernorm_fwd_kernel<4>);
  else rc = launch(layernorm_fwd_kernel<8>);
  if (rc) return rc;
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_layernorm_bwd_dxsum(const void* dy, const void* x, const void* gamma, const float* mean,
                                        const float* rstd, const void* dresid, void* dx, float* dgamma, float* dbeta,
                                        float* dxsum, int rows, int cols, void* stream);

extern "C" int b200_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* mean,
                                  const float* rstd, const void* dresid, void* dx, float* dgamma, float* dbeta,
                                  int rows, int cols, void* stream) {
  return b200_layernorm_bwd_dxsum(dy, x, gamma, mean, rstd, dresid, dx, dgamma, dbeta, nullptr, rows, cols, stream);
}

extern "C" int b200_layernorm_bwd_dxsum(const void* dy, const void* x, const void* gamma, const float* mean,
                                        const float* rstd, const void* dresid, void* dx, float* dgamma, float* dbeta,
                                        float* dxsum, int rows, int cols, void* stream) {
  B200_REQUIRE(rows > 0 && cols > 0 && cols % 8 == 0 && cols <= 2048, "layernorm: cols=%d must be a multiple of 8, <= 2048", cols);
  int blocks = (rows + LN_BWD_WARPS - 1) / LN_BWD_WARPS;
  int cap = num_sms() * (cols <= 768 ? 2 : 1);
  if (blocks > cap) blocks = cap;
  // two-stage ring of (x, dy, dresid) row groups (reused for the final block reduction) + its mbarriers
  const size_t smem = 2 * 3 * LN_BWD_WARPS * static_cast<size_t>(cols) * 2 + 4 * sizeof(uint64_t);
  static bool attr_set[5] = {false, false, false, false, false};
  int which = 0;
  auto launch = [&](auto kern) -> int {
    if (!attr_set[which]) {     // the instantiations share one function-pointer type: track them by index
      B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 3 * LN_BWD_WARPS * 2048 * 2 + 64));
      attr_set[which] = true;
    }
    B200_CUDA(launch_k(kern, dim3(blocks), dim3(256), smem, static_cast<cudaStream_t>(stream), 
        static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x),
        static_cast<const __nv_bfloat16*>(gamma), mean, rstd, static_cast<const __nv_bfloat16*>(dresid),
        static_cast<__nv_bfloat16*>(dx), dgamma, dbeta, dxsum, rows, cols));
    return 0;
  };
  int rc;
  if (cols <= 256) { which = 0; rc = launch(layernorm_bwd_kernel<1>); }
  else if (cols <= 512) { which = 1; rc = launch(layernorm_bwd_kernel<2>); }
  else if (cols <= 768) { which = 2; rc = launch(layernorm_bwd_kernel<3>); }
  else if (cols <= 1024) { which = 3; rc = launch(layernorm_bwd_kernel<4>); }
  else { which = 4; rc = launch(layernorm_bwd_kernel<8>); }
  if (rc) return rc;
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_embedding_fwd(const int64_t* tokens, const void* wte, const void* wpe, void* out, int batch,
                                  int seq, int dim, void* stream) {
  B200_REQUIRE(batch > 0 && seq > 0 && dim % 8 == 0, "embedding: dim=%d must be a multiple of 8", dim);
  int rows = batch * seq;
  int blocks = (rows + 3) / 4;
  int cap = num_sms() * 16;
  if (blocks > cap) blocks = cap;
  B200_CUDA(launch_k(embedding_fwd_kernel, dim3(blocks), dim3(128), 0, static_cast<cudaStream_t>(stream), 
      tokens, static_cast<const __nv_bfloat16*>(wte), static_cast<const __nv_bfloat16*>(wpe),
      static_cast<__nv_bfloat16*>(out), rows, seq, dim));
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_embedding_bwd(const int64_t* tokens, const void* dout, float* dwte, float* dwpe, int batch,
                                  int seq, int dim, void* stream) {
  B200_REQUIRE(batch > 0 && seq > 0 && dim % 8 == 0, "embedding: dim=%d must be a multiple of 8", dim);
  int rows = batch * seq;
  int blocks = (rows + 3) / 4;
  int cap = num_sms() * 16;
  if (blocks > cap) blocks = cap;
  B200_CUDA(launch_k(embedding_bwd_kernel, dim3(blocks), dim3(128), 0, static_cast<cudaStream_t>(stream), 
      tokens, static_cast<const __nv_bfloat16*>(dout), dwte, dwpe, rows, seq, dim));
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_embedding_fwd_sharded(const int64_t* tokens, void* const* wte_peers, int n_wte,
                                          int64_t wte_rows_per_shard, void* const* wpe_peers, int n_wpe,
                                          int64_t wpe_rows_per_shard, void* out, int batch, int seq, int dim,
                                          void* stream) {
  B200_REQUIRE(batch > 0 && seq > 0 && dim % 8 == 0, "embedding: dim=%d must be a multiple of 8", dim);
  B200_REQUIRE(n_wte >= 1 && n_wte <= 8 && n_wpe >= 1 && n_wpe <= 8, "embedding: 1..8 shards per table");
  TablePeers a, b;
  memset(&a, 0, sizeof(a));
  memset(&b, 0, sizeof(b));
  for (int i = 0; i < n_wte; ++i) a.p[i] = static_cast<const __nv_bfloat16*>(wte_peers[i]);
  for (int i = 0; i < n_wpe; ++i) b.p[i] = static_cast<const __nv_bfloat16*>(wpe_peers[i]);
  int rows = batch * seq;
  int blocks = (rows + 3) / 4;
  int cap = num_sms() * 16;
  if (blocks > cap) blocks = cap;
  B200_CUDA(launch_k(embedding_fwd_sharded_kernel, dim3(blocks), dim3(128), 0, static_cast<cudaStream_t>(stream), 
      tokens, a, wte_rows_per_shard, b, wpe_rows_per_shard, static_cast<__nv_bfloat16*>(out), rows, seq, dim));
  B200_CUDA(cudaGetLastError());
  return 0;
}
