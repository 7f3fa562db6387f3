// LayerNorm and embedding kernels for the GPT-2 configurations (BASELINE.json configs 2-5; the
// reference itself has no transformer). Memory-bound: one warp per row, 16-byte accesses, rows
// kept in registers between the statistics pass and the normalisation pass.
#include "../include/b200dist.h"
#include "common.cuh"

namespace b200 {

constexpr int LN_MAX_CHUNKS = 8;  // 8 chunks x 32 lanes x 8 elements = 2048 columns max

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

// cols % 8 == 0, cols <= 2048
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
    float v[LN_MAX_CHUNKS][8];
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < LN_MAX_CHUNKS; ++c) {
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
    for (int c = 0; c < LN_MAX_CHUNKS; ++c) {
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
    for (int c = 0; c < LN_MAX_CHUNKS; ++c) {
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
template <int NCH>   // 16-byte chunks per lane: cols <= NCH * 256
__global__ void __launch_bounds__(256, 2)
layernorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                     const __nv_bfloat16* __restrict__ gamma, const float* __restrict__ mean,
                     const float* __restrict__ rstd, const __nv_bfloat16* __restrict__ dresid,
                     __nv_bfloat16* __restrict__ dx, float* __restrict__ dgamma,
                     float* __restrict__ dbeta, float* __restrict__ dxsum, int rows, int cols) {
  pdl_sync();
  extern __shared__ float sm[];
  float* s_dg = sm;
  float* s_db = sm + cols;
  float* s_dx = sm + 2 * cols;
  for (int i = threadIdx.x; i < 3 * cols; i += blockDim.x) sm[i] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  const int nchunk = cols >> 3;
  float acc_g[NCH][8], acc_b[NCH][8], acc_x[NCH][8];   // acc_x: column sums of the emitted dx (optional)
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
#pragma unroll
    for (int j = 0; j < 8; ++j) { acc_g[c][j] = 0.f; acc_b[c][j] = 0.f; acc_x[c][j] = 0.f; }
  }
  for (int row = blockIdx.x * wpb + (threadIdx.x >> 5); row < rows; row += gridDim.x * wpb) {
    const uint4* xr = reinterpret_cast<const uint4*>(x + static_cast<size_t>(row) * cols);
    const uint4* dr = reinterpret_cast<const uint4*>(dy + static_cast<size_t>(row) * cols);
    const float mu = mean[row], rs = rstd[row];
    uint4 xw[NCH], dw[NCH], rw[NCH];       // the row stays in registers as packed bf16
#pragma unroll
    for (int c = 0; c < NCH; ++c) {        // all loads of the row first
      const int idx = c * 32 + lane;
      if (idx < nchunk) {
        xw[c] = xr[idx];
        dw[c] = dr[idx];
        if (dresid != nullptr) rw[c] = reinterpret_cast<const uint4*>(dresid + static_cast<size_t>(row) * cols)[idx];
      }
    }
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
        if (dresid != nullptr) {   // gradient arriving through the residual connection
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
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int idx = c * 32 + lane;
    if (idx < nchunk) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        atomicAdd(&s_dg[idx * 8 + j], acc_g[c][j]);
        atomicAdd(&s_db[idx * 8 + j], acc_b[c][j]);
        if (dxsum != nullptr) atomicAdd(&s_dx[idx * 8 + j], acc_x[c][j]);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < cols; i += blockDim.x) {
    atomicAdd(&dgamma[i], s_dg[i]);
    atomicAdd(&dbeta[i], s_db[i]);
    if (dxsum != nullptr) atomicAdd(&dxsum[i], s_dx[i]);
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
  B200_CUDA(launch_k(layernorm_fwd_kernel, dim3(blocks), dim3(128), 0, static_cast<cudaStream_t>(stream), 
      static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(gamma),
      static_cast<const __nv_bfloat16*>(beta), static_cast<__nv_bfloat16*>(y), mean, rstd, rows, cols, eps));
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
  int blocks = (rows + 7) / 8;
  int cap = num_sms() * 2;
  if (blocks > cap) blocks = cap;
  auto launch = [&](auto kern) -> int {
    B200_CUDA(launch_k(kern, dim3(blocks), dim3(256), 3 * cols * sizeof(float), static_cast<cudaStream_t>(stream), 
        static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x),
        static_cast<const __nv_bfloat16*>(gamma), mean, rstd, static_cast<const __nv_bfloat16*>(dresid),
        static_cast<__nv_bfloat16*>(dx), dgamma, dbeta, dxsum, rows, cols));
    return 0;
  };
  int rc;
  if (cols <= 256) rc = launch(layernorm_bwd_kernel<1>);
  else if (cols <= 512) rc = launch(layernorm_bwd_kernel<2>);
  else if (cols <= 768) rc = launch(layernorm_bwd_kernel<3>);
  else if (cols <= 1024) rc = launch(layernorm_bwd_kernel<4>);
  else rc = launch(layernorm_bwd_kernel<8>);
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
