// Memory-bound kernels: softmax cross-entropy (fwd+bwd fused), column sums, AdamW fused with
// the gradient scale and the bf16 working-copy cast, casts, RNG key evolution, stand-alone
// activation/dropout. All are sized to stream at HBM rate: 16-byte accesses, grid-stride loops
// with a grid that is a multiple of the SM count.
#include "../include/b200dist.h"
#include "comm.cuh"

namespace b200 {

// ------------------------------------------------------------------------------------------
// softmax cross-entropy: one warp per row when classes <= 1024 (the reference's 10-class head),
// one block per row otherwise (vocabulary-sized heads). Emits loss sum, correct count, dlogits.
// Reference: optax.softmax_cross_entropy_with_integer_labels + argmax accuracy
// (data_paral.py:181-187).
// ------------------------------------------------------------------------------------------
__global__ void xent_warp_kernel(const float* __restrict__ logits, int ld, const int64_t* __restrict__ labels,
                                 int rows, int classes, float grad_scale, __nv_bfloat16* __restrict__ dlogits,
                                 int ldd, float* __restrict__ metrics, float* __restrict__ row_loss,
                                 float* __restrict__ dbias) {
  pdl_sync();
  const int warps_per_block = blockDim.x >> 5;
  const int lane = threadIdx.x & 31;
  float loss_acc = 0.f, correct_acc = 0.f;
  for (int row = blockIdx.x * warps_per_block + (threadIdx.x >> 5); row < rows;
       row += gridDim.x * warps_per_block) {
    const float* x = logits + static_cast<size_t>(row) * ld;
    const int label = static_cast<int>(labels[row]);
    float mx = -INFINITY;
    int amax = 0x7fffffff;
    for (int c = lane; c < classes; c += 32) {
      float v = x[c];
      if (v > mx) { mx = v; amax = c; }
    }
    // warp argmax (first index wins ties, like jnp.argmax)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float om = __shfl_xor_sync(0xffffffffu, mx, o);
      int oa = __shfl_xor_sync(0xffffffffu, amax, o);
      if (om > mx || (om == mx && oa < amax)) { mx = om; amax = oa; }
    }
    float sum = 0.f;
    for (int c = lane; c < classes; c += 32) sum += __expf(x[c] - mx);
    sum = warp_sum(sum);
    const float lse = mx + __logf(sum);
    const float inv = 1.0f / sum;
    if (dlogits != nullptr) {
      __nv_bfloat16* d = dlogits + static_cast<size_t>(row) * ldd;
      for (int c = lane; c < classes; c += 32) {
        float pch = __expf(x[c] - mx) * inv;
        const __nv_bfloat16 gv = __float2bfloat16_rn((pch - (c == label ? 1.0f : 0.0f)) * grad_scale);
        d[c] = gv;
        if (dbias != nullptr) atomicAdd(&dbias[c], __bfloat162float(gv));   // few classes on this path: the head's bias gradient
      }
    }
    if (lane == 0) {
      float l = lse - x[label];
      loss_acc += l;
      correct_acc += (amax == label) ? 1.0f : 0.0f;
      if (row_loss != nullptr) row_loss[row] = l;
    }
  }
  // one atomic pair per block
  __shared__ float s_loss[32], s_corr[32];
  if (lane == 0) { s_loss[threadIdx.x >> 5] = loss_acc; s_corr[threadIdx.x >> 5] = correct_acc; }
  __syncthreads();
  if (threadIdx.x == 0 && metrics != nullptr) {
    float a = 0.f, b = 0.f;
    for (int w = 0; w < warps_per_block; ++w) { a += s_loss[w]; b += s_corr[w]; }
    if (a != 0.f || b != 0.f) { atomicAdd(&metrics[0], a); atomicAdd(&metrics[1], b); }
  }
}

// bf16 halves of a packed word as fp32, two equivalent encodings (shift/mask and byte permute).
// The row lives in registers as packed words across several passes; if two passes unpack with the
// same instructions, ptxas merges them, keeps a whole pass of unpacked floats alive and spills.
__device__ __forceinline__ float bf16_lo_a(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi_a(uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); }
__device__ __forceinline__ float bf16_lo_b(uint32_t w) { return __uint_as_float(__byte_perm(w, 0u, 0x1044)); }
__device__ __forceinline__ float bf16_hi_b(uint32_t w) { return __uint_as_float(__byte_perm(w, 0u, 0x3244)); }
__device__ __forceinline__ uint32_t max_bf16x2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("max.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ float ex2_fast(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int NV>
__global__ void __launch_bounds__(256, 2)
xent_row_bf16_kernel(__nv_bfloat16* __restrict__ logits, int ld, const int64_t* __restrict__ labels, int rows,
                     int classes, float grad_scale, float* __restrict__ metrics, float* __restrict__ row_loss) {
  pdl_sync();
  __shared__ float s_red[8];
  __shared__ int s_idx[8];
  __shared__ float s_bc[2];
  __shared__ int s_bi;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nvec = classes >> 3, tail = classes & 7;
  constexpr float kLog2e = 1.4426950408889634f;
  const float lscale = log2f(grad_scale);
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    __nv_bfloat16* x = logits + static_cast<size_t>(row) * ld;
    const uint4* xv = reinterpret_cast<const uint4*>(x);
    const int label = static_cast<int>(labels[row]);
    uint32_t w[NV][4];
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int i = tid + k * 256;
      const uint4 t = i < nvec ? xv[i] : make_uint4(0xFF80FF80u, 0xFF80FF80u, 0xFF80FF80u, 0xFF80FF80u);   // -inf pairs
      w[k][0] = t.x; w[k][1] = t.y; w[k][2] = t.z; w[k][3] = t.w;
    }
    const float vt = tid < tail ? __bfloat162float(x[nvec * 8 + tid]) : -INFINITY;
    const float x_label = (tid == 0) ? __bfloat162float(x[label]) : 0.f;     // read before the row is overwritten
    // ---- maximum (packed bf16x2) ----
    uint32_t pm0 = w[0][0], pm1 = w[0][1];
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      pm0 = max_bf16x2(pm0, w[k][0]);
      pm1 = max_bf16x2(pm1, w[k][1]);
      pm0 = max_bf16x2(pm0, w[k][2]);
      pm1 = max_bf16x2(pm1, w[k][3]);
    }
    pm0 = max_bf16x2(pm0, pm1);
    const float mloc = fmaxf(fmaxf(__uint_as_float(pm0 << 16), __uint_as_float(pm0 & 0xFFFF0000u)), vt);
    float mx = warp_max(mloc);
    if (lane == 0) s_red[warp] = mx;
    __syncthreads();
    if (warp == 0) {
      float m2 = warp_max(lane < 8 ? s_red[lane] : -INFINITY);
      if (lane == 0) s_bc[0] = m2;
    }
    __syncthreads();
    const float m = s_bc[0];
    // first index of the maximum: only threads that hold it look for it
    int cand = 0x7fffffff;
    const uint32_t mbits = __float_as_uint(m) >> 16;
    if (mloc == m) {
      if (vt == m) cand = nvec * 8 + tid;
#pragma unroll
      for (int k = NV - 1; k >= 0; --k) {
#pragma unroll
        for (int j = 3; j >= 0; --j) {       // integer compares on the bf16 bit patterns (m is one of them)
          if ((w[k][j] >> 16) == mbits) cand = (tid + k * 256) * 8 + 2 * j + 1;
          if ((w[k][j] & 0xFFFFu) == mbits) cand = (tid + k * 256) * 8 + 2 * j;
        }
      }
    }
    // ---- sum of exponentials ----
    const float ml2 = m * kLog2e;
    float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        sum0 += ex2_fast(fmaf(bf16_lo_a(w[k][j]), kLog2e, -ml2));
        sum1 += ex2_fast(fmaf(bf16_hi_a(w[k][j]), kLog2e, -ml2));
      }
    }
    float sum = sum0 + sum1 + (tid < tail ? ex2_fast(fmaf(vt, kLog2e, -ml2)) : 0.f);
    sum = warp_sum(sum);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cand = min(cand, __shfl_xor_sync(0xffffffffu, cand, o));
    if (lane == 0) { s_red[warp] = sum; s_idx[warp] = cand; }
    __syncthreads();
    if (warp == 0) {
      float s2 = warp_sum(lane < 8 ? s_red[lane] : 0.f);
      int a2 = lane < 8 ? s_idx[lane] : 0x7fffffff;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) a2 = min(a2, __shfl_xor_sync(0xffffffffu, a2, o));
      if (lane == 0) { s_bc[1] = s2; s_bi = a2; }
    }
    __syncthreads();
    const float ssum = s_bc[1];
    if (tid == 0) {
      const float l = m + __logf(ssum) - x_label;
      if (row_loss != nullptr) row_loss[row] = l;
      if (metrics != nullptr) {
        atomicAdd(&metrics[0], l);
        if (s_bi == label) atomicAdd(&metrics[1], 1.0f);
      }
    }
    // ---- gradient: softmax * scale, minus scale at the label ----
    const float c = lscale - ml2 - log2f(ssum);
    const int lchunk = label >> 3, lj = label & 7;
    uint4* xo = reinterpret_cast<uint4*>(x);
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int i = tid + k * 256;
      if (i < nvec) {
        float g[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          g[2 * j] = ex2_fast(fmaf(bf16_lo_b(w[k][j]), kLog2e, c));
          g[2 * j + 1] = ex2_fast(fmaf(bf16_hi_b(w[k][j]), kLog2e, c));
        }
        if (i == lchunk) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (j == lj) g[j] -= grad_scale;
        }
        uint4 o;
        __nv_bfloat162 t;
        t = __floats2bfloat162_rn(g[0], g[1]); o.x = *reinterpret_cast<uint32_t*>(&t);
        t = __floats2bfloat162_rn(g[2], g[3]); o.y = *reinterpret_cast<uint32_t*>(&t);
        t = __floats2bfloat162_rn(g[4], g[5]); o.z = *reinterpret_cast<uint32_t*>(&t);
        t = __floats2bfloat162_rn(g[6], g[7]); o.w = *reinterpret_cast<uint32_t*>(&t);
        xo[i] = o;
      }
    }
    if (tid < tail) {
      const int cidx = nvec * 8 + tid;
      x[cidx] = __float2bfloat16_rn(ex2_fast(fmaf(vt, kLog2e, c)) - (cidx == label ? grad_scale : 0.f));
    }
    __syncthreads();     // s_red / s_bc are reused by the next row
  }
}

// Large-vocabulary variant on bf16 logits, in place, any vocabulary size (fallback of the register-resident kernel above):
// logits are replaced by dlogits.
// One block per row; the row is read twice (max+sum online in pass 1, gradient in pass 2);
// the second read hits L2 (a 50k-class bf16 row is 100 KB).
__global__ void __launch_bounds__(512)
xent_block_bf16_kernel(__nv_bfloat16* __restrict__ logits, int ld, const int64_t* __restrict__ labels,
                       int rows, int classes, float grad_scale, float* __restrict__ metrics,
                       float* __restrict__ row_loss) {
  pdl_sync();
  __shared__ float s_m[16], s_s[16];
  __shared__ int s_a[16];
  __shared__ float s_bcast[2];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    __nv_bfloat16* x = logits + static_cast<size_t>(row) * ld;
    const int label = static_cast<int>(labels[row]);
    float mx = -INFINITY, sum = 0.f;
    int amax = 0x7fffffff;
    const bool vec = (ld % 8 == 0) && ((reinterpret_cast<uintptr_t>(logits) & 15) == 0);
    const int nvec = vec ? classes / 8 : 0;
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
      uint4 w = *reinterpret_cast<const uint4*>(x + i * 8);
      const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float v = __uint_as_float(((ws[j >> 1] >> ((j & 1) * 16)) & 0xFFFFu) << 16);
        if (v > mx) { sum = sum * __expf(mx - v) + 1.0f; mx = v; amax = i * 8 + j; }
        else sum += __expf(v - mx);
      }
    }
    for (int c = nvec * 8 + threadIdx.x; c < classes; c += blockDim.x) {
      float v = __bfloat162float(x[c]);
      if (v > mx) { sum = sum * __expf(mx - v) + 1.0f; mx = v; amax = c; }
      else sum += __expf(v - mx);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float om = __shfl_xor_sync(0xffffffffu, mx, o);
      float os = __shfl_xor_sync(0xffffffffu, sum, o);
      int oa = __shfl_xor_sync(0xffffffffu, amax, o);
      float nm = fmaxf(mx, om);
      float ns = (mx == -INFINITY ? 0.f : sum * __expf(mx - nm)) + (om == -INFINITY ? 0.f : os * __expf(om - nm));
      if (om > mx || (om == mx && oa < amax)) amax = oa;
      mx = nm; sum = ns;
    }
    if (lane == 0) { s_m[warp] = mx; s_s[warp] = sum; s_a[warp] = amax; }
    __syncthreads();
    if (warp == 0) {
      float m2 = lane < nwarp ? s_m[lane] : -INFINITY;
      float s2 = lane < nwarp ? s_s[lane] : 0.f;
      int a2 = lane < nwarp ? s_a[lane] : 0x7fffffff;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        float om = __shfl_xor_sync(0xffffffffu, m2, o);
        float os = __shfl_xor_sync(0xffffffffu, s2, o);
        int oa = __shfl_xor_sync(0xffffffffu, a2, o);
        float nm = fmaxf(m2, om);
        float ns = (m2 == -INFINITY ? 0.f : s2 * __expf(m2 - nm)) + (om == -INFINITY ? 0.f : os * __expf(om - nm));
        if (om > m2 || (om == m2 && oa < a2)) a2 = oa;
        m2 = nm; s2 = ns;
      }
      if (lane == 0) {
        s_bcast[0] = m2; s_bcast[1] = s2;
        float l = m2 + __logf(s2) - __bfloat162float(x[label]);
        if (row_loss != nullptr) row_loss[row] = l;
        if (metrics != nullptr) {
          atomicAdd(&metrics[0], l);
          if (a2 == label) atomicAdd(&metrics[1], 1.0f);
        }
      }
    }
    __syncthreads();
    const float m = s_bcast[0];
    const float inv = 1.0f / s_bcast[1];
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
      uint4 w = *reinterpret_cast<const uint4*>(x + i * 8);
      uint32_t ws[4] = {w.x, w.y, w.z, w.w};
      uint32_t out[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float v0 = __uint_as_float((ws[j] & 0xFFFFu) << 16);
        float v1 = __uint_as_float(ws[j] & 0xFFFF0000u);
        int c0 = i * 8 + j * 2;
        float g0 = (__expf(v0 - m) * inv - (c0 == label ? 1.f : 0.f)) * grad_scale;
        float g1 = (__expf(v1 - m) * inv - (c0 + 1 == label ? 1.f : 0.f)) * grad_scale;
        __nv_bfloat162 t = __floats2bfloat162_rn(g0, g1);
        out[j] = *reinterpret_cast<uint32_t*>(&t);
      }
      *reinterpret_cast<uint4*>(x + i * 8) = make_uint4(out[0], out[1], out[2], out[3]);
    }
    for (int c = nvec * 8 + threadIdx.x; c < classes; c += blockDim.x) {
      float v = __bfloat162float(x[c]);
      x[c] = __float2bfloat16_rn((__expf(v - m) * inv - (c == label ? 1.f : 0.f)) * grad_scale);
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------
// column sums (bias gradients). Vector path (ld % 8 == 0): a thread owns 8 consecutive columns
// (one 16-byte load per row), 32 threads cover 256 columns, 8 row-lanes per block walk the
// block's row range; partial sums meet in shared memory and leave with one atomicAdd per column
// per block. Scalar path for narrow / unaligned matrices (the 10-class head).
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
colsum_bf16_vec_kernel(const __nv_bfloat16* __restrict__ x, int ld, int rows, int cols, float* __restrict__ out,
                       int row_chunks, int direct) {
  pdl_sync();
  __shared__ float s[8][256 + 8];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int col = blockIdx.x * 256 + cx * 8;
  const int rows_per = (rows + row_chunks - 1) / row_chunks;
  const int r0 = blockIdx.y * rows_per;
  const int r1 = min(rows, r0 + rows_per);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (col < cols) {   // cols % 8 == 0 on this path, so the 8 columns are all valid
    int r = r0 + ry;
    for (; r + 24 < r1; r += 32) {   // four independent loads in flight
      uint4 w[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) w[u] = *reinterpret_cast<const uint4*>(x + static_cast<size_t>(r + u * 8) * ld + col);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        acc[0] += __uint_as_float(w[u].x << 16); acc[1] += __uint_as_float(w[u].x & 0xFFFF0000u);
        acc[2] += __uint_as_float(w[u].y << 16); acc[3] += __uint_as_float(w[u].y & 0xFFFF0000u);
        acc[4] += __uint_as_float(w[u].z << 16); acc[5] += __uint_as_float(w[u].z & 0xFFFF0000u);
        acc[6] += __uint_as_float(w[u].w << 16); acc[7] += __uint_as_float(w[u].w & 0xFFFF0000u);
      }
    }
    for (; r < r1; r += 8) {
      const uint4 w = *reinterpret_cast<const uint4*>(x + static_cast<size_t>(r) * ld + col);
      acc[0] += __uint_as_float(w.x << 16); acc[1] += __uint_as_float(w.x & 0xFFFF0000u);
      acc[2] += __uint_as_float(w.y << 16); acc[3] += __uint_as_float(w.y & 0xFFFF0000u);
      acc[4] += __uint_as_float(w.z << 16); acc[5] += __uint_as_float(w.z & 0xFFFF0000u);
      acc[6] += __uint_as_float(w.w << 16); acc[7] += __uint_as_float(w.w & 0xFFFF0000u);
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) s[ry][cx * 8 + j] = acc[j];
  __syncthreads();
  const int c = threadIdx.x;   // 256 threads = 256 columns
  if (blockIdx.x * 256 + c < cols) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += s[i][c];
    if (direct) out[blockIdx.x * 256 + c] = t;      // one block per column stripe, no accumulation
    else atomicAdd(&out[blockIdx.x * 256 + c], t);
  }
}

__global__ void colsum_bf16_kernel(const __nv_bfloat16* __restrict__ x, int ld, int rows, int cols,
                                   float* __restrict__ out, int accumulate, int row_chunks) {
  pdl_sync();
  __shared__ float s[8][33];
  const int col = blockIdx.x * 32 + threadIdx.x;
  const int chunk = blockIdx.y;
  const int rows_per = (rows + row_chunks - 1) / row_chunks;
  const int r0 = chunk * rows_per;
  const int r1 = min(rows, r0 + rows_per);
  float acc = 0.f;
  if (col < cols)
    for (int r = r0 + threadIdx.y; r < r1; r += 8)
      acc += __bfloat162float(x[static_cast<size_t>(r) * ld + col]);
  s[threadIdx.y][threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y == 0 && col < cols) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += s[i][threadIdx.x];
    if (row_chunks == 1 && !accumulate) out[col] = t;
    else atomicAdd(&out[col], t);
  }
}

// ------------------------------------------------------------------------------------------
// AdamW, optax.adamw semantics: m,v update; bias correction with t = count+1; eps outside the
// sqrt; decoupled weight decay on every leaf. Fused: gradient scale (mean over
// minibatches / devices), bf16 working copy, optional gradient zeroing for the next step.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void adamw_elem(float& p, float& m, float& v, float g, float lr, float b1,
                                           float b2, float eps, float wd, float c1, float c2) {
  m = b1 * m + (1.0f - b1) * g;
  v = b2 * v + (1.0f - b2) * g * g;
  float mh = m * c1;
  float vh = v * c2;
  p = p - lr * (mh / (sqrtf(vh) + eps) + wd * p);
}

__global__ void adamw_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v,
                             float* __restrict__ g, __nv_bfloat16* __restrict__ pb, int64_t n,
                             b200_adamw_hparams hp, const int32_t* __restrict__ step_count, int zero_grad) {
  pdl_sync();
  const float t = static_cast<float>(*step_count + 1);
  const float c1 = 1.0f / (1.0f - __powf(hp.beta1, t));
  const float c2 = 1.0f / (1.0f - __powf(hp.beta2, t));
  const int64_t n4 = n >> 2;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float4 gg = reinterpret_cast<const float4*>(g)[i];
    adamw_elem(pp.x, mm.x, vv.x, gg.x * hp.grad_scale, hp.lr, hp.beta1, hp.beta2, hp.eps, hp.weight_decay, c1, c2);
    adamw_elem(pp.y, mm.y, vv.y, gg.y * hp.grad_scale, hp.lr, hp.beta1, hp.beta2, hp.eps, hp.weight_decay, c1, c2);
    adamw_elem(pp.z, mm.z, vv.z, gg.z * hp.grad_scale, hp.lr, hp.beta1, hp.beta2, hp.eps, hp.weight_decay, c1, c2);
    adamw_elem(pp.w, mm.w, vv.w, gg.w * hp.grad_scale, hp.lr, hp.beta1, hp.beta2, hp.eps, hp.weight_decay, c1, c2);
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
    if (zero_grad) reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (pb != nullptr) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(pp.x, pp.y);
      __nv_bfloat162 hi = __floats2bfloat162_rn(pp.z, pp.w);
      uint2 w;
      w.x = *reinterpret_cast<uint32_t*>(&lo);
      w.y = *reinterpret_cast<uint32_t*>(&hi);
      reinterpret_cast<uint2*>(pb)[i] = w;
    }
  }
  // tail
  for (int64_t i = (n4 << 2) + static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    float pp = p[i], mm = m[i], vv = v[i];
    adamw_elem(pp, mm, vv, g[i] * hp.grad_scale, hp.lr, hp.beta1, hp.beta2, hp.eps, hp.weight_decay, c1, c2);
    p[i] = pp; m[i] = mm; v[i] = vv;
    if (zero_grad) g[i] = 0.f;
    if (pb != nullptr) pb[i] = __float2bfloat16_rn(pp);
  }
}

__global__ void bump_step_kernel(int32_t* step) {
  pdl_sync(); *step += 1; }

__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t n) {
  pdl_sync();
  const int64_t n4 = n >> 2;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 a = reinterpret_cast<const float4*>(src)[i];
    __nv_bfloat162 lo = __floats2bfloat162_rn(a.x, a.y), hi = __floats2bfloat162_rn(a.z, a.w);
    uint2 w;
    w.x = *reinterpret_cast<uint32_t*>(&lo);
    w.y = *reinterpret_cast<uint32_t*>(&hi);
    reinterpret_cast<uint2*>(dst)[i] = w;
  }
  for (int64_t i = (n4 << 2) + static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    dst[i] = __float2bfloat16_rn(src[i]);
}

__global__ void cast_bf16_f32_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst, int64_t n) {
  pdl_sync();
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    dst[i] = __bfloat162float(src[i]);
}

__global__ void scale_f32_kernel(float* __restrict__ x, int64_t n, float s) {
  pdl_sync();
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    x[i] = s == 0.f ? 0.f : x[i] * s;
}

__global__ void rng_split_kernel(uint64_t* key, uint64_t* out_step, uint64_t salt_next, uint64_t salt_step,
                                 int32_t* step_count) {
  pdl_sync();
  uint64_t k = *key;
  if (out_step != nullptr) *out_step = mix64(k ^ salt_step);
  *key = mix64(k ^ salt_next);
  if (step_count != nullptr) *step_count += 1;   // end-of-step bookkeeping rides along
}

// Start-of-step housekeeping in one launch: fp32 -> bf16 cast of the inputs and zeroing of up to two
// small fp32 regions (step metrics, bias-gradient vectors).
__global__ void cast_zero_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t n,
                                 float* __restrict__ z0, int64_t n0, float* __restrict__ z1, int64_t n1,
                                 const uint32_t* __restrict__ sig, int rank, int world, uint64_t timeout_ns,
                                 const SignalPeers peers, int deferred) {
  pdl_sync();
  // First kernel of a step: the peers must have finished the previous step's update (they publish
  // into our parameters and read our gradients) before anything here clears a gradient region or a
  // later kernel reads a weight. See comm.cuh.
  if (sig != nullptr) {
    if (deferred) publish_and_wait(sig, peers, rank, world, timeout_ns);
    else wait_published(sig, rank, world, timeout_ns);
  }
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t n4 = n >> 2;
  for (int64_t i = tid; i < n4; i += stride) {
    float4 a = reinterpret_cast<const float4*>(src)[i];
    __nv_bfloat162 lo = __floats2bfloat162_rn(a.x, a.y), hi = __floats2bfloat162_rn(a.z, a.w);
    uint2 w;
    w.x = *reinterpret_cast<uint32_t*>(&lo);
    w.y = *reinterpret_cast<uint32_t*>(&hi);
    reinterpret_cast<uint2*>(dst)[i] = w;
  }
  for (int64_t i = (n4 << 2) + tid; i < n; i += stride) dst[i] = __float2bfloat16_rn(src[i]);
  for (int64_t i = tid; i < n0; i += stride) z0[i] = 0.f;
  for (int64_t i = tid; i < n1; i += stride) z1[i] = 0.f;
}

// stand-alone activation + dropout (rows x cols, pitch ld)
__global__ void act_dropout_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                                   __nv_bfloat16* __restrict__ out, int64_t rows, int cols, int ld, int act,
                                   const uint64_t* __restrict__ key, uint64_t salt, uint32_t thresh16, float scale,
                                   int bwd) {
  pdl_sync();
  uint64_t seed = 0;
  if (key != nullptr) seed = mix64(*key ^ salt);
  const int64_t total = rows * cols;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    int64_t r = i / cols;
    int c = static_cast<int>(i - r * cols);
    size_t off = static_cast<size_t>(r) * ld + c;
    float xv = __bfloat162float(x[off]);
    float v;
    if (!bwd) {
      v = act == B200_ACT_SILU ? silu_f(xv) : act == B200_ACT_GELU ? gelu_f(xv) : xv;
    } else {
      float gy = __bfloat162float(dy[off]);
      float d = act == B200_ACT_SILU ? silu_grad_f(xv) : act == B200_ACT_GELU ? gelu_grad_f(xv) : 1.0f;
      v = gy * d;
    }
    if (key != nullptr) v = dropout_keep(seed, static_cast<uint32_t>(r), static_cast<uint32_t>(c), thresh16) ? v * scale : 0.f;
    out[off] = __float2bfloat16_rn(v);
  }
}

static inline int stream_grid(int64_t work_items, int threads) {
  int64_t blocks = (work_items + threads - 1) / threads;
  int64_t cap = static_cast<int64_t>(num_sms()) * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

}  // namespace b200

using namespace b200;

extern "C" int b200_softmax_xent_dbias(const float* logits, int ld_logits, const int64_t* labels, int rows,
                                       int classes, float grad_scale, void* dlogits, int ld_dlogits,
                                       float* metrics, float* row_loss, float* dbias, void* stream);

extern "C" int b200_softmax_xent(const float* logits, int ld_logits, const int64_t* labels, int rows,
                                 int classes, float grad_scale, void* dlogits, int ld_dlogits,
                                 float* metrics, float* row_loss, void* stream) {
  return b200_softmax_xent_dbias(logits, ld_logits, labels, rows, classes, grad_scale, dlogits, ld_dlogits, metrics,
                                 row_loss, nullptr, stream);
}

extern "C" int b200_softmax_xent_dbias(const float* logits, int ld_logits, const int64_t* labels, int rows,
                                       int classes, float grad_scale, void* dlogits, int ld_dlogits,
                                       float* metrics, float* row_loss, float* dbias, void* stream) {
  B200_REQUIRE(rows > 0 && classes > 0, "xent: empty input");
  const int threads = 128;
  int blocks = (rows + 3) / 4;
  int cap = num_sms() * 16;
  if (blocks > cap) blocks = cap;
  B200_CUDA(launch_k(xent_warp_kernel, dim3(blocks), dim3(threads), 0, static_cast<cudaStream_t>(stream), 
      logits, ld_logits, labels, rows, classes, grad_scale, static_cast<__nv_bfloat16*>(dlogits),
      ld_dlogits, metrics, row_loss, dbias));
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_softmax_xent_bf16_inplace(void* logits_bf16, int ld, const int64_t* labels, int rows,
                                              int classes, float grad_scale, float* metrics, float* row_loss,
                                              void* stream) {
  B200_REQUIRE(rows > 0 && classes > 0, "xent: empty input");
  const bool vec_ok = (ld % 8 == 0) && ((reinterpret_cast<uintptr_t>(logits_bf16) & 15) == 0) && grad_scale > 0.f;
  if (vec_ok && classes >= 4096 && classes <= 256 * 8 * 25) {      // row fits the block's registers
    int rblocks = rows < num_sms() * 2 ? rows : num_sms() * 2;
    B200_CUDA(launch_k(xent_row_bf16_kernel<25>, dim3(rblocks), dim3(256), 0, static_cast<cudaStream_t>(stream),
        static_cast<__nv_bfloat16*>(logits_bf16), ld, labels, rows, classes, grad_scale, metrics, row_loss));
    B200_CUDA(cudaGetLastError());
    return 0;
  }
  int blocks = rows < num_sms() * 4 ? rows : num_sms() * 4;
  B200_CUDA(launch_k(xent_block_bf16_kernel, dim3(blocks), dim3(512), 0, static_cast<cudaStream_t>(stream), 
      static_cast<__nv_bfloat16*>(logits_bf16), ld, labels, rows, classes, grad_scale, metrics, row_loss));
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_colsum_bf16(const void* x, int ld, int rows, int cols, float* out, int accumulate,
                                void* stream) {
  B200_REQUIRE(rows > 0 && cols > 0, "colsum: empty input");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const bool vec = (ld % 8 == 0) && (cols % 8 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0) && cols >= 64;
  if (vec) {
    int col_blocks = (cols + 255) / 256;
    int row_chunks = 1;
    while (col_blocks * row_chunks < num_sms() * 2 && rows / (row_chunks * 2) >= 32) row_chunks *= 2;
    const int direct = (row_chunks == 1 && !accumulate) ? 1 : 0;
    if (!accumulate && !direct) B200_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * cols, s));
    B200_CUDA(launch_k(colsum_bf16_vec_kernel, dim3(col_blocks, row_chunks), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(x), ld, rows,
                                                                      cols, out, row_chunks, direct));
  } else {
    int col_blocks = (cols + 31) / 32;
    int row_chunks = 1;
    while (col_blocks * row_chunks < num_sms() * 2 && rows / (row_chunks * 2) >= 64) row_chunks *= 2;
    if (row_chunks > 1 && !accumulate) B200_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * cols, s));
    B200_CUDA(launch_k(colsum_bf16_kernel, dim3(col_blocks, row_chunks), dim3(32, 8), 0, s, 
        static_cast<const __nv_bfloat16*>(x), ld, rows, cols, out, accumulate, row_chunks));
  }
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_adamw(float* p, float* m, float* v, const float* g, void* p_bf16, int64_t n,
                          const b200_adamw_hparams* hp, int32_t* step_count, int bump_step, int zero_grad,
                          void* stream) {
  B200_REQUIRE(n > 0, "adamw: empty buffer");
  B200_REQUIRE(((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(m) |
                 reinterpret_cast<uintptr_t>(v) | reinterpret_cast<uintptr_t>(g)) & 15) == 0,
               "adamw: buffers must be 16-byte aligned");
  B200_REQUIRE(p_bf16 == nullptr || (reinterpret_cast<uintptr_t>(p_bf16) & 7) == 0,
               "adamw: bf16 copy must be 8-byte aligned");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  B200_CUDA(launch_k(adamw_kernel, dim3(stream_grid(n / 4 + 1, 256)), dim3(256), 0, s, p, m, v, const_cast<float*>(g),
                                                          static_cast<__nv_bfloat16*>(p_bf16), n, *hp,
                                                          step_count, zero_grad));
  if (bump_step) {
    B200_CUDA(launch_k(bump_step_kernel, dim3(1), dim3(1), 0, s, step_count));
  }
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_cast_f32_bf16(const float* src, void* dst, int64_t n, void* stream) {
  B200_REQUIRE(n > 0, "cast: empty buffer");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 7) == 0,
               "cast: misaligned buffers");
  B200_CUDA(launch_k(cast_f32_bf16_kernel, dim3(stream_grid(n / 4 + 1, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), 
      src, static_cast<__nv_bfloat16*>(dst), n));
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_cast_bf16_f32(const void* src, float* dst, int64_t n, void* stream) {
  B200_REQUIRE(n > 0, "cast: empty buffer");
  B200_CUDA(launch_k(cast_bf16_f32_kernel, dim3(stream_grid(n, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), 
      static_cast<const __nv_bfloat16*>(src), dst, n));
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_scale_f32(float* x, int64_t n, float scale, void* stream) {
  B200_REQUIRE(n > 0, "scale: empty buffer");
  B200_CUDA(launch_k(scale_f32_kernel, dim3(stream_grid(n, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), x, n, scale));
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_rng_split_bump(uint64_t* key, uint64_t* out_step, uint64_t salt_next, uint64_t salt_step,
                                   int32_t* step_count, void* stream) {
  B200_CUDA(launch_k(rng_split_kernel, dim3(1), dim3(1), 0, static_cast<cudaStream_t>(stream), key, out_step, salt_next,
                     salt_step, step_count));
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_cast_zero(const float* src, void* dst, int64_t n, float* z0, int64_t n0, float* z1, int64_t n1,
                              void* stream) {
  return b200_cast_zero_wait(src, dst, n, z0, n0, z1, n1, nullptr, nullptr, 0, 1, stream);
}

extern "C" int b200_cast_zero_wait(const float* src, void* dst, int64_t n, float* z0, int64_t n0, float* z1, int64_t n1,
                                   const void* signal_local, void* const* peer_signal, int rank, int world,
                              void* stream) {
  B200_REQUIRE(n >= 0 && n0 >= 0 && n1 >= 0, "cast_zero: negative size");
  B200_REQUIRE(n == 0 || ((reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 7) == 0),
               "cast_zero: misaligned buffers");
  int64_t work = n / 4 + 1;
  if (n0 > work) work = n0;
  if (n1 > work) work = n1;
  SignalPeers peers;
  memset(&peers, 0, sizeof(peers));
  const bool deferred = world > 1 && signal_local != nullptr && comm_defer_publish();
  if (deferred) {
    B200_REQUIRE(peer_signal != nullptr && world <= MAX_WORLD, "cast_zero_wait: deferred publish needs the peers' signal buffers");
    for (int i = 0; i < world; ++i) peers.p[i] = static_cast<uint32_t*>(peer_signal[i]);
  }
  B200_CUDA(launch_k(cast_zero_kernel, dim3(stream_grid(work, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), src,
                     static_cast<__nv_bfloat16*>(dst), n, z0, n0, z1, n1,
                     world > 1 ? static_cast<const uint32_t*>(signal_local) : nullptr, rank, world, comm_timeout_ns(),
                     peers, deferred ? 1 : 0));
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_rng_split(uint64_t* key, uint64_t* out_step, uint64_t salt_next, uint64_t salt_step,
                              void* stream) {
  B200_CUDA(launch_k(rng_split_kernel, dim3(1), dim3(1), 0, static_cast<cudaStream_t>(stream), key, out_step, salt_next, salt_step,
                     static_cast<int32_t*>(nullptr)));
  B200_CUDA(cudaGetLastError());
  return 0;
}

static int act_dropout_launch(const void* x, const void* dy, void* out, int64_t rows, int cols, int ld,
                              int act, const uint64_t* key, uint64_t salt, float rate, int bwd, void* stream) {
  B200_REQUIRE(rows > 0 && cols > 0, "act_dropout: empty input");
  float keep = 1.0f - rate;
  const uint64_t* k = rate > 0.f ? key : nullptr;
  B200_CUDA(launch_k(act_dropout_kernel, dim3(stream_grid(rows * cols, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), 
      static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(dy),
      static_cast<__nv_bfloat16*>(out), rows, cols, ld, act, k, salt,
      static_cast<uint32_t>(keep * 65536.0f + 0.5f), keep > 0.f ? 1.0f / keep : 0.f, bwd));
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200_act_dropout_fwd(const void* x, void* y, int64_t rows, int cols, int ld, int act,
                                    const uint64_t* drop_key, uint64_t drop_salt, float drop_rate,
                                    void* stream) {
  return act_dropout_launch(x, nullptr, y, rows, cols, ld, act, drop_key, drop_salt, drop_rate, 0, stream);
}
extern "C" int b200_act_dropout_bwd(const void* x, const void* dy, void* dx, int64_t rows, int cols, int ld,
                                    int act, const uint64_t* drop_key, uint64_t drop_salt, float drop_rate,
                                    void* stream) {
  return act_dropout_launch(x, dy, dx, rows, cols, ld, act, drop_key, drop_salt, drop_rate, 1, stream);
}
