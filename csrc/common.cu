// Host-side runtime pieces shared by every kernel file: error strings, launch counter,
// tensor-map encoding via the driver entry point, device queries.
#include "common.cuh"

#include <stdarg.h>
#include <stdlib.h>

#include <atomic>
#include <mutex>

#include "../include/b200dist.h"

namespace b200 {

static thread_local char g_err[1024] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
  return 1;
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

static std::atomic<int> g_pdl_override{-1};     // -1: follow the environment, 0 / 1: forced by b200_set_pdl

bool pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200DIST_NO_PDL");
    v = (e != nullptr && e[0] == '1') ? 0 : 1;
  }
  const int o = g_pdl_override.load(std::memory_order_relaxed);
  return v == 1 && o != 0;      // the environment switch always wins when it says "off"
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

int make_tmap_2d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t inner, uint64_t outer,
                 uint64_t outer_stride_bytes, uint32_t box_inner, uint32_t box_outer,
                 int swizzle_bytes) {
  EncodeTiledFn enc = get_encode();
  B200_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available (no CUDA driver?)");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0,
               "TMA operand base %p is not 16-byte aligned", base);
  B200_REQUIRE((outer_stride_bytes & 15) == 0,
               "TMA operand row pitch %llu bytes is not a multiple of 16",
               (unsigned long long)outer_stride_bytes);
  CUtensorMapDataType dt =
      elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {outer_stride_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMapSwizzle sw = swizzle_bytes == 128  ? CU_TENSOR_MAP_SWIZZLE_128B
                          : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                          : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                                                : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUresult r = enc(out, dt, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS,
               "cuTensorMapEncodeTiled failed (%d): base=%p inner=%llu outer=%llu pitch=%llu box=%ux%u",
               (int)r, base, (unsigned long long)inner, (unsigned long long)outer,
               (unsigned long long)outer_stride_bytes, box_inner, box_outer);
  return 0;
}

int num_sms() {
  static int n[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (n[dev] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0)
      v = 148;
    n[dev] = v;
  }
  return n[dev];
}

}  // namespace b200

extern "C" {
const char* b200_last_error(void) { return b200::g_err; }
long long b200_launch_count(void) { return b200::g_launches.load(); }
void b200_reset_launch_count(void) { b200::g_launches.store(0); }
int b200_abi_version(void) { return B200_ABI_VERSION; }
void b200_set_pdl(int mode) { b200::g_pdl_override.store(mode < 0 ? -1 : (mode ? 1 : 0)); }
}

#ifndef B200_SOURCE_HASH
#define B200_SOURCE_HASH "unstamped"
#endif
extern "C" const char* b200_source_hash(void) { return B200_SOURCE_HASH; }
