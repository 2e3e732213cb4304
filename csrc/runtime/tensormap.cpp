#include "tensormap.h"

#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "../common.cuh"

namespace ddpb {

static thread_local char g_last_error[1024] = "";

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
  va_end(ap);
}

const char* last_error() { return g_last_error; }

// programmatic dependent launch switch: environment DDPB_PDL (default on), overridable at run time
static int g_pdl = -1;
bool pdl_enabled() {
  if (g_pdl < 0) {
    const char* e = getenv("DDPB_PDL");
    g_pdl = (e && e[0] == '0') ? 0 : 1;
  }
  return g_pdl == 1;
}
int pdl_mask() {
  static int m = -2;
  if (m == -2) {
    const char* e = getenv("DDPB_PDL_MASK");
    m = e ? atoi(e) : -1;
  }
  return m;
}

static int g_ktrace_on = 0, g_ktrace_id = 0;
void ktrace_enable(int on) { g_ktrace_on = on ? 1 : 0; g_ktrace_id = 0; }
int ktrace_last_id() { return g_ktrace_id; }
int pdl_late_mask() {
  static int m = -2;
  if (m == -2) {
    const char* e = getenv("DDPB_PDL_LATE");
    m = e ? atoi(e) : DDPB_PDL_LATE_DEFAULT;
  }
  return m;
}

int kernel_flags(cudaStream_t stream, int family, int late_family) {
  int pdl = pdl_active(stream, family) ? 1 : 0;
  if (pdl && (pdl_late_mask() & (late_family ? late_family : family))) pdl = 2;       // trigger the successor near the end of the CTA's work
  if (!g_ktrace_on) return pdl;
  return pdl | (++g_ktrace_id << 2);
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { (void)cudaGetLastError(); return 148; }
  if (!cached[dev]) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) {
      (void)cudaGetLastError();
      n = 148;
    }
    cached[dev] = n;
  }
  return cached[dev];
}

bool first_on_device(unsigned long long* mask) {
  static std::mutex mu;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) { (void)cudaGetLastError(); dev = 0; }
  const unsigned long long bit = 1ull << (dev & 63);
  std::lock_guard<std::mutex> lock(mu);
  if (*mask & bit) return false;
  *mask |= bit;
  return true;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

static int encode(CUtensorMap* out, const void* ptr, int rank, const cuuint64_t* dims, const cuuint64_t* strides,
                  const cuuint32_t* box, CUtensorMapDataType dtype = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16,
                  CUtensorMapSwizzle swizzle = CU_TENSOR_MAP_SWIZZLE_128B) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_last_error("cuTensorMapEncodeTiled is not available from this driver");
    return DDPB_ERR_UNSUPPORTED;
  }
  cuuint32_t ones[5] = {1, 1, 1, 1, 1};
  CUresult r = fn(out, dtype, rank, const_cast<void*>(ptr), dims, strides, box, ones,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %d, dims %llu %llu %llu %llu, box %u %u %u %u)",
                   (int)r, rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
                   (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 3 ? dims[3] : 0),
                   box[0], rank > 1 ? box[1] : 0, rank > 2 ? box[2] : 0, rank > 3 ? box[3] : 0);
    return DDPB_ERR_CUDA;
  }
  return DDPB_OK;
}

int make_tmap_nhwc_bf16(CUtensorMap* out, const void* ptr, int N, int H, int W, int C, int box_c, int box_w,
                        int box_h, int box_n, int c_pitch) {
  if (c_pitch <= 0) c_pitch = C;          // c_pitch > C: a channel range [ptr .. ptr + C) of a wider tensor
  if (box_c != 64 || C % 64 != 0 || c_pitch % 8 != 0 || c_pitch < C) {
    set_last_error("tensor map: channel box must be 64 and C a multiple of 64 (C=%d box=%d pitch=%d)", C, box_c, c_pitch);
    return DDPB_ERR_ARG;
  }
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)c_pitch * 2, (cuuint64_t)W * c_pitch * 2, (cuuint64_t)H * W * c_pitch * 2};
  cuuint32_t box[4] = {(cuuint32_t)box_c, (cuuint32_t)box_w, (cuuint32_t)box_h, (cuuint32_t)box_n};
  return encode(out, ptr, 4, dims, strides, box);
}

int make_tmap_nhwc_f32(CUtensorMap* out, const void* ptr, int N, int H, int W, int C, int box_c, int box_w, int box_h,
                       int box_n, bool atom32) {
  if (box_c != 32 || C % 32 != 0) {
    set_last_error("tensor map: fp32 channel box must be 32 and C a multiple of 32 (C=%d box=%d)", C, box_c);
    return DDPB_ERR_ARG;
  }
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)C * 4, (cuuint64_t)W * C * 4, (cuuint64_t)H * W * C * 4};
  cuuint32_t box[4] = {(cuuint32_t)box_c, (cuuint32_t)box_w, (cuuint32_t)box_h, (cuuint32_t)box_n};
  return encode(out, ptr, 4, dims, strides, box, CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
                atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B);
}

int make_tmap_2d_bf16(CUtensorMap* out, const void* ptr, long long rows, long long cols, int box_cols,
                      int box_rows) {
  if (box_cols != 64 || cols % 8 != 0) {
    set_last_error("tensor map: 2-D box must be 64 columns wide and the row pitch 16-byte aligned");
    return DDPB_ERR_ARG;
  }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  return encode(out, ptr, 2, dims, strides, box);
}

int make_tmap_2d_f32(CUtensorMap* out, const void* ptr, long long rows, long long cols, int box_cols,
                     int box_rows, bool atom32) {
  if (box_cols != 32 || cols % 4 != 0) {
    set_last_error("tensor map: fp32 2-D box must be 32 columns (128 bytes) wide, row pitch 16-byte aligned");
    return DDPB_ERR_ARG;
  }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols * 4};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  return encode(out, ptr, 2, dims, strides, box, CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
                atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B);
}

}  // namespace ddpb

extern "C" const char* ddpb_last_error() { return ddpb::last_error(); }

// 1 / 0 = allow / forbid programmatic dependent launches from now on; returns the previous setting.
extern "C" int ddpb_set_pdl(int on) {
  const int prev = ddpb::pdl_enabled() ? 1 : 0;
  ddpb::g_pdl = on ? 1 : 0;
  return prev;
}
