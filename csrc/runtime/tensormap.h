// Host-side construction of TMA tensor maps (cuTensorMapEncodeTiled fetched at run time through
// cudaGetDriverEntryPoint, so the library does not link against libcuda).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>

namespace ddpb {

// NHWC bf16 activation tensor seen as 4-D (C, W, H, N); box = (box_c, box_w, box_h, box_n) elements,
// 128-byte swizzle (box_c must be 64), out-of-bounds elements read as zero.
int make_tmap_nhwc_bf16(CUtensorMap* out, const void* ptr, int N, int H, int W, int C, int box_c, int box_w,
                        int box_h, int box_n, int c_pitch = 0);      // c_pitch: elements between pixels (0 = C)

// Row-major bf16 matrix [rows][cols]; box = box_cols (must be 64) x box_rows, 128-byte swizzle.
int make_tmap_2d_bf16(CUtensorMap* out, const void* ptr, long long rows, long long cols, int box_cols,
                      int box_rows);

// Row-major fp32 matrix [rows][cols]; box = 32 columns (128 B) x box_rows, 128-byte swizzle.
// atom32: the "128-byte swizzle with 32-byte atoms" pattern (32-byte chunks XORed with the row, period 4 rows),
// the only layout tcgen05.mma accepts for MN-major 32-bit (tf32) operands.
int make_tmap_2d_f32(CUtensorMap* out, const void* ptr, long long rows, long long cols, int box_cols,
                     int box_rows, bool atom32 = false);

// NHWC fp32 activation tensor seen as 4-D (C, W, H, N); box_c must be 32 (128 bytes), 128-byte swizzle.
int make_tmap_nhwc_f32(CUtensorMap* out, const void* ptr, int N, int H, int W, int C, int box_c, int box_w, int box_h,
                       int box_n, bool atom32 = false);

}  // namespace ddpb
