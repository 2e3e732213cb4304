// y[P, 64k] = x[P, 64] * w[Cout, 64]^T + bias, bf16 in / out, fp32 accumulation in TMEM, plus the
// per-channel sum / sum-of-squares of y for BatchNorm.
//
// This is the forward of the first VGG layer (ref Part 1/model.py:18-23, the 3 -> 64 convolution) after
// ddpb_im2col3x3_c3 has laid its 3x3x3 patches out as a [pixels][64] matrix.  K is a single 64-wide
// chunk, so a "one CTA per 128-row tile" kernel is nothing but prologue and epilogue (2048 CTAs, 33 us
// through the generic conv kernel).  Here one CTA walks MT consecutive row tiles:
//   warp 0     TMA: the weight tile once, then one 16 KB activation tile per row tile (ring of 3)
//   warp 1     one thread: 4 x tcgen05.mma (128 x 64 x 16) per row tile into one of TWO TMEM
//              accumulators, so the MMAs of tile i+1 run while tile i is in the epilogue
//   warps 2-5  epilogue per tile: TMEM -> +bias -> bf16 -> swizzled shared-memory tile -> coalesced 16-byte
//              stores; BatchNorm statistics stay in registers across the tiles and cost one atomicAdd
//              per channel and CTA at the very end
#include "../common.cuh"
#include "../../include/ddp_b200.h"
#include "../runtime/tensormap.h"

namespace ddpb {

constexpr int GR_N = 64;                 // output channels per CTA
constexpr int GR_A = 128 * 128;          // activation tile: 128 rows x 64 bf16
constexpr int GR_B = GR_N * 128;         // weight tile: 64 rows x 64 bf16
constexpr int GR_STAGES = 3;          // 3 x 16 KB + tiles: ~98 KB, two CTAs per SM
constexpr int GR_TILE = 128 * GR_N * 2;  // bf16 output tile

struct GemmRowsParams {
  long long rows;          // P
  int n_out;               // Cout (row pitch of y)
  int tiles;               // ceil(P / 128)
  int mt;                  // row tiles per CTA
  int pdl;                 // 1: execute griddepcontrol.launch_dependents (captured launches only)
  bf16* y;
};

__global__ void __launch_bounds__(192)
gemm_rows_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const float* __restrict__ bias, float* __restrict__ stats, const GemmRowsParams prm) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sB = smem + GR_STAGES * GR_A;
  uint8_t* tile = sB + GR_B;                                     // [2][128 rows][64] bf16 (double-buffered)
  float* red = reinterpret_cast<float*>(tile + 2 * GR_TILE);      // [16][2 * 64]
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(red + 16 * 2 * GR_N);
  uint64_t* empty_bar = full_bar + GR_STAGES;
  uint64_t* b_bar = empty_bar + GR_STAGES;
  uint64_t* acc_full = b_bar + 1;         // [2]
  uint64_t* acc_empty = acc_full + 2;     // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  float* bias_s = reinterpret_cast<float*>(tmem_slot + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tile = blockIdx.y;
  const int tile0 = blockIdx.x * prm.mt;
  const int nsub = min(prm.mt, prm.tiles - tile0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int i = 0; i < GR_STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    mbar_init(b_bar, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 4); }   // 4 epilogue warps
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<2 * GR_N>(tmem_slot);
  if (prm.pdl & 1) pdl_launch_dependents();     // programmatic dependent launch: see common.cuh
  pdl_wait();
  ktrace_begin(prm.pdl);      // stamped once the predecessor has finished: the start of this kernel's own work
  for (int i = threadIdx.x; i < GR_N; i += blockDim.x) bias_s[i] = bias ? ld_coh(&bias[n_tile * GR_N + i]) : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(b_bar, GR_B);
      tma_load_2d(sB, &tmB, b_bar, 0, n_tile * GR_N);
      for (int sub = 0; sub < nsub; ++sub) {
        const int stage = sub % GR_STAGES, use = sub / GR_STAGES;
        if (use > 0) mbar_wait(&empty_bar[stage], (use - 1) & 1);
        mbar_expect_tx(&full_bar[stage], GR_A);
        tma_load_2d(smem + stage * GR_A, &tmA, &full_bar[stage], 0, (tile0 + sub) * 128);   // rows past P read as zero
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, GR_N, false, false);
      mbar_wait(b_bar, 0);
      const uint32_t sb = smem_u32(sB);
      for (int sub = 0; sub < nsub; ++sub) {
        const int stage = sub % GR_STAGES, a = sub & 1;
        if (sub >= 2) mbar_wait(&acc_empty[a], ((sub >> 1) - 1) & 1);     // the epilogue has drained this accumulator
        mbar_wait(&full_bar[stage], (sub / GR_STAGES) & 1);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * GR_A);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          umma_bf16(tmem_base + a * GR_N, umma_desc(sa + kk * 32, 16, 1024), umma_desc(sb + kk * 32, 16, 1024), idesc, kk != 0);
        umma_commit(&empty_bar[stage]);
        umma_commit(&acc_full[a]);
      }
    }
  } else {
    const int q = warp & 3, row = q * 32 + lane, t = threadIdx.x - 64;
    const int ch = t & 7, part = t >> 3;              // 8 chunks of 8 channels per row, 16 row groups
    const int col0 = n_tile * GR_N + ch * 8;
    float s1[8], s2[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { s1[j] = 0.f; s2[j] = 0.f; }
    for (int sub = 0; sub < nsub; ++sub) {
      const int a = sub & 1;
      uint8_t* tl = tile + a * GR_TILE;
      mbar_wait(&acc_full[a], (sub >> 1) & 1);
      tc_fence_after();
      // 1. accumulator -> bf16 tile (row per thread, 16-byte chunks XOR-swizzled by the row)
#pragma unroll
      for (int c = 0; c < GR_N; c += 32) {
        float v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + a * GR_N + c, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int jc = (c >> 3) + j;
          *reinterpret_cast<uint4*>(tl + row * 128 + ((jc ^ (row & 7)) << 4)) =
              make_uint4(pack_bf16x2(v[8 * j] + bias_s[c + 8 * j], v[8 * j + 1] + bias_s[c + 8 * j + 1]),
                         pack_bf16x2(v[8 * j + 2] + bias_s[c + 8 * j + 2], v[8 * j + 3] + bias_s[c + 8 * j + 3]),
                         pack_bf16x2(v[8 * j + 4] + bias_s[c + 8 * j + 4], v[8 * j + 5] + bias_s[c + 8 * j + 5]),
                         pack_bf16x2(v[8 * j + 6] + bias_s[c + 8 * j + 6], v[8 * j + 7] + bias_s[c + 8 * j + 7]));
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[a]);                 // this warp's quarter of the accumulator is free
      asm volatile("bar.sync 1, 128;" ::: "memory");
      // 2. coalesced store (8 consecutive threads = one 128-byte row) + statistics of the rounded values
      const long long row0 = (long long)(tile0 + sub) * 128;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int r = part + k * 16;
        const long long grow = row0 + r;
        if (grow < prm.rows) {
          const uint4 u = *reinterpret_cast<const uint4*>(tl + r * 128 + ((ch ^ (r & 7)) << 4));
          *reinterpret_cast<uint4*>(prm.y + grow * prm.n_out + col0) = u;
          const float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y), f2 = unpack_bf16x2(u.z), f3 = unpack_bf16x2(u.w);
          const float d[8] = {f0.x, f0.y, f1.x, f1.y, f2.x, f2.y, f3.x, f3.y};
#pragma unroll
          for (int j = 0; j < 8; ++j) { s1[j] += d[j]; s2[j] = fmaf(d[j], d[j], s2[j]); }
        }
      }
      // the tile buffer `a` is written again two tiles later, after the bar.sync of the tile in between
    }
    if (stats) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        red[part * 2 * GR_N + ch * 8 + j] = s1[j];
        red[part * 2 * GR_N + GR_N + ch * 8 + j] = s2[j];
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      float acc = 0.f;
#pragma unroll
      for (int pz = 0; pz < 16; ++pz) acc += red[pz * 2 * GR_N + t];
      atomicAdd(&stats[(t < GR_N ? 0 : prm.n_out) + n_tile * GR_N + (t & (GR_N - 1))], acc);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<2 * GR_N>(tmem_base);
  ktrace_end(prm.pdl);
}

}  // namespace ddpb
DDPB_KTRACE_SETTER(gemm_rows_tc)

using namespace ddpb;

// x bf16 [P][64], w bf16 [Cout][64] (Cout % 64 == 0), y bf16 [P][Cout]; stats (nullable, fp32 [2*Cout]) accumulates.
extern "C" int ddpb_gemm_rows_fwd_tc(const void* x, const void* w_bf16, const float* bias, void* y, float* stats,
                                     long long P, int Cout, cudaStream_t stream) {
  DDPB_REQUIRE(P > 0 && Cout > 0 && Cout % GR_N == 0, "gemm_rows: bad shape (P %lld, Cout %d)", P, Cout);
  CUtensorMap tmA, tmB;
  int rc = make_tmap_2d_bf16(&tmA, x, P, 64, 64, 128);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tmB, w_bf16, Cout, 64, 64, GR_N);
  if (rc) return rc;
  GemmRowsParams prm;
  prm.rows = P; prm.n_out = Cout; prm.y = (bf16*)y;
  prm.tiles = (int)((P + 127) / 128);
  // two CTAs per SM: enough row tiles per CTA that the whole grid is one wave
  const int n_tiles = Cout / GR_N;
  const int slots = sm_count() * 2 / n_tiles > 0 ? sm_count() * 2 / n_tiles : 1;
  prm.mt = (prm.tiles + slots - 1) / slots;
  if (prm.mt < 1) prm.mt = 1;
  dim3 grid((prm.tiles + prm.mt - 1) / prm.mt, n_tiles);
  constexpr int smem = GR_STAGES * GR_A + GR_B + 2 * GR_TILE + 16 * 2 * GR_N * 4 + (2 * GR_STAGES + 5) * 8 + 16 + GR_N * 4 + 1024;
  static unsigned long long configured = 0;
  if (first_on_device(&configured)) {
    DDPB_CUDA_CHECK(cudaFuncSetAttribute(gemm_rows_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    DDPB_CUDA_CHECK(cudaFuncSetAttribute(gemm_rows_tc_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                         cudaSharedmemCarveoutMaxShared));
  }
  prm.pdl = kernel_flags(stream, 1);
  DDPB_CUDA_CHECK(launch_ex(gemm_rows_tc_kernel, grid, dim3(192), smem, stream, dim3(1, 1, 1), 1, tmA, tmB, bias, stats, prm));
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}
