// Diagnostic: how many bytes per clock can ONE CTA pull from L2 into shared memory?
//
// The conv kernels' main loops run at ~27-32 B/clk per CTA (tools/kernel_bench.py, profiles/).  This
// probe runs the same producer / consumer mbarrier ring with different load flavours and consumers,
// so that the limiter (TMA box shape, stage hand-off, or the MMA side) can be told apart:
//   mode 0  conv A box (4-D NHWC, 64 ch x 128 pixels, shifted per tap) + weight box (2-D, 64 x 64)
//   mode 1  conv A box only
//   mode 2  2-D matrix box (64 columns x 128 rows of a [rows][cols] matrix over `flat`; cols sets the row stride)
//   mode 3  1-D bulk copy of 16 KB contiguous
//   mode 4  mode 0 with the real consumer: 4 x tcgen05.mma (128x64x16) + tcgen05.commit per stage
//   mode 5  mode 4 with TWO producer warps (even stages / odd stages)
//   mode 6  mode 4, the A box issued as two 64-row halves by two different warps
// No reference counterpart; used by tools/tma_probe.py only.
#include "../common.cuh"
#include "../../include/ddp_b200.h"
#include "../runtime/tensormap.h"

namespace ddpb {

constexpr int PB_A = 128 * 128;      // 16 KB
constexpr int PB_B = 64 * 128;       // 8 KB

template <int STAGES>
__global__ void __launch_bounds__(128)
tma_probe_kernel(const __grid_constant__ CUtensorMap tmA4, const __grid_constant__ CUtensorMap tmA2,
                 const __grid_constant__ CUtensorMap tmB, const uint8_t* __restrict__ flat, long long flat_bytes,
                 int mode, int iters, int k_chunks, int m_tiles, int m2_chunks, int m2_tiles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr int STAGE = PB_A + PB_B;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE);
  uint64_t* empty_bar = full_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(empty_bar + STAGES);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_tile = blockIdx.x % m_tiles, n_tile = (blockIdx.x / m_tiles) % 8;
  if (warp == 0 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    fence_mbar_init();
  }
  const bool mma = mode >= 4;
  if (warp == 1 && mma) tmem_alloc<64>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const bool with_b = mode == 0 || mma;
  const uint32_t bytes = PB_A + (with_b ? PB_B : 0);
  if (mode == 5 && (warp == 0 || warp == 2) && lane == 0) {
    // two independent producers: warp 0 fills the even stages, warp 2 the odd ones
    for (int it = warp >> 1; it < iters; it += 2) {
      const int stage = it % STAGES, use = it / STAGES;
      if (use > 0) mbar_wait(&empty_bar[stage], (use - 1) & 1);
      uint8_t* sa = smem + stage * STAGE;
      mbar_expect_tx(&full_bar[stage], bytes);
      const int kk = it % (9 * k_chunks);
      const int tap = kk / k_chunks, kc = kk % k_chunks;
      tma_load_4d(sa, &tmA4, &full_bar[stage], kc * 64, tap % 3 - 1, tap / 3 - 1, m_tile * 8);
      tma_load_2d(sa + PB_A, &tmB, &full_bar[stage], tap * k_chunks * 64 + kc * 64, n_tile * 64);
    }
  } else if (mode == 6 && (warp == 0 || warp == 2) && lane == 0) {
    // warp 0: first half of the A box + B; warp 2: second half of the A box (tmA2 map reused as a 64-row box is
    // not available, so both halves use the 4-D map with 4 of the 8 images: coordinates n and n + 4)
    for (int it = 0; it < iters; ++it) {
      const int stage = it % STAGES, use = it / STAGES;
      if (use > 0) mbar_wait(&empty_bar[stage], (use - 1) & 1);
      uint8_t* sa = smem + stage * STAGE;
      const int kk = it % (9 * k_chunks);
      const int tap = kk / k_chunks, kc = kk % k_chunks;
      if (warp == 0) {
        mbar_expect_tx(&full_bar[stage], bytes);
        tma_load_4d(sa, &tmA4, &full_bar[stage], kc * 64, tap % 3 - 1, tap / 3 - 1, m_tile * 8);
        tma_load_2d(sa + PB_A, &tmB, &full_bar[stage], tap * k_chunks * 64 + kc * 64, n_tile * 64);
      } else {
        tma_load_4d(sa + PB_A / 2, &tmA4, &full_bar[stage], kc * 64, tap % 3 - 1, tap / 3 - 1, m_tile * 8 + 4);
      }
    }
  } else if (mode < 5 && warp == 0 && lane == 0) {
    for (int it = 0; it < iters; ++it) {
      const int stage = it % STAGES, use = it / STAGES;
      if (use > 0) mbar_wait(&empty_bar[stage], (use - 1) & 1);
      uint8_t* sa = smem + stage * STAGE;
      mbar_expect_tx(&full_bar[stage], bytes);
      const int kk = it % (9 * k_chunks);
      const int tap = kk / k_chunks, kc = kk % k_chunks;
      if (mode == 0 || mode == 1 || mode == 4) {
        tma_load_4d(sa, &tmA4, &full_bar[stage], kc * 64, tap % 3 - 1, tap / 3 - 1, m_tile * 8);
      } else if (mode == 2) {
        tma_load_2d(sa, &tmA2, &full_bar[stage], (it % m2_chunks) * 64, ((blockIdx.x * 131 + it / m2_chunks) % m2_tiles) * 128);
      } else {
        const long long off = (((long long)blockIdx.x * 977 + it) * PB_A) % (flat_bytes - PB_A);
        const uint8_t* src = flat + (off & ~1023LL);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(sa)), "l"(src), "r"((uint32_t)PB_A), "r"(smem_u32(&full_bar[stage])) : "memory");
      }
      if (with_b) tma_load_2d(sa + PB_A, &tmB, &full_bar[stage], tap * k_chunks * 64 + kc * 64, n_tile * 64);
    }
  } else if (warp == 1 && lane == 0) {
    const uint32_t tmem_base = mma ? *tmem_slot : 0;
    constexpr uint32_t idesc = umma_idesc_bf16(128, 64, false, false);
    for (int it = 0; it < iters; ++it) {
      const int stage = it % STAGES, use = it / STAGES;
      mbar_wait(&full_bar[stage], use & 1);
      if (mma) {
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * STAGE), sb = sa + PB_A;
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4)
          umma_bf16(tmem_base, umma_desc(sa + k4 * 32, 16, 1024), umma_desc(sb + k4 * 32, 16, 1024), idesc, (it | k4) != 0);
        umma_commit(&empty_bar[stage]);
      } else {
        mbar_arrive(&empty_bar[stage]);
      }
    }
    if (mma) {
      // drain: the last commit must have arrived before TMEM is released
      const int last = (iters - 1) % STAGES;
      mbar_wait(&empty_bar[last], ((iters - 1) / STAGES) & 1);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1 && mma) tmem_dealloc<64>(*tmem_slot);
}

}  // namespace ddpb

using namespace ddpb;

// act: bf16 [256,4,4,512] (>= 4 MB), wgt: bf16 [512][4608], flat: any buffer of flat_bytes.
extern "C" int ddpb_probe_tma(int mode, int stages, int iters, int ctas, int smem_pad_kb, int cols, const void* act,
                              const void* wgt, const void* flat, long long flat_bytes, cudaStream_t stream) {
  DDPB_REQUIRE(mode >= 0 && mode <= 6 && (stages == 2 || stages == 4 || stages == 8) && iters > 0 && ctas > 0, "probe_tma: bad arguments");
  CUtensorMap tmA4, tmA2, tmB;
  // mode 6 loads the A tile as two boxes of 4 images each
  int rc = make_tmap_nhwc_bf16(&tmA4, act, 256, 4, 4, 512, 64, 4, 4, mode == 6 ? 4 : 8);
  if (rc) return rc;
  DDPB_REQUIRE(cols >= 64 && cols % 64 == 0, "probe_tma: cols must be a multiple of 64");
  // at most 16 MB of `flat` so that the matrix stays L2-resident
  long long m2_rows = (flat_bytes < (16LL << 20) ? flat_bytes : (16LL << 20)) / ((long long)cols * 2) / 128 * 128;
  DDPB_REQUIRE(m2_rows >= 128, "probe_tma: flat buffer too small");
  rc = make_tmap_2d_bf16(&tmA2, flat, m2_rows, cols, 64, 128);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tmB, wgt, 512, 4608, 64, 64);
  if (rc) return rc;
  // smem_pad_kb > 0 inflates the allocation so that only one CTA fits per SM
  const int smem = stages * (PB_A + PB_B) + 2 * stages * 8 + 16 + 1024 + smem_pad_kb * 1024;
  DDPB_REQUIRE(smem <= 227 * 1024, "probe_tma: too much shared memory");
  if (stages == 2) {
    DDPB_CUDA_CHECK(cudaFuncSetAttribute(tma_probe_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    tma_probe_kernel<2><<<ctas, 128, smem, stream>>>(tmA4, tmA2, tmB, (const uint8_t*)flat, flat_bytes, mode, iters, 8, 32, cols / 64, (int)(m2_rows / 128));
  } else if (stages == 4) {
    DDPB_CUDA_CHECK(cudaFuncSetAttribute(tma_probe_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    tma_probe_kernel<4><<<ctas, 128, smem, stream>>>(tmA4, tmA2, tmB, (const uint8_t*)flat, flat_bytes, mode, iters, 8, 32, cols / 64, (int)(m2_rows / 128));
  } else {
    DDPB_CUDA_CHECK(cudaFuncSetAttribute(tma_probe_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    tma_probe_kernel<8><<<ctas, 128, smem, stream>>>(tmA4, tmA2, tmB, (const uint8_t*)flat, flat_bytes, mode, iters, 8, 32, cols / 64, (int)(m2_rows / 128));
  }
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}
