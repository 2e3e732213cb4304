// conv3x3 forward / dgrad on CTA PAIRS: tcgen05.mma.cta_group::2 (UMMA M = 256).
//
// Same implicit GEMM as conv_tc.cu (TMA box loads whose out-of-image zero fill is the padding,
// K-major A, K-major or MN-major B, TMEM accumulators, TMA-store epilogue with the BatchNorm
// statistics), but two CTAs on the two SMs of a TPC form one MMA:
//   * each CTA loads ITS OWN 128-pixel A tile and only HALF of the weight tile (N/2 channels);
//   * the leader CTA issues tcgen05.mma.cta_group::2, which reads A and B from both CTAs' shared
//     memory and writes each CTA's 128 x N accumulator into that CTA's TMEM;
//   * tcgen05.commit ... multicast releases the pipeline stage in both CTAs.
// The single-CTA kernel needs 128 B of operands per tensor-core cycle for a 128x128 tile and is
// bound by shared-memory fill (~42 B/clk/SM measured).  A pair with N = 256 needs 62 B/clk per SM:
// the weight traffic per SM is halved and the A traffic is amortised over twice as many columns.
#include "../common.cuh"
#include "../../include/ddp_b200.h"
#include "../runtime/tensormap.h"

namespace ddpb {

void pixel_box(int H, int W, int pixels, int* bw, int* bh, int* bn);   // conv_tc.cu

namespace {

constexpr int T2_M = 128;            // pixels per CTA (256 per pair)
constexpr int T2_K = 64;             // bf16 elements per 128-byte row
constexpr int T2_A_BYTES = T2_M * T2_K * 2;

struct Conv2Params {
  int tiles_per_img, imgs_per_tile, rows_per_tile;
  long long total_pixels;
  int n_out, k_chunks, k_total_cols;
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of the LEADER CTA's copy of a barrier (bit 24 selects the CTA of the pair)
__device__ __forceinline__ uint32_t leader_bar(const uint64_t* bar) { return smem_u32(bar) & 0xFEFFFFFFu; }

// arrive on the barrier at this shared-memory offset in BOTH CTAs once the pair's MMAs have retired
__device__ __forceinline__ void umma2_commit_both(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(smem_u32(bar)), "h"((uint16_t)3)
      : "memory");
}

template <int BN, int STAGES, bool DGRAD>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(192)
conv3x3_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                   const __grid_constant__ CUtensorMap tmC, const float* __restrict__ bias,
                   float* __restrict__ stats, const Conv2Params prm) {
  constexpr int BH = BN / 2;                     // weight-tile channels held by each CTA
  constexpr int B_BYTES = BH * T2_K * 2;
  constexpr int STAGE_BYTES = T2_A_BYTES + B_BYTES;
  constexpr int TILE_BYTES = T2_M * BN * 2;      // epilogue staging tile
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* accum_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);
  float* bias_s = reinterpret_cast<float*>(tmem_slot + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();
  const bool leader = cta == 0;
  const int m_tile = blockIdx.x, n_tile = blockIdx.y;
  int n0, h0;
  if (prm.tiles_per_img > 0) { n0 = m_tile / prm.tiles_per_img; h0 = (m_tile % prm.tiles_per_img) * prm.rows_per_tile; }
  else { n0 = m_tile * prm.imgs_per_tile; h0 = 0; }
  const int k_iters = 9 * prm.k_chunks;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    mbar_init(accum_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem2_alloc<BN>(tmem_slot);
  if (!DGRAD) for (int i = threadIdx.x; i < BN; i += blockDim.x) bias_s[i] = bias ? bias[n_tile * BN + i] : 0.f;
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                 // the peer's barriers exist before anything arrives on them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs; bytes are accounted on the leader's barrier) =====
    if (lane == 0) {
      for (int it = 0; it < k_iters; ++it) {
        const int stage = it % STAGES, use = it / STAGES;
        if (use > 0) mbar_wait(&empty_bar[stage], (use - 1) & 1);
        const int tap = it / prm.k_chunks, kc = it % prm.k_chunks;
        const int r = tap / 3, s = tap % 3;
        uint8_t* sa = smem + stage * STAGE_BYTES;
        uint8_t* sb = sa + T2_A_BYTES;
        if (leader) mbar_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);
        const uint32_t bar = leader_bar(&full_bar[stage]);
        const int dh = DGRAD ? 1 - r : r - 1, dw = DGRAD ? 1 - s : s - 1;
        tma2_load_4d(sa, &tmA, bar, kc * T2_K, dw, h0 + dh, n0);
        if (!DGRAD) {
          tma2_load_2d(sb, &tmB, bar, tap * prm.k_total_cols + kc * T2_K, n_tile * BN + (int)cta * BH);
        } else {
#pragma unroll
          for (int j = 0; j < BH / 64; ++j)
            tma2_load_2d(sb + j * (64 * 128), &tmB, bar,
                         tap * prm.k_total_cols + n_tile * BN + (int)cta * BH + j * 64, kc * T2_K);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer: one thread of the LEADER CTA drives both SMs =====================
    if (leader && lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(2 * T2_M, BN, false, DGRAD);
      for (int it = 0; it < k_iters; ++it) {
        const int stage = it % STAGES, use = it / STAGES;
        mbar_wait(&full_bar[stage], use & 1);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
        const uint32_t sb = sa + T2_A_BYTES;
#pragma unroll
        for (int kk = 0; kk < T2_K / 16; ++kk) {
          const uint64_t da = umma_desc(sa + kk * 32, 16, 1024);
          const uint64_t db = DGRAD ? umma_desc(sb + kk * 2048, 64 * 128, 1024) : umma_desc(sb + kk * 32, 16, 1024);
          umma2_bf16(tmem_base, da, db, idesc, (it | kk) != 0);
        }
        umma2_commit_both(&empty_bar[stage]);
      }
      umma2_commit_both(accum_bar);
    }
  } else {
    // ===================== epilogue (each CTA stores its own 128 x BN tile) =====================
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int t = threadIdx.x - 64;
    uint8_t* tile = smem;                                             // [BN/64][128 rows][128 B]
    float* red = reinterpret_cast<float*>(smem + TILE_BYTES);         // [parts][2*BN]
    mbar_wait(accum_bar, 0);
    tc_fence_after();
#pragma unroll 1
    for (int c = 0; c < BN; c += 32) {
      float v[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + c, v);
      tmem_ld_wait();
      if (!DGRAD) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] += bias_s[c + j];
      }
      uint8_t* rbase = tile + (c >> 6) * (T2_M * 128) + row * 128;
      const int j0 = (c & 63) >> 3;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint4 pk = make_uint4(pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                              pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
        *reinterpret_cast<uint4*>(rbase + (((j0 + j) ^ (row & 7)) << 4)) = pk;
      }
    }
    fence_async_proxy();
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (t == 0) {
#pragma unroll
      for (int hblk = 0; hblk < BN / 64; ++hblk)
        tma_store_2d(&tmC, tile + hblk * (T2_M * 128), n_tile * BN + hblk * 64, m_tile * T2_M);
      tma_store_commit();
    }
    if (!DGRAD && stats) {
      constexpr int GROUPS = BN / 4, PARTS = 128 / GROUPS, ROWS = T2_M / PARTS;
      const int g = t % GROUPS, part = t / GROUPS;
      const long long left = prm.total_pixels - (long long)m_tile * T2_M;
      const int valid_rows = left < T2_M ? (left < 0 ? 0 : (int)left) : T2_M;
      const uint8_t* cbase = tile + ((4 * g) >> 6) * (T2_M * 128);
      const int jc = ((4 * g) & 63) >> 3, inner = ((4 * g) & 7) * 2;
      float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
      const int r_end = min(valid_rows, (part + 1) * ROWS);
#pragma unroll 4
      for (int r = part * ROWS; r < r_end; ++r) {
        const uint2 u = *reinterpret_cast<const uint2*>(cbase + r * 128 + ((jc ^ (r & 7)) << 4) + inner);
        const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y);
        s1[0] += a.x; s1[1] += a.y; s1[2] += b.x; s1[3] += b.y;
        s2[0] = fmaf(a.x, a.x, s2[0]); s2[1] = fmaf(a.y, a.y, s2[1]);
        s2[2] = fmaf(b.x, b.x, s2[2]); s2[3] = fmaf(b.y, b.y, s2[3]);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        red[part * 2 * BN + 4 * g + j] = s1[j];
        red[part * 2 * BN + BN + 4 * g + j] = s2[j];
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      for (int i = t; i < 2 * BN; i += 128) {
        float acc = 0.f;
#pragma unroll
        for (int pz = 0; pz < PARTS; ++pz) acc += red[pz * 2 * BN + i];
        const int col = i < BN ? i : i - BN;
        atomicAdd(&stats[(i < BN ? 0 : prm.n_out) + n_tile * BN + col], acc);
      }
    }
    if (t == 0) tma_store_wait_read();
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                 // the peer may still be reading our shared memory / TMEM through the pair MMA
  if (warp == 1) tmem2_dealloc<BN>(tmem_base);
}

template <int BN, int STAGES>
constexpr int conv2_smem() {
  return STAGES * (T2_A_BYTES + (BN / 2) * T2_K * 2) + (2 * STAGES + 1) * 8 + 16 + BN * 4 + 1024;
}

template <int BN, int STAGES, bool DGRAD>
int launch2(const bf16* in, const bf16* w, bf16* out, const float* bias, float* stats, int B, int H, int W, int Cin,
            int Cout, cudaStream_t stream) {
  static_assert(STAGES * (T2_A_BYTES + (BN / 2) * T2_K * 2) >= T2_M * BN * 2 + 2 * 2 * BN * 4 * (128 / (BN / 4)),
                "pipeline smem must hold the epilogue staging tile and the stats partials");
  const int Kc = DGRAD ? Cout : Cin, Nc = DGRAD ? Cin : Cout;
  int bw, bh, bn;
  pixel_box(H, W, T2_M, &bw, &bh, &bn);
  CUtensorMap tmA, tmB, tmC;
  int rc = make_tmap_nhwc_bf16(&tmA, in, B, H, W, Kc, T2_K, bw, bh, bn);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tmC, out, (long long)B * H * W, Nc, 64, T2_M);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tmB, w, Cout, 9LL * Cin, T2_K, DGRAD ? 64 : BN / 2);
  if (rc) return rc;
  Conv2Params prm;
  prm.tiles_per_img = H * W >= T2_M ? H * W / T2_M : 0;
  prm.imgs_per_tile = H * W >= T2_M ? 0 : T2_M / (H * W);
  prm.rows_per_tile = bh;
  prm.total_pixels = (long long)B * H * W;
  prm.n_out = Nc;
  prm.k_chunks = Kc / T2_K;
  prm.k_total_cols = Cin;
  int m_tiles = prm.tiles_per_img > 0 ? B * prm.tiles_per_img : ceil_div(B, prm.imgs_per_tile);
  m_tiles = (m_tiles + 1) & ~1;                  // pairs; a surplus tile is fully out of range (clipped stores)
  dim3 grid(m_tiles, Nc / BN);
  constexpr int smem = conv2_smem<BN, STAGES>();
  auto kern = conv3x3_tc2_kernel<BN, STAGES, DGRAD>;
  static unsigned long long configured = 0;
  if (first_on_device(&configured)) {
    DDPB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  }
  kern<<<grid, 192, smem, stream>>>(tmA, tmB, tmC, bias, stats, prm);
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}

}  // namespace

// bn = 128 or 256 output channels per CTA pair (each CTA computes 128 pixels x bn channels)
int launch_conv_tc2(bool dgrad, int bn, const void* in, const void* w, void* out, const float* bias, float* stats,
                    int B, int H, int W, int Cin, int Cout, cudaStream_t stream) {
  const bf16* i = (const bf16*)in;
  const bf16* ww = (const bf16*)w;
  bf16* o = (bf16*)out;
  if (bn == 256) {
    return dgrad ? launch2<256, 6, true>(i, ww, o, nullptr, nullptr, B, H, W, Cin, Cout, stream)
                 : launch2<256, 6, false>(i, ww, o, bias, stats, B, H, W, Cin, Cout, stream);
  }
  if (bn == 128) {
    return dgrad ? launch2<128, 4, true>(i, ww, o, nullptr, nullptr, B, H, W, Cin, Cout, stream)
                 : launch2<128, 4, false>(i, ww, o, bias, stats, B, H, W, Cin, Cout, stream);
  }
  set_last_error("conv_tc2: tile must be 128 or 256 channels per CTA pair (got %d)", bn);
  return DDPB_ERR_ARG;
}

}  // namespace ddpb
