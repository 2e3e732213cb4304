// fp32 convolutions on the tensor cores: tcgen05.mma kind::tf32 (fp32 in shared memory, read with a 10-bit
// mantissa, fp32 accumulation in TMEM).  The fp32 training path of the engine - the reference is fp32
// throughout (ref Part 1/main.py:10; Part 1/model.py:18-23) - used to run on SIMT kernels, two orders of
// magnitude slower than the bf16 path; TF32 is also what stock PyTorch uses for fp32 convolutions on a GPU.
//
// Same implicit GEMM as conv_tc.cu with 4-byte elements: a 128-byte shared-memory row now holds 32 channels,
// so one pipeline stage is 128 pixels x 32 channels of activations and BN x 32 of weights, K per MMA is 8
// (still 32 bytes along a K-major row), and the MN-major operands (dgrad weights, both wgrad operands) come
// in 32-element-wide blocks.  Kept deliberately simple - one CTA per 128-pixel x BN-channel tile, no split-K,
// no fused BatchNorm epilogues: this path exists for fp32 parity, the bf16 kernels are the fast path.
//   fprop:  Y[p, co]  = sum_{tap,ci} X[p + shift(tap), ci]  * W[co, tap, ci]
//   dgrad:  dX[p, ci] = sum_{tap,co} dY[p - shift(tap), co] * W[co, tap, ci]
//   wgrad:  dW[co, tap, ci] = sum_p dY[p, co] * X[p + shift(tap), ci]
#include "../common.cuh"
#include "../../include/ddp_b200.h"
#include "../runtime/tensormap.h"

namespace ddpb {

void pixel_box(int H, int W, int pixels, int* bw, int* bh, int* bn);   // conv_tc.cu

constexpr int TF_M = 128;            // pixels per CTA
constexpr int TF_K = 32;             // fp32 elements per 128-byte row
constexpr int TF_A = TF_M * 128;     // 16 KB

struct ConvTfParams {
  int tiles_per_img, imgs_per_tile, rows_per_tile;
  long long total_pixels;
  int n_out, k_chunks, k_total_cols, taps;
  float* out;
  int flags;
};

template <int BN, int STAGES, bool DGRAD>
__global__ void __launch_bounds__(192)
conv3x3_tf32_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const float* __restrict__ bias, float* __restrict__ stats, const ConvTfParams prm) {
  constexpr int B_BYTES = BN * 128;
  constexpr int STAGE = TF_A + B_BYTES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* accum_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);
  float* bias_s = reinterpret_cast<float*>(tmem_slot + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_tile = blockIdx.x, n_tile = blockIdx.y;
  int n0, h0;
  if (prm.tiles_per_img > 0) { n0 = m_tile / prm.tiles_per_img; h0 = (m_tile % prm.tiles_per_img) * prm.rows_per_tile; }
  else { n0 = m_tile * prm.imgs_per_tile; h0 = 0; }
  const int k_iters = prm.taps * prm.k_chunks;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    mbar_init(accum_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<BN>(tmem_slot);
  if (prm.flags & 1) pdl_launch_dependents();
  pdl_wait();
  ktrace_begin(prm.flags);
  if (!DGRAD) for (int i = threadIdx.x; i < BN; i += blockDim.x) bias_s[i] = bias ? ld_coh(&bias[n_tile * BN + i]) : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      for (int it = 0; it < k_iters; ++it) {
        const int stage = it % STAGES, use = it / STAGES;
        if (use > 0) mbar_wait(&empty_bar[stage], (use - 1) & 1);
        const int tap = it / prm.k_chunks, kc = it % prm.k_chunks;
        const int r = prm.taps == 1 ? 1 : tap / 3, s = prm.taps == 1 ? 1 : tap % 3;
        uint8_t* sa = smem + stage * STAGE;
        uint8_t* sb = sa + TF_A;
        mbar_expect_tx(&full_bar[stage], STAGE);
        const int dh = DGRAD ? 1 - r : r - 1, dw = DGRAD ? 1 - s : s - 1;
        tma_load_4d(sa, &tmA, &full_bar[stage], kc * TF_K, dw, h0 + dh, n0);
        if (!DGRAD) {
          // weights as a [Cout][9*Cin] fp32 matrix: BN rows (co), 32 contiguous K columns
          tma_load_2d(sb, &tmB, &full_bar[stage], tap * prm.k_total_cols + kc * TF_K, n_tile * BN);
        } else {
          // the same matrix read sideways: 32 rows of K (= co), BN / 32 boxes of 32 contiguous ci
#pragma unroll
          for (int j = 0; j < BN / 32; ++j)
            tma_load_2d(sb + j * (TF_K * 128), &tmB, &full_bar[stage], tap * prm.k_total_cols + n_tile * BN + j * 32,
                        kc * TF_K);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_tf32(TF_M, BN, false, DGRAD);
      for (int it = 0; it < k_iters; ++it) {
        const int stage = it % STAGES, use = it / STAGES;
        mbar_wait(&full_bar[stage], use & 1);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * STAGE);
        const uint32_t sb = sa + TF_A;
#pragma unroll
        for (int kk = 0; kk < TF_K / 8; ++kk) {
          // K-major: 8 K-elements = 32 bytes along the row; 8-row groups 1024 B apart
          const uint64_t da = umma_desc(sa + kk * 32, 16, 1024);
          // dgrad B is MN-major (32-byte-atom swizzle): 8 K-rows = 1024 bytes = two 4-row atoms 512 B apart (SBO);
          // 32-wide N blocks are TF_K * 128 bytes apart (LBO)
          const uint64_t db = DGRAD ? umma_desc_base32(sb + kk * 1024, TF_K * 128, 512) : umma_desc(sb + kk * 32, 16, 1024);
          umma_tf32(tmem_base, da, db, idesc, (it | kk) != 0);
        }
        umma_commit(&empty_bar[stage]);
      }
      umma_commit(accum_bar);
    }
  } else {
    mbar_wait(accum_bar, 0);
    tc_fence_after();
  }
  // Epilogue: TMEM -> (+bias) -> fp32 tile in shared memory (16-byte chunks XOR-swizzled by the row) ->
  // coalesced 16-byte stores, with the per-channel sum / sum of squares for BatchNorm taken from the stored values
  constexpr int CH = BN / 4;                  // 16-byte chunks per tile row
  constexpr int PARTS = 128 / CH;             // row groups of the store pass
  static_assert(CH >= 8 && CH <= 128 && 128 % CH == 0, "tile width");
  float* tile = reinterpret_cast<float*>(smem);                               // [128 rows][BN] fp32
  float* red = tile + TF_M * BN;                                              // [PARTS][2 * BN]
  if (warp >= 2) {
    const int q = warp & 3, row = q * 32 + lane, t = threadIdx.x - 64;
    const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
    for (int c = 0; c < BN; c += 16) {
      float v[16];
      tmem_ld_32x32_x16(lane_base + c, v);
      tmem_ld_wait();
      if (!DGRAD) {
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] += bias_s[c + j];
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int chunk = (c >> 2) + j;
        *reinterpret_cast<float4*>(tile + row * BN + ((chunk ^ (row & 7)) << 2)) =
            make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      }
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    const int ch = t % CH, part = t / CH;
    const int col0 = n_tile * BN + ch * 4;
    const long long row0 = (long long)m_tile * TF_M;
    const bool want_stats = !DGRAD && stats != nullptr;
    float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
    for (int k = 0; k < CH; ++k) {
      const int r = part + k * PARTS;
      const long long grow = row0 + r;
      if (grow >= prm.total_pixels) break;
      const float4 u = *reinterpret_cast<const float4*>(tile + r * BN + ((ch ^ (r & 7)) << 2));
      *reinterpret_cast<float4*>(prm.out + grow * prm.n_out + col0) = u;
      if (want_stats) {
        s1[0] += u.x; s1[1] += u.y; s1[2] += u.z; s1[3] += u.w;
        s2[0] = fmaf(u.x, u.x, s2[0]); s2[1] = fmaf(u.y, u.y, s2[1]); s2[2] = fmaf(u.z, u.z, s2[2]); s2[3] = fmaf(u.w, u.w, s2[3]);
      }
    }
    if (want_stats) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        red[part * 2 * BN + ch * 4 + j] = s1[j];
        red[part * 2 * BN + BN + ch * 4 + j] = s2[j];
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      for (int i = t; i < 2 * BN; i += 128) {
        float acc = 0.f;
        for (int pz = 0; pz < PARTS; ++pz) acc += red[pz * 2 * BN + i];
        const int cc = i < BN ? i : i - BN;
        atomicAdd(&stats[(i < BN ? 0 : prm.n_out) + n_tile * BN + cc], acc);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<BN>(tmem_base);
  ktrace_end(prm.flags);
}

// ---- weight gradient ------------------------------------------------------------------------------------------
// One CTA: 128 output channels x 64 input channels x the three taps of one filter row, over a slice of the pixels
// (split-K); both operands MN-major in 32-wide blocks; partial results added to the fp32 gradient with TMA
// reduce-add, exactly as in wgrad_tc.cu.
constexpr int TW_K = 32;             // pixels per pipeline stage
constexpr int TW_BOX = TW_K * 128;   // one 32-pixel x 32-channel fp32 box: 4 KB

struct WgradTfParams {
  int chunks_per_img, imgs_per_chunk, rows_per_chunk, total_chunks, Cin;
  int flags;
};

template <int STAGES>
__global__ void __launch_bounds__(192)
wgrad_tf32_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmX,
                  const __grid_constant__ CUtensorMap tmW, const WgradTfParams prm) {
  constexpr int STAGE = (4 + 3 * 2) * TW_BOX;                 // 4 dY boxes (128 co) + 3 taps x 2 X boxes (64 ci) = 40 KB
  constexpr int EPI = 3 * 2 * 128 * 128;                      // fp32 staging of the epilogue: 96 KB
  constexpr int DATA = STAGES * STAGE > EPI ? STAGES * STAGE : EPI;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + DATA);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* accum_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int co_blk = blockIdx.x / (prm.Cin / 64), ci_blk = blockIdx.x % (prm.Cin / 64);
  const int r = blockIdx.y;
  const int split = blockIdx.z, nsplit = gridDim.z;
  const int chunk_lo = (int)((long long)prm.total_chunks * split / nsplit);
  const int chunk_hi = (int)((long long)prm.total_chunks * (split + 1) / nsplit);
  const int k_iters = chunk_hi - chunk_lo;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmDY); tma_prefetch_desc(&tmX); tma_prefetch_desc(&tmW);
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    mbar_init(accum_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<256>(tmem_slot);
  if (prm.flags & 1) pdl_launch_dependents();
  pdl_wait();
  ktrace_begin(prm.flags);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (k_iters > 0) {
    if (warp == 0) {
      if (lane == 0) {
        for (int it = 0; it < k_iters; ++it) {
          const int stage = it % STAGES, use = it / STAGES;
          if (use > 0) mbar_wait(&empty_bar[stage], (use - 1) & 1);
          const int chunk = chunk_lo + it;
          int n0, h0;
          if (prm.chunks_per_img > 0) { n0 = chunk / prm.chunks_per_img; h0 = (chunk % prm.chunks_per_img) * prm.rows_per_chunk; }
          else { n0 = chunk * prm.imgs_per_chunk; h0 = 0; }
          uint8_t* s = smem + stage * STAGE;
          mbar_expect_tx(&full_bar[stage], STAGE);
#pragma unroll
          for (int a = 0; a < 4; ++a) tma_load_4d(s + a * TW_BOX, &tmDY, &full_bar[stage], co_blk * 128 + a * 32, 0, h0, n0);
#pragma unroll
          for (int t = 0; t < 3; ++t)
#pragma unroll
            for (int b = 0; b < 2; ++b)
              tma_load_4d(s + (4 + t * 2 + b) * TW_BOX, &tmX, &full_bar[stage], ci_blk * 64 + b * 32, t - 1, h0 + r - 1, n0);
        }
      }
    } else if (warp == 1) {
      if (lane == 0) {
        constexpr uint32_t idesc = umma_idesc_tf32(128, 64, true, true);
        for (int it = 0; it < k_iters; ++it) {
          const int stage = it % STAGES, use = it / STAGES;
          mbar_wait(&full_bar[stage], use & 1);
          tc_fence_after();
          const uint32_t s = smem_u32(smem + stage * STAGE);
#pragma unroll
          for (int kk = 0; kk < TW_K / 8; ++kk) {
            // MN-major (32-byte-atom swizzle): 8 pixels (K) = 1024 bytes = two 4-row atoms 512 B apart (SBO);
            // 32-channel blocks one box (4 KB) apart (LBO)
            const uint64_t da = umma_desc_base32(s + kk * 1024, TW_BOX, 512);
#pragma unroll
            for (int t = 0; t < 3; ++t)
              umma_tf32(tmem_base + t * 64, da, umma_desc_base32(s + (4 + t * 2) * TW_BOX + kk * 1024, TW_BOX, 512), idesc,
                        (it | kk) != 0);
          }
          umma_commit(&empty_bar[stage]);
        }
        umma_commit(accum_bar);
      }
    } else {
      const int q = warp & 3, row = q * 32 + lane;
      mbar_wait(accum_bar, 0);
      tc_fence_after();
#pragma unroll 1
      for (int t = 0; t < 3; ++t)
#pragma unroll 1
        for (int c = 0; c < 64; c += 32) {
          float v[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + t * 64 + c, v);
          tmem_ld_wait();
          uint8_t* rbase = smem + (t * 2 + (c >> 5)) * (128 * 128) + row * 128;
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(rbase + ((j ^ (row & 7)) << 4)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        }
      fence_async_proxy();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (threadIdx.x == 64) {
#pragma unroll
        for (int t = 0; t < 3; ++t)
#pragma unroll
          for (int hb = 0; hb < 2; ++hb)
            tma_reduce_add_2d(&tmW, smem + (t * 2 + hb) * (128 * 128), (r * 3 + t) * prm.Cin + ci_blk * 64 + hb * 32, co_blk * 128);
        tma_store_commit();
        tma_store_wait_read();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<256>(tmem_base);
  ktrace_end(prm.flags);
}

}  // namespace ddpb
DDPB_KTRACE_SETTER(conv_tf32)

using namespace ddpb;

namespace {

template <int BN, int STAGES, bool DGRAD>
int launch_tf32(const float* in, const float* w, float* out, const float* bias, float* stats, int B, int H, int W, int Cin,
                int Cout, cudaStream_t stream) {
  const int Kc = DGRAD ? Cout : Cin, Nc = DGRAD ? Cin : Cout;
  int bw, bh, bn;
  pixel_box(H, W, TF_M, &bw, &bh, &bn);
  CUtensorMap tmA, tmB;
  int rc = make_tmap_nhwc_f32(&tmA, in, B, H, W, Kc, TF_K, bw, bh, bn);
  if (rc) return rc;
  rc = make_tmap_2d_f32(&tmB, w, Cout, 9LL * Cin, TF_K, DGRAD ? TF_K : BN, DGRAD);     // dgrad reads the weights MN-major
  if (rc) return rc;
  ConvTfParams prm;
  prm.tiles_per_img = H * W >= TF_M ? H * W / TF_M : 0;
  prm.imgs_per_tile = H * W >= TF_M ? 0 : TF_M / (H * W);
  prm.rows_per_tile = bh;
  prm.total_pixels = (long long)B * H * W;
  prm.n_out = Nc; prm.k_chunks = Kc / TF_K; prm.k_total_cols = Cin; prm.taps = 9; prm.out = out;
  static_assert(TF_M * BN * 4 + (128 / (BN / 4)) * 2 * BN * 4 <= STAGES * (TF_A + BN * 128), "epilogue tile must fit in the stages");
  const int m_tiles = prm.tiles_per_img > 0 ? B * prm.tiles_per_img : ceil_div(B, prm.imgs_per_tile);
  constexpr int smem = STAGES * (TF_A + BN * 128) + (2 * STAGES + 1) * 8 + 16 + BN * 4 + 1024;
  auto kern = conv3x3_tf32_kernel<BN, STAGES, DGRAD>;
  static unsigned long long configured = 0;
  if (first_on_device(&configured)) {
    DDPB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    DDPB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  }
  prm.flags = kernel_flags(stream, 1);
  DDPB_CUDA_CHECK(launch_ex(kern, dim3(m_tiles, Nc / BN), dim3(192), smem, stream, dim3(1, 1, 1), 1, tmA, tmB, bias, stats, prm));
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}

int check_tf32(const char* what, int B, int H, int W, int Cin, int Cout) {
  DDPB_REQUIRE(B > 0 && H == W && (H == 32 || H == 16 || H == 8 || H == 4 || H == 2), "%s: unsupported spatial size %dx%d", what, H, W);
  DDPB_REQUIRE(Cin % 64 == 0 && Cout % 64 == 0, "%s: channels must be multiples of 64 (got %d, %d)", what, Cin, Cout);
  return DDPB_OK;
}

}  // namespace

// y[B,H,W,Cout] = conv3x3(x[B,H,W,Cin], w[Cout,3,3,Cin]) + bias, all fp32, TF32 tensor-core math.
// `stats` (fp32 [2*Cout], may be null) ACCUMULATES per-channel sum and sum of squares of y.
extern "C" int ddpb_conv3x3_fwd_tf32(const float* x, const float* w, const float* bias, float* y, float* stats, int B,
                                     int H, int W, int Cin, int Cout, cudaStream_t stream) {
  int rc = check_tf32("conv_tf32 fwd", B, H, W, Cin, Cout);
  if (rc) return rc;
  const long long tiles128 = Cout % 128 == 0 ? (long long)ceil_div((long long)B * H * W, TF_M) * (Cout / 128) : 0;
  if (tiles128 >= sm_count()) return launch_tf32<128, 3, false>(x, w, y, bias, stats, B, H, W, Cin, Cout, stream);
  return launch_tf32<64, 4, false>(x, w, y, bias, stats, B, H, W, Cin, Cout, stream);
}

// dx[B,H,W,Cin] = conv3x3_transpose(dy[B,H,W,Cout], w[Cout,3,3,Cin]); fp32, TF32 math.
extern "C" int ddpb_conv3x3_dgrad_tf32(const float* dy, const float* w, float* dx, int B, int H, int W, int Cin, int Cout,
                                       cudaStream_t stream) {
  int rc = check_tf32("conv_tf32 dgrad", B, H, W, Cin, Cout);
  if (rc) return rc;
  const long long tiles128 = Cin % 128 == 0 ? (long long)ceil_div((long long)B * H * W, TF_M) * (Cin / 128) : 0;
  if (tiles128 >= sm_count()) return launch_tf32<128, 3, true>(dy, w, dx, nullptr, nullptr, B, H, W, Cin, Cout, stream);
  return launch_tf32<64, 4, true>(dy, w, dx, nullptr, nullptr, B, H, W, Cin, Cout, stream);
}

// dw (fp32 [Cout,3,3,Cin]) += dy^T * shifted(x); fp32 operands, TF32 math; split_k = 0 picks.
extern "C" int ddpb_conv3x3_wgrad_tf32(const float* dy, const float* x, float* dw, int B, int H, int W, int Cin, int Cout,
                                       int split_k, cudaStream_t stream) {
  DDPB_REQUIRE(B > 0 && H == W && (H == 32 || H == 16 || H == 8 || H == 4 || H == 2), "wgrad_tf32: unsupported spatial size %dx%d", H, W);
  DDPB_REQUIRE(Cin % 64 == 0 && Cout % 128 == 0, "wgrad_tf32: need Cin %% 64 == 0 and Cout %% 128 == 0 (got %d, %d)", Cin, Cout);
  int bw, bh, bn;
  pixel_box(H, W, TW_K, &bw, &bh, &bn);
  CUtensorMap tmDY, tmX, tmW;
  int rc = make_tmap_2d_f32(&tmW, dw, Cout, 9LL * Cin, 32, 128);
  if (rc) return rc;
  rc = make_tmap_nhwc_f32(&tmDY, dy, B, H, W, Cout, 32, bw, bh, bn, true);      // both operands are MN-major
  if (rc) return rc;
  rc = make_tmap_nhwc_f32(&tmX, x, B, H, W, Cin, 32, bw, bh, bn, true);
  if (rc) return rc;
  WgradTfParams prm;
  prm.chunks_per_img = H * W >= TW_K ? H * W / TW_K : 0;
  prm.imgs_per_chunk = H * W >= TW_K ? 0 : TW_K / (H * W);
  prm.rows_per_chunk = bh;
  prm.total_chunks = prm.chunks_per_img > 0 ? B * prm.chunks_per_img : ceil_div(B, prm.imgs_per_chunk);
  prm.Cin = Cin;
  const int base_ctas = (Cout / 128) * (Cin / 64) * 3;
  if (split_k <= 0) {
    long long best = -1;
    const int max_split = prm.total_chunks < 128 ? prm.total_chunks : 128;
    for (int sp = 1; sp <= max_split; ++sp) {
      const long long waves = ((long long)base_ctas * sp + 2 * sm_count() - 1) / (2 * sm_count());
      const long long iters = (prm.total_chunks + sp - 1) / sp;
      const long long cost = waves * (iters * 2 + 8);
      if (best < 0 || cost < best) { best = cost; split_k = sp; }
    }
  }
  if (split_k > prm.total_chunks) split_k = prm.total_chunks;
  constexpr int ST = 2;
  constexpr int pipe = ST * 10 * TW_BOX, epi = 3 * 2 * 128 * 128;
  constexpr int smem = (pipe > epi ? pipe : epi) + (2 * ST + 1) * 8 + 16 + 1024;
  auto kern = wgrad_tf32_kernel<ST>;
  static unsigned long long configured = 0;
  if (first_on_device(&configured)) {
    DDPB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    DDPB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  }
  prm.flags = kernel_flags(stream, 2);
  DDPB_CUDA_CHECK(launch_ex(kern, dim3(base_ctas / 3, 3, split_k), dim3(192), smem, stream, dim3(1, 1, 1), 2, tmDY, tmX, tmW, prm));
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}
