// conv3x3 (stride 1, pad 1) forward and data-gradient as implicit GEMM on tcgen05 tensor cores.
//
// Replaces nn.Conv2d forward and its dgrad (ref Part 1/model.py:18-23; SURVEY K-C1/K-C2).
//
// GEMM view (NHWC bf16 activations, KRSC bf16 weights, fp32 accumulation in TMEM):
//   fprop:  Y[p, co]  = sum_{tap,ci} X[p + shift(tap), ci]  * W[co, tap, ci]      M=pixels N=Cout K=9*Cin
//   dgrad:  dX[p, ci] = sum_{tap,co} dY[p - shift(tap), co] * W[co, tap, ci]      M=pixels N=Cin  K=9*Cout
//
// * One CTA = one 128-pixel x BN-channel output tile.  128 pixels are a TMA *box* of the 4-D
//   activation tensor (C, W, H, N); the filter tap is applied by shifting the box start by
//   (s-1, r-1).  Coordinates that fall outside the image (including negative ones) are
//   zero-filled by TMA, which IS the convolution's zero padding: no im2col buffer, no halo code.
// * A tiles land in shared memory as 128 rows x 128 B with the 128-byte swizzle = the canonical
//   K-major UMMA layout.  fprop B tiles (weights, K contiguous) are K-major too; dgrad reads the
//   SAME weight tensor as an MN-major B operand (rows = K = Cout, 128 B of Cin contiguous), so no
//   transposed weight copy exists anywhere.
// * Warp roles: warp 0 = TMA producer, warp 1 = TMEM allocator + single-thread MMA issuer,
//   warps 2..5 = epilogue (tcgen05.ld -> bias -> bf16 -> global, plus the BatchNorm
//   sum / sum-of-squares reduction for fprop).  Stages are handed over with mbarriers;
//   tcgen05.commit releases a stage the moment the MMAs that read it retire.
// * smem is sized so two CTAs fit per SM: one CTA's epilogue overlaps the other's main loop.
#include "../common.cuh"
#include "../../include/ddp_b200.h"
#include "../runtime/tensormap.h"

namespace ddpb {

constexpr int TILE_M = 128;          // pixels per CTA
constexpr int KCHUNK = 64;           // bf16 elements per 128-byte row
constexpr int A_BYTES = TILE_M * KCHUNK * 2;

struct ConvTcParams {
  int H, W;                // spatial size
  int tiles_per_img;       // >= 1 when H*W >= 128, else 0
  int imgs_per_tile;       // >= 1 when H*W <= 128
  int rows_per_tile;       // box height when H*W >= 128
  long long total_pixels;  // B*H*W (rows of the output matrix)
  int n_out;               // output channels (row pitch of `out`)
  int k_chunks;            // input channels / 64
  int k_total_cols;        // fprop: Cin (column stride between taps in the weight matrix)
};

template <int BN, int STAGES, bool DGRAD>
__global__ void __launch_bounds__(192)
conv3x3_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  bf16* __restrict__ out, const float* __restrict__ bias, float* __restrict__ stats,
                  const ConvTcParams prm) {
  constexpr int B_BYTES = BN * KCHUNK * 2;
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment is required by the 128-byte swizzle (TMA and UMMA agree on address bits)
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* accum_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);
  float* red = reinterpret_cast<float*>(tmem_slot + 2);   // [2*BN] per-CTA stats partials (fprop)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_tile = blockIdx.x, n_tile = blockIdx.y;

  // output tile -> (image, row) origin of the pixel box
  int n0, h0;
  if (prm.tiles_per_img > 0) { n0 = m_tile / prm.tiles_per_img; h0 = (m_tile % prm.tiles_per_img) * prm.rows_per_tile; }
  else { n0 = m_tile * prm.imgs_per_tile; h0 = 0; }
  const int k_iters = 9 * prm.k_chunks;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    mbar_init(accum_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<BN>(tmem_slot);
  if (!DGRAD) for (int i = threadIdx.x; i < 2 * BN; i += blockDim.x) red[i] = 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      for (int it = 0; it < k_iters; ++it) {
        const int stage = it % STAGES, use = it / STAGES;
        if (use > 0) mbar_wait(&empty_bar[stage], (use - 1) & 1);
        const int tap = it / prm.k_chunks, kc = it % prm.k_chunks;
        const int r = tap / 3, s = tap % 3;
        uint8_t* sa = smem + stage * STAGE_BYTES;
        uint8_t* sb = sa + A_BYTES;
        mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
        // fprop reads X at p + (r-1, s-1); dgrad reads dY at p - (r-1, s-1)
        const int dh = DGRAD ? 1 - r : r - 1, dw = DGRAD ? 1 - s : s - 1;
        tma_load_4d(sa, &tmA, &full_bar[stage], kc * KCHUNK, dw, h0 + dh, n0);
        if (!DGRAD) {
          // weights as a [Cout][9*Cin] matrix: BN rows (co), 64 contiguous K columns
          tma_load_2d(sb, &tmB, &full_bar[stage], tap * prm.k_total_cols + kc * KCHUNK, n_tile * BN);
        } else {
          // same matrix read "sideways": 64 rows of K (= co), BN/64 boxes of 64 contiguous ci
#pragma unroll
          for (int j = 0; j < BN / 64; ++j)
            tma_load_2d(sb + j * (64 * 128), &tmB, &full_bar[stage],
                        tap * prm.k_total_cols + n_tile * BN + j * 64, kc * KCHUNK);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread) =====================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(TILE_M, BN, false, DGRAD);
      for (int it = 0; it < k_iters; ++it) {
        const int stage = it % STAGES, use = it / STAGES;
        mbar_wait(&full_bar[stage], use & 1);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
        const uint32_t sb = sa + A_BYTES;
#pragma unroll
        for (int kk = 0; kk < KCHUNK / 16; ++kk) {
          // K-major, 128B swizzle: 16 K-elements = 32 bytes along the row; 8-row groups 1024 B apart
          const uint64_t da = umma_desc(sa + kk * 32, 16, 1024);
          // dgrad B is MN-major: 16 K-rows = 2048 bytes; 64-wide N blocks are 8192 B apart (LBO)
          const uint64_t db = DGRAD ? umma_desc(sb + kk * 2048, 64 * 128, 1024)
                                    : umma_desc(sb + kk * 32, 16, 1024);
          umma_bf16(tmem_base, da, db, idesc, (it | kk) != 0);
        }
        umma_commit(&empty_bar[stage]);        // frees the stage when these MMAs have read it
      }
      umma_commit(accum_bar);                  // accumulator complete
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int q = warp & 3;                    // TMEM lane quarter this warp may access
    const int row = q * 32 + lane;             // pixel row inside the tile
    const long long p = (long long)m_tile * TILE_M + row;
    const bool valid = p < prm.total_pixels;
    mbar_wait(accum_bar, 0);
    tc_fence_after();
    bf16* orow = out + p * prm.n_out + n_tile * BN;
#pragma unroll 1
    for (int c = 0; c < BN; c += 32) {
      float v[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + c, v);
      tmem_ld_wait();
      if (!DGRAD && bias) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] += __ldg(&bias[n_tile * BN + c + j]);
      }
      uint32_t pk[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) pk[j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
      if (valid) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          reinterpret_cast<uint4*>(orow + c)[j] = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
      }
      if (!DGRAD && stats) {
        // BatchNorm statistics of the ROUNDED outputs; column sums over the warp's 32 rows via a
        // halving butterfly (31 shuffles per 32 columns): lane l ends with column l's total.
        float a[32], b[32];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          float2 f = unpack_bf16x2(pk[j]);
          a[2 * j] = valid ? f.x : 0.f; a[2 * j + 1] = valid ? f.y : 0.f;
          b[2 * j] = a[2 * j] * a[2 * j]; b[2 * j + 1] = a[2 * j + 1] * a[2 * j + 1];
        }
#pragma unroll
        for (int step = 16, n = 32; step >= 1; step >>= 1, n >>= 1) {
          const bool upper = (lane & step) != 0;
#pragma unroll
          for (int i = 0; i < n / 2; ++i) {
            float sa_ = upper ? a[i] : a[i + n / 2], ka = upper ? a[i + n / 2] : a[i];
            float sb_ = upper ? b[i] : b[i + n / 2], kb = upper ? b[i + n / 2] : b[i];
            a[i] = ka + __shfl_xor_sync(0xffffffffu, sa_, step);
            b[i] = kb + __shfl_xor_sync(0xffffffffu, sb_, step);
          }
        }
        atomicAdd(&red[c + lane], a[0]);
        atomicAdd(&red[BN + c + lane], b[0]);
      }
    }
    if (!DGRAD && stats) {
      asm volatile("bar.sync 1, 128;" ::: "memory");     // the four epilogue warps only
      const int t = threadIdx.x - 64;
      for (int i = t; i < BN; i += 128) {
        atomicAdd(&stats[n_tile * BN + i], red[i]);
        atomicAdd(&stats[prm.n_out + n_tile * BN + i], red[BN + i]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<BN>(tmem_base);
}

template <int BN, int STAGES>
constexpr int conv_tc_smem() { return STAGES * (A_BYTES + BN * KCHUNK * 2) + (2 * STAGES + 1) * 8 + 16 + 2 * BN * 4 + 1024; }

// pixel box (w, h, n) holding `pixels` (128 or 64) consecutive NHWC pixels of an HxW image
void pixel_box(int H, int W, int pixels, int* bw, int* bh, int* bn) {
  if (H * W >= pixels) { *bw = W; *bh = pixels / W; *bn = 1; }
  else { *bw = W; *bh = H; *bn = pixels / (H * W); }
}

template <int BN, int STAGES, bool DGRAD>
static int launch(const bf16* in, const bf16* w, bf16* out, const float* bias, float* stats, int B, int H, int W,
                  int Cin, int Cout, cudaStream_t stream) {
  const int Kc = DGRAD ? Cout : Cin;     // reduction channels
  const int Nc = DGRAD ? Cin : Cout;     // output channels
  int bw, bh, bn;
  pixel_box(H, W, TILE_M, &bw, &bh, &bn);
  CUtensorMap tmA, tmB;
  int rc = make_tmap_nhwc_bf16(&tmA, in, B, H, W, Kc, KCHUNK, bw, bh, bn);
  if (rc) return rc;
  // weight matrix [Cout rows][9*Cin cols]; fprop box = 64 cols x BN rows, dgrad box = 64 cols x 64 rows
  rc = make_tmap_2d_bf16(&tmB, w, Cout, 9LL * Cin, KCHUNK, DGRAD ? 64 : BN);
  if (rc) return rc;
  ConvTcParams prm;
  prm.H = H; prm.W = W;
  prm.tiles_per_img = H * W >= TILE_M ? H * W / TILE_M : 0;
  prm.imgs_per_tile = H * W >= TILE_M ? 0 : TILE_M / (H * W);
  prm.rows_per_tile = bh;
  prm.total_pixels = (long long)B * H * W;
  prm.n_out = Nc;
  prm.k_chunks = Kc / KCHUNK;
  prm.k_total_cols = Cin;
  int m_tiles = prm.tiles_per_img > 0 ? B * prm.tiles_per_img : ceil_div(B, prm.imgs_per_tile);
  dim3 grid(m_tiles, Nc / BN);
  constexpr int smem = conv_tc_smem<BN, STAGES>();
  auto kern = conv3x3_tc_kernel<BN, STAGES, DGRAD>;
  static bool configured = false;
  if (!configured) {
    DDPB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  kern<<<grid, 192, smem, stream>>>(tmA, tmB, out, bias, stats, prm);
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}

static int check_shapes(int B, int H, int W, int Cin, int Cout) {
  DDPB_REQUIRE(B > 0 && H == W && (H == 32 || H == 16 || H == 8 || H == 4 || H == 2),
               "conv_tc: unsupported spatial size %dx%d", H, W);
  DDPB_REQUIRE(Cin % 64 == 0 && Cout % 64 == 0, "conv_tc: channels must be multiples of 64 (got %d, %d)", Cin,
               Cout);
  return DDPB_OK;
}

}  // namespace ddpb

using namespace ddpb;

// y[B,H,W,Cout] = conv3x3(x[B,H,W,Cin], w[Cout,3,3,Cin]) + bias; bf16 in/out, fp32 accumulate.
// `stats` (fp32 [2*Cout], may be null) ACCUMULATES per-channel sum and sum of squares of y.
extern "C" int ddpb_conv3x3_fwd_tc(const void* x, const void* w_bf16, const float* bias, void* y, float* stats,
                                   int B, int H, int W, int Cin, int Cout, int tile_n, cudaStream_t stream) {
  int rc = check_shapes(B, H, W, Cin, Cout);
  if (rc) return rc;
  if (tile_n == 0) tile_n = (Cout % 128 == 0 && (long long)B * H * W * Cout >= 128LL * 128 * 148) ? 128 : 64;
  DDPB_REQUIRE(Cout % tile_n == 0, "conv_tc fwd: Cout %d not a multiple of tile %d", Cout, tile_n);
  if (tile_n == 128)
    return launch<128, 3, false>((const bf16*)x, (const bf16*)w_bf16, (bf16*)y, bias, stats, B, H, W, Cin, Cout, stream);
  if (tile_n == 64)
    return launch<64, 4, false>((const bf16*)x, (const bf16*)w_bf16, (bf16*)y, bias, stats, B, H, W, Cin, Cout, stream);
  DDPB_REQUIRE(false, "conv_tc fwd: tile_n must be 64 or 128");
}

// dx[B,H,W,Cin] = conv3x3_transpose(dy[B,H,W,Cout], w[Cout,3,3,Cin]); bf16 in/out.
extern "C" int ddpb_conv3x3_dgrad_tc(const void* dy, const void* w_bf16, void* dx, int B, int H, int W, int Cin,
                                     int Cout, int tile_n, cudaStream_t stream) {
  int rc = check_shapes(B, H, W, Cin, Cout);
  if (rc) return rc;
  if (tile_n == 0) tile_n = (Cin % 128 == 0 && (long long)B * H * W * Cin >= 128LL * 128 * 148) ? 128 : 64;
  DDPB_REQUIRE(Cin % tile_n == 0, "conv_tc dgrad: Cin %d not a multiple of tile %d", Cin, tile_n);
  if (tile_n == 128)
    return launch<128, 3, true>((const bf16*)dy, (const bf16*)w_bf16, (bf16*)dx, nullptr, nullptr, B, H, W, Cin, Cout, stream);
  if (tile_n == 64)
    return launch<64, 4, true>((const bf16*)dy, (const bf16*)w_bf16, (bf16*)dx, nullptr, nullptr, B, H, W, Cin, Cout, stream);
  DDPB_REQUIRE(false, "conv_tc dgrad: tile_n must be 64 or 128");
}
