// conv3x3 weight gradient as a split-K implicit GEMM on tcgen05 tensor cores.
//
// Replaces autograd's conv weight gradient (ref Part 1/model.py:18-23 backward; SURVEY K-C3).
//
//   dW[co, tap, ci] = sum_p dY[p, co] * X[p + shift(tap), ci]          M = Cout, N = Cin (per tap), K = pixels
//
// The reduction dimension (pixels) is the SLOW dimension of both NHWC operands, so both are
// MN-major UMMA operands: a TMA box of 64 pixels x 64 channels is 64 rows of 128 B, which with the
// 128-byte swizzle is exactly the canonical MN-major layout (8-row K groups 1024 B apart).  The
// shifted X tile comes from the same trick as fprop: move the box by (s-1, r-1) and let TMA
// zero-fill what falls outside the image.
//
// One CTA owns: 128 output channels x 64 input channels x the three taps of one filter row
// (3 accumulators of 64 TMEM columns that share every dY tile), over one slice of the pixels
// (split-K).  Partial results are added to the fp32 gradient with TMA reduce-add
// (cp.reduce.async.bulk.tensor ... .add) straight into the gradient arena in KRSC order, i.e. the
// buffer the peer-to-peer reduce kernels and the fused SGD read - there is no staging copy.
#include <stdlib.h>

#include "../common.cuh"
#include "../../include/ddp_b200.h"
#include "../runtime/tensormap.h"

namespace ddpb {

constexpr int WG_M = 128;            // output channels per CTA
constexpr int WG_N = 64;             // input channels per CTA (per tap)
constexpr int WG_K = 64;             // pixels per pipeline stage
constexpr int WG_TAPS = 3;           // taps per CTA (one filter row)
constexpr int WG_BOX = WG_K * 128;   // bytes of one 64x64 bf16 box

struct WgradParams {
  int H, W;
  int chunks_per_img;   // >= 1 when H*W >= 64 else 0
  int imgs_per_chunk;   // >= 1 when H*W <= 64
  int rows_per_chunk;
  int total_chunks;     // ceil(B*H*W / 64)
  int Cin, Cout;
  float* dw_raw;        // M == 64 variant: gradient matrix [64][ld_raw], columns < cols_raw are accumulated
  int ld_raw, cols_raw;
  int pdl;              // 1: execute griddepcontrol.launch_dependents (captured launches only)
  int cl;               // M == 64 variant: cluster size along the split; ranks > 0 hand their partial tile to rank 0
};

void pixel_box(int H, int W, int pixels, int* bw, int* bh, int* bn);   // conv_tc.cu

// M = 128, TAPS = 3: the 3x3 layers.  M = 64, TAPS = 1: the first layer as a plain GEMM over its patch
// matrix (x = [pixels][64], 27 valid columns); the 64x27 result is added with scalar atomics.
template <int STAGES, int M, int TAPS>
__global__ void __launch_bounds__(192)
wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmX,
                const __grid_constant__ CUtensorMap tmW, const WgradParams prm) {
  constexpr int A_BOXES = M / 64;
  constexpr int WG_STAGE = (A_BOXES + TAPS) * WG_BOX;       // 3x3: 2 dY boxes + 3 X boxes = 40 KB
  constexpr int WG_TMEM_COLS = TAPS == 3 ? 256 : 64;        // TAPS x 64 used
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  // barriers sit behind whichever is larger: the pipeline stages or the epilogue's fp32 staging tile
  constexpr int EPI_BYTES = M == 128 ? TAPS * 2 * M * 128 : 0;
  constexpr int DATA_BYTES = STAGES * WG_STAGE > EPI_BYTES ? STAGES * WG_STAGE : EPI_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + DATA_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* accum_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);
  uint64_t* red_bar = reinterpret_cast<uint64_t*>(tmem_slot + 2);          // M == 64, cl > 1: partial tiles have arrived
  float* recv = reinterpret_cast<float*>(smem + DATA_BYTES + 512);         // [cl - 1][64 rows][32 cols]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int co_blk = blockIdx.x / (prm.Cin / WG_N), ci_blk = blockIdx.x % (prm.Cin / WG_N);
  const int r = TAPS == 3 ? blockIdx.y : 1;      // filter row (1x1: the centre)
  const int split = blockIdx.z, nsplit = gridDim.z;
  const int chunk_lo = (int)((long long)prm.total_chunks * split / nsplit);
  const int chunk_hi = (int)((long long)prm.total_chunks * (split + 1) / nsplit);
  const int k_iters = chunk_hi - chunk_lo;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmDY);
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmW);
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    mbar_init(accum_bar, 1);
    if (M == 64 && prm.cl > 1) mbar_init(red_bar, (prm.cl - 1) * 4);       // one arrival per epilogue warp of every peer
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<WG_TMEM_COLS>(tmem_slot);
  if (prm.pdl & 1) pdl_launch_dependents();     // programmatic dependent launch: see common.cuh
  pdl_wait();
  ktrace_begin(prm.pdl);      // stamped once the predecessor has finished: the start of this kernel's own work
  tc_fence_before();
  __syncthreads();
  if (M == 64 && prm.cl > 1) cluster_barrier();     // rank 0's red_bar exists before any peer can arrive on it
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
          uint8_t* s = smem + stage * WG_STAGE;
          mbar_expect_tx(&full_bar[stage], WG_STAGE);
#pragma unroll
          for (int a = 0; a < A_BOXES; ++a)
            tma_load_4d(s + a * WG_BOX, &tmDY, &full_bar[stage], co_blk * M + a * 64, 0, h0, n0);
#pragma unroll
          for (int t = 0; t < TAPS; ++t)
            tma_load_4d(s + (A_BOXES + t) * WG_BOX, &tmX, &full_bar[stage], ci_blk * WG_N, TAPS == 3 ? t - 1 : 0,
                        h0 + r - 1, n0);
        }
      }
    } else if (warp == 1) {
      if (lane == 0) {
        constexpr uint32_t idesc = umma_idesc_bf16(M, WG_N, true, true);
        for (int it = 0; it < k_iters; ++it) {
          const int stage = it % STAGES, use = it / STAGES;
          mbar_wait(&full_bar[stage], use & 1);
          tc_fence_after();
          const uint32_t s = smem_u32(smem + stage * WG_STAGE);
#pragma unroll
          for (int kk = 0; kk < WG_K / 16; ++kk) {
            // MN-major A: two 64-channel blocks WG_BOX apart (LBO), 8-pixel groups 1024 B apart (SBO)
            const uint64_t da = umma_desc(s + kk * 2048, WG_BOX, 1024);
#pragma unroll
            for (int t = 0; t < TAPS; ++t) {
              const uint64_t db = umma_desc(s + (A_BOXES + t) * WG_BOX + kk * 2048, WG_BOX, 1024);
              umma_bf16(tmem_base + t * WG_N, da, db, idesc, (it | kk) != 0);
            }
          }
          umma_commit(&empty_bar[stage]);
        }
        umma_commit(accum_bar);
      }
    } else if (M == 64) {
      // 64-row accumulator: row i lives in TMEM lane (i % 16) + 32 * (i / 16), i.e. the lower half of
      // each warp's lane quarter.  Only cols_raw (27) of the 64 columns are real.
      const int q = warp & 3;
      const int crank = prm.cl > 1 ? (int)cluster_rank() : 0;
      mbar_wait(accum_bar, 0);
      tc_fence_after();
      // cl > 1: the split-K partials of a cluster are summed in rank 0's shared memory first, so only one CTA in
      // `cl` ends with the 64 x cols atomic adds (they serialise on 1728 addresses and were the tail of this
      // kernel when more than one CTA per SM was tried)
      if (prm.cl > 1 && crank == 0) mbar_wait_cluster(red_bar, 0);
#pragma unroll 1
      for (int c = 0; c < (prm.cl > 1 ? 32 : WG_N); c += 32) {
        float v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + c, v);
        tmem_ld_wait();
        if (lane < 16) {
          const int row = q * 16 + lane;
          if (crank != 0) {
            const uint32_t remote = cluster_map_shared(smem_u32(recv + ((crank - 1) * 64 + row) * 32), 0u);
#pragma unroll
            for (int j = 0; j < 32; ++j) st_cluster_f32(remote + 4u * j, v[j]);
          } else {
            for (int sl = 0; sl < prm.cl - 1; ++sl) {
              const float* rp = recv + (sl * 64 + row) * 32;
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] += rp[j];
            }
            float* dst = prm.dw_raw + (size_t)(co_blk * 64 + row) * prm.ld_raw;
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (c + j < prm.cols_raw) atomicAdd(dst + c + j, v[j]);
          }
        }
      }
      if (crank != 0) {
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(cluster_map_shared(smem_u32(red_bar), 0u));
      }
    } else {
      // epilogue: TMEM -> registers -> shared memory (128B-swizzled [128 co][32 ci] fp32 boxes), then
      // six TMA reduce-add operations accumulate the partial tile into the gradient arena with
      // whole 128-byte lines (instead of 6144 scattered 16-byte RED instructions per CTA).
      const int q = warp & 3;
      const int row = q * 32 + lane;             // output channel inside the tile
      mbar_wait(accum_bar, 0);
      tc_fence_after();
#pragma unroll 1
      for (int t = 0; t < TAPS; ++t) {
#pragma unroll 1
        for (int c = 0; c < WG_N; c += 32) {
          float v[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + t * WG_N + c, v);
          tmem_ld_wait();
          uint8_t* rbase = smem + (t * 2 + (c >> 5)) * (WG_M * 128) + row * 128;
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(rbase + ((j ^ (row & 7)) << 4)) =
                make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        }
      }
      fence_async_proxy();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (threadIdx.x == 64) {
#pragma unroll
        for (int t = 0; t < TAPS; ++t)
#pragma unroll
          for (int hblk = 0; hblk < 2; ++hblk)
            tma_reduce_add_2d(&tmW, smem + (t * 2 + hblk) * (WG_M * 128),
                              (r * 3 + t) * prm.Cin + ci_blk * WG_N + hblk * 32, co_blk * WG_M);
        tma_store_commit();
        tma_store_wait_read();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<WG_TMEM_COLS>(tmem_base);
  ktrace_end(prm.pdl);
}

}  // namespace ddpb
DDPB_KTRACE_SETTER(wgrad_tc)

using namespace ddpb;

// dw (fp32 [Cout,3,3,Cin], KRSC) += dy^T * shifted(x).  dw is ACCUMULATED into (split-K partials
// and the caller's zeroed gradient arena).  split_k = 0 picks a split that fills ~148 SMs.
// `dy` points at the first of the `Cout` output channels handled here; dy_pitch = channels per pixel of the dY tensor
static int wgrad_tc_launch(const void* dy, int dy_pitch, const void* x, float* dw, int B, int H, int W, int Cin, int Cout,
                           int split_k, cudaStream_t stream) {
  DDPB_REQUIRE(B > 0 && H == W && (H == 32 || H == 16 || H == 8 || H == 4 || H == 2),
               "wgrad_tc: unsupported spatial size %dx%d", H, W);
  DDPB_REQUIRE(Cin % 64 == 0 && Cout % 128 == 0, "wgrad_tc: need Cin %% 64 == 0 and Cout %% 128 == 0 (got %d, %d)",
               Cin, Cout);
  int bw, bh, bn;
  pixel_box(H, W, WG_K, &bw, &bh, &bn);
  CUtensorMap tmDY, tmX, tmW;
  int rc = make_tmap_2d_f32(&tmW, dw, Cout, 9LL * Cin, 32, WG_M);
  if (rc) return rc;
  rc = make_tmap_nhwc_bf16(&tmDY, dy, B, H, W, Cout, 64, bw, bh, bn, dy_pitch);
  if (rc) return rc;
  rc = make_tmap_nhwc_bf16(&tmX, x, B, H, W, Cin, 64, bw, bh, bn);
  if (rc) return rc;
  WgradParams prm;
  prm.H = H; prm.W = W;
  prm.chunks_per_img = H * W >= WG_K ? H * W / WG_K : 0;
  prm.imgs_per_chunk = H * W >= WG_K ? 0 : WG_K / (H * W);
  prm.rows_per_chunk = bh;
  prm.total_chunks = prm.chunks_per_img > 0 ? B * prm.chunks_per_img : ceil_div(B, prm.imgs_per_chunk);
  prm.Cin = Cin; prm.Cout = Cout;
  int base_ctas = (Cout / WG_M) * (Cin / WG_N) * 3;
  // Pipeline depth: a CTA gets only about one stage's worth of TMA loads in flight whatever the depth
  // (tools/tma_probe.py), so two stages lose nothing against four - and at ~98 KB (the epilogue's fp32
  // staging tile sets the floor) two CTAs, or one of these and one dgrad CTA, share an SM instead of
  // this kernel evicting the critical path's kernels from every SM it touches.
  static int stages = 0;
  if (!stages) {
    const char* e = getenv("DDPB_WGRAD_STAGES");
    stages = e ? atoi(e) : 2;
    if (stages != 2 && stages != 3 && stages != 4) stages = 2;
  }
  const int per_sm = stages == 2 ? 2 : 1;
  if (split_k <= 0) {
    // pick the split that minimises (waves) x (k-iterations per CTA + fixed cost)
    long long best_cost = -1;
    const int max_split = prm.total_chunks < 64 ? prm.total_chunks : 64;
    for (int sp = 1; sp <= max_split; ++sp) {
      long long waves = ((long long)base_ctas * sp + sm_count() * per_sm - 1) / (sm_count() * per_sm);
      long long iters = (prm.total_chunks + sp - 1) / sp;
      long long cost = waves * (iters * per_sm + 6);        // CTAs sharing an SM share its load bandwidth
      if (best_cost < 0 || cost < best_cost) { best_cost = cost; split_k = sp; }
    }
  }
  if (split_k > prm.total_chunks) split_k = prm.total_chunks;
  prm.dw_raw = nullptr; prm.ld_raw = 0; prm.cols_raw = 0; prm.cl = 1;
  prm.pdl = kernel_flags(stream, 2);
  dim3 grid(base_ctas / 3, 3, split_k);
  constexpr int EPI_BYTES = WG_TAPS * 2 * WG_M * 128;       // fp32 staging of the epilogue (96 KB)
#define DDPB_WGRAD_LAUNCH(ST)                                                                                  \
  do {                                                                                                         \
    constexpr int pipe = ST * (WG_M / 64 + WG_TAPS) * WG_BOX;                                                  \
    constexpr int smem = (pipe > EPI_BYTES ? pipe : EPI_BYTES) + (2 * ST + 1) * 8 + 16 + 1024;                 \
    auto kern = wgrad_tc_kernel<ST, WG_M, WG_TAPS>;                                                            \
    static unsigned long long configured = 0;                                                                            \
    if (first_on_device(&configured)) {                                                                                         \
      DDPB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));          \
      DDPB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout,               \
                                           cudaSharedmemCarveoutMaxShared));                                   \
    }                                                                                                          \
    DDPB_CUDA_CHECK(launch_ex(kern, grid, dim3(192), smem, stream, dim3(1, 1, 1), 2, tmDY, tmX, tmW, prm));   \
  } while (0)
  if (stages == 2) DDPB_WGRAD_LAUNCH(2);
  else if (stages == 3) DDPB_WGRAD_LAUNCH(3);
  else DDPB_WGRAD_LAUNCH(4);
#undef DDPB_WGRAD_LAUNCH
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}

extern "C" int ddpb_conv3x3_wgrad_tc(const void* dy, const void* x, float* dw, int B, int H, int W, int Cin,
                                     int Cout, int split_k, cudaStream_t stream) {
  return wgrad_tc_launch(dy, Cout, x, dw, B, H, W, Cin, Cout, split_k, stream);
}

// The same for the output channels [co_lo, co_lo + co_n) only (multiples of 128): rows co_lo.. of dw.  The engine
// launches a large layer's weight gradient as two such halves so that the first half of the layer's bucket can be
// reduced across the GPUs while the second half is still being computed (sub-bucket hand-off).
extern "C" int ddpb_conv3x3_wgrad_tc_range(const void* dy, const void* x, float* dw, int B, int H, int W, int Cin,
                                           int Cout, int co_lo, int co_n, int split_k, cudaStream_t stream) {
  DDPB_REQUIRE(co_lo >= 0 && co_n > 0 && co_lo % 128 == 0 && co_n % 128 == 0 && co_lo + co_n <= Cout,
               "wgrad_tc_range: bad channel range [%d, %d) of %d", co_lo, co_lo + co_n, Cout);
  return wgrad_tc_launch((const bf16*)dy + co_lo, Cout, x, dw + (size_t)co_lo * 9 * Cin, B, H, W, Cin, co_n, split_k, stream);
}

// dw (fp32 [64][ld], columns < cols) += dy[pixels][64]^T * x[pixels][64]: the first layer's weight
// gradient from its patch matrix (ddpb_im2col3x3_c3), split over ~148 CTAs along the pixels.
extern "C" int ddpb_patch_wgrad_tc(const void* dy, const void* x_cols, float* dw, int B, int H, int W, int Cout,
                                   int ld, int cols, int split_k, cudaStream_t stream) {
  DDPB_REQUIRE(B > 0 && H == W && (H == 32 || H == 16 || H == 8), "patch_wgrad_tc: unsupported spatial size %dx%d", H, W);
  DDPB_REQUIRE(Cout % 64 == 0 && cols >= 1 && cols <= 64 && ld >= cols, "patch_wgrad_tc: bad shape (Cout %d, cols %d, ld %d)",
               Cout, cols, ld);
  int bw, bh, bn;
  pixel_box(H, W, WG_K, &bw, &bh, &bn);
  CUtensorMap tmDY, tmX;
  int rc = make_tmap_nhwc_bf16(&tmDY, dy, B, H, W, Cout, 64, bw, bh, bn);
  if (rc) return rc;
  rc = make_tmap_nhwc_bf16(&tmX, x_cols, B, H, W, 64, 64, bw, bh, bn);
  if (rc) return rc;
  WgradParams prm;
  prm.H = H; prm.W = W;
  prm.chunks_per_img = H * W / WG_K;
  prm.imgs_per_chunk = 0;
  prm.rows_per_chunk = bh;
  prm.total_chunks = B * prm.chunks_per_img;
  prm.Cin = 64; prm.Cout = Cout;
  prm.dw_raw = dw; prm.ld_raw = ld; prm.cols_raw = cols;
  prm.pdl = kernel_flags(stream, 2);
  const int base_ctas = Cout / 64;
  // CTA count: more CTAs per SM put more loads in flight (both operands stream from HBM), but every CTA
  // ends with 64 x 27 atomic adds on the same addresses; measured best around one CTA per SM
  static int ctas = 0, cl = 0;
  if (!ctas) {
    const char* e = getenv("DDPB_WGRAD0_CTAS");
    ctas = e ? atoi(e) : sm_count();
    if (ctas < 1 || ctas > 1024) ctas = sm_count();
    const char* c = getenv("DDPB_WGRAD0_CLUSTER");
    cl = c ? atoi(c) : 1;
    if (cl != 1 && cl != 2 && cl != 4) cl = 1;
  }
  if (split_k <= 0) split_k = ctas / base_ctas;
  int cluster = cols <= 32 ? cl : 1;
  if (split_k % cluster) split_k = split_k / cluster * cluster;
  if (split_k < cluster) cluster = 1;
  prm.cl = cluster;
  if (split_k > prm.total_chunks) split_k = prm.total_chunks;
  if (split_k < 1) split_k = 1;
  dim3 grid(base_ctas, 1, split_k);
  static int stages = 0;
  if (!stages) {
    const char* e = getenv("DDPB_WGRAD0_STAGES");
    stages = e ? atoi(e) : 8;
    if (stages != 2 && stages != 4 && stages != 8 && stages != 12) stages = 8;
  }
#define DDPB_WGRAD0_LAUNCH(ST)                                                                                 \
  do {                                                                                                         \
    constexpr int smem = ST * 2 * WG_BOX + 512 + 3 * 64 * 32 * 4 + 1024;   /* stages, barriers, receive area */   \
    auto kern = wgrad_tc_kernel<ST, 64, 1>;                                                                    \
    static unsigned long long configured = 0;                                                                            \
    if (first_on_device(&configured)) {                                                                                         \
      DDPB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));          \
      DDPB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout,               \
                                           cudaSharedmemCarveoutMaxShared));                                   \
    }                                                                                                          \
    DDPB_CUDA_CHECK(launch_ex(kern, grid, dim3(192), smem, stream, dim3(1, 1, cluster), 2, tmDY, tmX, tmDY, prm)); \
  } while (0)
  // both operands stream from HBM (~1 us latency): here a deep ring does pay (8 stages 27 us, 2 stages 35 us)
  if (stages == 2) DDPB_WGRAD0_LAUNCH(2);
  else if (stages == 4) DDPB_WGRAD0_LAUNCH(4);
  else if (stages == 12) DDPB_WGRAD0_LAUNCH(12);
  else DDPB_WGRAD0_LAUNCH(8);
#undef DDPB_WGRAD0_LAUNCH
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}
