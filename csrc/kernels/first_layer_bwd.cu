// Backward of the first VGG block in ONE kernel: BatchNorm + ReLU + 2x2 max-pool backward (apply pass)
// fused with the weight gradient of the 3 -> 64 convolution.
//
// Replaces, for block 0, autograd's backward of Conv2d(3, 64) -> BatchNorm2d -> ReLU -> MaxPool2d
// (ref Part 1/model.py:16,18-25; SURVEY K-N2 + K-C3).  Block 0 has no data gradient (its input is the
// image), so the gradient dY of its convolution output - 33.5 MB at batch 256, the largest tensor of the
// whole backward pass - is consumed by exactly one reader, the weight-gradient GEMM
//     dW[co, k] = sum_p dY[p, co] * cols[p, k]          (cols = the 3x3x3 patch matrix, ddpb_im2col3x3_c3).
// The two separate kernels (ddpb_bn_relu_pool_bwd_apply, then ddpb_patch_wgrad_tc) wrote dY to HBM and read
// it back, and sat one after the other at the very end of the step, where nothing overlaps them.  Here dY
// never leaves the SM:
//   warp 0      TMA: per chunk of 64 pixels (two image rows = 16 complete pool windows) the conv output y,
//               the pooled incoming gradient and the patch-matrix tile, 2 stages deep (template parameter)
//   warps 2-5   one thread per (pool window, 8 channels): recomputes the ReLU / arg-max routing from y,
//               applies dy = sc * dz + cb - cy * y and writes the 64 x 64 bf16 tile straight into shared
//               memory in the 128-byte-swizzled MN-major layout tcgen05.mma reads (what TMA would have
//               produced had dY come from global memory); also accumulates the conv-bias gradient
//   warp 1      one thread: 4 x tcgen05.mma (M = 64 channels, N = 64 patch columns, K = 16 pixels) per chunk
//               into one TMEM accumulator; tcgen05.commit frees the stage and the dY buffer
// and at the end the 64 x 27 partial result is added to the gradient arena.
#include "../common.cuh"
#include "../../include/ddp_b200.h"
#include "../runtime/tensormap.h"

namespace ddpb {

constexpr int FL_C = 64;                 // channels of block 0
constexpr int FL_PIX = 64;               // pixels per chunk: two rows of a 32-wide image
constexpr int FL_WIN = 16;               // pool windows per chunk
constexpr int FL_STAGES_DEFAULT = 2;     // pipeline depth ST of the kernel template (DDPB_FL_STAGES = 2 | 3 | 4)
constexpr int FL_Y = FL_PIX * 128;       // y tile: 64 rows x 128 B
constexpr int FL_D = FL_WIN * 128;       // pooled gradient tile: 16 rows x 128 B
constexpr int FL_X = FL_PIX * 128;       // patch tile: 64 rows x 128 B
constexpr int FL_STAGE = FL_Y + FL_X + FL_D;          // 18 KB (every part a multiple of 1024 B)
constexpr int FL_ABUF = 2;               // dY tiles in flight between the transform warps and the MMA
constexpr int FL_A = FL_PIX * 128;

struct FirstBwdParams {
  long long chunks;        // B * 1024 / 64
  float inv_p;             // 1 / (B * 1024)
  const float *mean, *invstd, *gamma, *beta, *sums;
  float *dw, *dgamma, *dbeta, *dbias;
  int ld, cols;            // dw row pitch and number of real patch columns (27)
  int flags;               // launch flags (common.cuh): bits 0-1 programmatic launch trigger, bits 2.. trace id
  int bulk_add;            // 1: dw is contiguous (ld == cols), 16-byte aligned: one bulk reduce-add per CTA
};

template <int FL_STAGES>
__global__ void __launch_bounds__(192)
first_layer_bwd_kernel(const __grid_constant__ CUtensorMap tmY, const __grid_constant__ CUtensorMap tmD,
                       const __grid_constant__ CUtensorMap tmX, const FirstBwdParams prm) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem + FL_STAGES * FL_STAGE;                                   // [FL_ABUF][64 x 128 B]
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(sA + FL_ABUF * FL_A);
  uint64_t* empty_bar = full_bar + FL_STAGES;
  uint64_t* afull_bar = empty_bar + FL_STAGES;
  uint64_t* aempty_bar = afull_bar + FL_ABUF;
  uint64_t* accum_bar = aempty_bar + FL_ABUF;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);
  float* red = reinterpret_cast<float*>(tmem_slot + 2);                        // [16 windows][64] conv-bias partials

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long lo = prm.chunks * blockIdx.x / gridDim.x, hi = prm.chunks * (blockIdx.x + 1) / gridDim.x;
  const int n_iter = (int)(hi - lo);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmY); tma_prefetch_desc(&tmD); tma_prefetch_desc(&tmX);
    for (int i = 0; i < FL_STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 5); }   // 4 transform warps + the MMA commit
    for (int i = 0; i < FL_ABUF; ++i) { mbar_init(&afull_bar[i], 4); mbar_init(&aempty_bar[i], 1); }
    mbar_init(accum_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<64>(tmem_slot);
  if (prm.flags & 1) pdl_launch_dependents();
  pdl_wait();
  ktrace_begin(prm.flags);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      for (int it = 0; it < n_iter; ++it) {
        const int stage = it % FL_STAGES, use = it / FL_STAGES;
        if (use > 0) mbar_wait(&empty_bar[stage], (use - 1) & 1);
        const long long chunk = lo + it;
        uint8_t* s = smem + stage * FL_STAGE;
        mbar_expect_tx(&full_bar[stage], FL_STAGE);
        tma_load_2d(s, &tmY, &full_bar[stage], 0, (int)(chunk * FL_PIX));
        tma_load_2d(s + FL_Y, &tmX, &full_bar[stage], 0, (int)(chunk * FL_PIX));
        tma_load_2d(s + FL_Y + FL_X, &tmD, &full_bar[stage], 0, (int)(chunk * FL_WIN));
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(64, 64, true, true);
      for (int it = 0; it < n_iter; ++it) {
        const int stage = it % FL_STAGES, ab = it % FL_ABUF;
        mbar_wait(&full_bar[stage], (it / FL_STAGES) & 1);        // the patch tile has landed
        mbar_wait(&afull_bar[ab], (it / FL_ABUF) & 1);            // the dY tile has been written
        tc_fence_after();
        const uint32_t a = smem_u32(sA + ab * FL_A), b = smem_u32(smem + stage * FL_STAGE + FL_Y);
#pragma unroll
        for (int kk = 0; kk < FL_PIX / 16; ++kk)
          // both operands MN-major: 16 pixels (K) = 2048 bytes; one 64-wide block each (LBO unused), 8-pixel groups 1024 B apart
          umma_bf16(tmem_base, umma_desc(a + kk * 2048, FL_A, 1024), umma_desc(b + kk * 2048, FL_X, 1024), idesc,
                    (it | kk) != 0);
        umma_commit(&empty_bar[stage]);
        umma_commit(&aempty_bar[ab]);
      }
      umma_commit(accum_bar);
    }
  } else {
    // ---- transform warps: thread = (pool window w, channel group g of 8) ----
    const int t = threadIdx.x - 64, w = t >> 3, g = t & 7, c0 = g * 8;
    float sc[8], sh[8], cb[8], cy[8], ab_sum[8];
    {
      float mu[8], is[8], s1[8], s2[8];
      ld_coh_vec<8>(prm.gamma + c0, sc); ld_coh_vec<8>(prm.beta + c0, sh);
      ld_coh_vec<8>(prm.mean + c0, mu); ld_coh_vec<8>(prm.invstd + c0, is);
      ld_coh_vec<8>(prm.sums + c0, s1); ld_coh_vec<8>(prm.sums + FL_C + c0, s2);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        sc[j] *= is[j];
        sh[j] -= mu[j] * sc[j];
        cy[j] = sc[j] * is[j] * (s2[j] * prm.inv_p);
        cb[j] = cy[j] * mu[j] - sc[j] * (s1[j] * prm.inv_p);
        ab_sum[j] = 0.f;
        if (blockIdx.x == 0 && w == 0) { prm.dbeta[c0 + j] = s1[j]; prm.dgamma[c0 + j] = s2[j]; }
      }
    }
    // the four pixels of this thread's window inside the chunk (two rows of 32): same order as the BN kernels
    const int pix[4] = {2 * w, 2 * w + 1, 32 + 2 * w, 32 + 2 * w + 1};
    for (int it = 0; it < n_iter; ++it) {
      const int stage = it % FL_STAGES, ab = it % FL_ABUF;
      mbar_wait(&full_bar[stage], (it / FL_STAGES) & 1);
      if (it >= FL_ABUF) mbar_wait(&aempty_bar[ab], ((it / FL_ABUF) - 1) & 1);
      const uint8_t* sy = smem + stage * FL_STAGE;
      const uint8_t* sd = sy + FL_Y + FL_X;
      uint8_t* sa = sA + ab * FL_A;
      float d[8], v[4][8];
      { const uint4 u = *reinterpret_cast<const uint4*>(sd + w * 128 + ((g ^ (w & 7)) << 4));
        const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), e = unpack_bf16x2(u.w);
        d[0] = a.x; d[1] = a.y; d[2] = b.x; d[3] = b.y; d[4] = c.x; d[5] = c.y; d[6] = e.x; d[7] = e.y; }
      float bz[8];
      int bq[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) { bz[j] = 0.f; bq[j] = -1; }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int p = pix[q];
        const uint4 u = *reinterpret_cast<const uint4*>(sy + p * 128 + ((g ^ (p & 7)) << 4));
        const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), e = unpack_bf16x2(u.w);
        v[q][0] = a.x; v[q][1] = a.y; v[q][2] = b.x; v[q][3] = b.y; v[q][4] = c.x; v[q][5] = c.y; v[q][6] = e.x; v[q][7] = e.y;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float z = fmaf(v[q][j], sc[j], sh[j]);
          if (z > bz[j]) { bz[j] = z; bq[j] = q; }          // strict >: the first maximum wins, z <= 0 never does
        }
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int p = pix[q];
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float dz = (bq[j] == q) ? d[j] : 0.f;
          o[j] = fmaf(sc[j], dz, fmaf(-cy[j], v[q][j], cb[j]));
          ab_sum[j] += o[j];
        }
        *reinterpret_cast<uint4*>(sa + p * 128 + ((g ^ (p & 7)) << 4)) =
            make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
      }
      fence_async_proxy();             // the tile was written with ordinary stores; the tensor core reads it through the async proxy
      __syncwarp();
      if (lane == 0) { mbar_arrive(&afull_bar[ab]); mbar_arrive(&empty_bar[stage]); }
    }
    // conv-bias gradient: column sums of dY over this CTA's pixels
#pragma unroll
    for (int j = 0; j < 8; ++j) red[w * FL_C + c0 + j] = ab_sum[j];
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (t < FL_C && prm.dbias) {
      float s = 0.f;
#pragma unroll
      for (int k = 0; k < FL_WIN; ++k) s += red[k * FL_C + t];
      atomicAdd(&prm.dbias[t], s);
    }
    // weight gradient: 64-row accumulator, row i in TMEM lane (i % 16) + 32 * (i / 16)
    if (n_iter > 0) {
      const int q = warp & 3;
      mbar_wait(accum_bar, 0);
      tc_fence_after();
      float acc[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16), acc);
      tmem_ld_wait();
      if (prm.bulk_add) {
        // the 64 x cols partial result goes to the gradient arena as ONE bulk reduce-add from shared memory (every
        // stage of the ring is free by now: the accumulator barrier follows the last MMA) instead of 64 x cols
        // RED instructions per CTA - 296 CTAs x 1728 atomics on the same 54 lines throttled the LSU
        float* stg = reinterpret_cast<float*>(smem);
        if (lane < 16) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < prm.cols) stg[(q * 16 + lane) * prm.cols + j] = acc[j];
        }
        fence_async_proxy();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (t == 0) {
          bulk_reduce_add_f32(prm.dw, stg, (uint32_t)(FL_C * prm.cols * 4));
          tma_store_commit();
          tma_store_wait_all();
        }
      } else if (lane < 16) {
        float* dst = prm.dw + (size_t)(q * 16 + lane) * prm.ld;
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (j < prm.cols) atomicAdd(dst + j, acc[j]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<64>(tmem_base);
  ktrace_end(prm.flags);
}

}  // namespace ddpb
DDPB_KTRACE_SETTER(first_layer_bwd)

using namespace ddpb;

// Block 0 backward in one launch.  dout: gradient of the block's (pooled) output [B,16,16,64] bf16; y: the block's
// conv output [B,32,32,64] bf16; x_cols: its patch matrix [B*1024][64] bf16 (ddpb_im2col3x3_c3); sums: [128] the
// BatchNorm-backward reduction (sum dz, sum dz * xhat), e.g. from ddpb_conv3x3_dgrad_bnred_tc.  Outputs: dw (fp32
// [64][ld], columns < cols) and dbias (nullable) are ACCUMULATED, dgamma / dbeta are written.
extern "C" int ddpb_first_layer_bwd_fused(const void* dout, const void* y, const void* x_cols, const float* save_mean,
                                          const float* save_invstd, const float* gamma, const float* beta,
                                          const float* sums, float* dw, float* dgamma, float* dbeta, float* dbias,
                                          int B, int H, int W, int C, int ld, int cols, int ctas, cudaStream_t stream) {
  DDPB_REQUIRE(B > 0 && H == 32 && W == 32 && C == FL_C, "first_layer_bwd: needs 32x32 images and 64 channels (got %dx%d, %d)",
               H, W, C);
  DDPB_REQUIRE(cols >= 1 && cols <= 32 && ld >= cols, "first_layer_bwd: bad weight-gradient shape (cols %d, ld %d)", cols, ld);
  DDPB_REQUIRE(dout && y && x_cols && save_mean && save_invstd && gamma && beta && sums && dw && dgamma && dbeta,
               "first_layer_bwd: null pointer");
  const long long pixels = (long long)B * H * W;
  CUtensorMap tmY, tmD, tmX;
  int rc = make_tmap_2d_bf16(&tmY, y, pixels, FL_C, 64, FL_PIX);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tmD, dout, pixels / 4, FL_C, 64, FL_WIN);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tmX, x_cols, pixels, 64, 64, FL_PIX);
  if (rc) return rc;
  FirstBwdParams prm;
  prm.chunks = pixels / FL_PIX;
  prm.inv_p = 1.f / (float)pixels;
  prm.mean = save_mean; prm.invstd = save_invstd; prm.gamma = gamma; prm.beta = beta; prm.sums = sums;
  prm.dw = dw; prm.dgamma = dgamma; prm.dbeta = dbeta; prm.dbias = dbias;
  prm.ld = ld; prm.cols = cols;
  static const bool bulk_ok = !(getenv("DDPB_FL_BULK") && atoi(getenv("DDPB_FL_BULK")) == 0);
  prm.bulk_add = (bulk_ok && ld == cols && (FL_C * cols * 4) % 16 == 0 && reinterpret_cast<uintptr_t>(dw) % 16 == 0) ? 1 : 0;
  // Everything streams from HBM once.  Measured in the step (bench.py, 200 steps; stages x CTAs per SM): 4 x 2
  // 0.5111 ms, 3 x 2 0.5109, 2 x 2 0.5080, 2 x 3 0.5092, 2 x 4 0.5065 - the transform warps, not the loads, are the
  // bottleneck, so more (smaller) CTAs win: two stages (58 KB, three CTAs resident per SM) and a grid of 4 per SM.
  // DDPB_FL_STAGES / DDPB_FL_PER_SM override.
  static int stages = 0, per_sm = 0;
  if (!stages) {
    const char* e = getenv("DDPB_FL_STAGES");
    stages = e ? atoi(e) : FL_STAGES_DEFAULT;
    if (stages < 2 || stages > 4) stages = FL_STAGES_DEFAULT;
    const char* q = getenv("DDPB_FL_PER_SM");
    per_sm = q ? atoi(q) : 4;
    if (per_sm < 1 || per_sm > 4) per_sm = 4;
  }
  if (ctas <= 0) ctas = per_sm * sm_count();
  if (ctas > prm.chunks) ctas = (int)prm.chunks;
  prm.flags = kernel_flags(stream, 16);
#define DDPB_FL_LAUNCH(ST)                                                                                            \
  do {                                                                                                                \
    constexpr int smem = ST * FL_STAGE + FL_ABUF * FL_A + (2 * ST + 2 * FL_ABUF + 1) * 8 + 16 + FL_WIN * FL_C * 4 + 1024; \
    static unsigned long long configured = 0;                                                                         \
    if (first_on_device(&configured)) {                                                                               \
      DDPB_CUDA_CHECK(cudaFuncSetAttribute(first_layer_bwd_kernel<ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); \
      DDPB_CUDA_CHECK(cudaFuncSetAttribute(first_layer_bwd_kernel<ST>, cudaFuncAttributePreferredSharedMemoryCarveout,     \
                                           cudaSharedmemCarveoutMaxShared));                                          \
    }                                                                                                                 \
    DDPB_CUDA_CHECK(launch_ex(first_layer_bwd_kernel<ST>, dim3(ctas), dim3(192), smem, stream, dim3(1, 1, 1), 16, tmY, tmD, \
                              tmX, prm));                                                                             \
  } while (0)
  if (stages == 2) DDPB_FL_LAUNCH(2);
  else if (stages == 3) DDPB_FL_LAUNCH(3);
  else DDPB_FL_LAUNCH(4);
#undef DDPB_FL_LAUNCH
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}
