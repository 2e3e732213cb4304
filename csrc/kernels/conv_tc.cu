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
#include <stdlib.h>

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
  int slice_dim;           // NC > 1: tensor dimension (2 = h, 3 = n) along which the A box is cut into NC slices
  int slice_extent;        // NC > 1: extent of one slice along slice_dim
  int taps;                // 9 = 3x3 filter, 1 = 1x1 (plain GEMM over the pixel rows; first-layer patch matrix)
  bf16* out_raw;           // KS > 1: output matrix [total_pixels][n_out] (written with plain stores)
};

// NC > 1: the NC CTAs of a cluster (consecutive n-tiles, same pixel tile) need the SAME activation
// tile.  Each loads 1/NC of it and multicasts that slice into all NC shared memories, so the tile is
// read from L2 once per cluster instead of once per CTA (the kernels are bound by L2 -> SM traffic).
//
// KS > 1 (cluster split-K): the deep layers have few output tiles (64 for the 2x2 layers) but a long
// reduction (K = 4608), and one CTA cannot pull more than ~30-50 B/clk through TMA.  KS CTAs of a
// cluster (blockIdx.z) each accumulate 1/KS of the K range of the SAME tile; then every CTA sends the
// other CTAs' column blocks of its fp32 partial tile through distributed shared memory and finishes
// BN/KS columns itself (reduce-scatter), so the epilogue work is split KS ways too.
template <int BN, int STAGES, bool DGRAD, int NC, int KS>
__global__ void __launch_bounds__(192)
conv3x3_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const __grid_constant__ CUtensorMap tmC, const float* __restrict__ bias,
                  float* __restrict__ stats, const ConvTcParams prm) {
  constexpr int B_BYTES = BN * KCHUNK * 2;
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment is required by the 128-byte swizzle (TMA and UMMA agree on address bits)
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* accum_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);
  float* bias_s = reinterpret_cast<float*>(tmem_slot + 2);   // [BN] bias of this tile's channels (fprop)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_tile = blockIdx.x, n_tile = blockIdx.y;

  // output tile -> (image, row) origin of the pixel box
  int n0, h0;
  if (prm.tiles_per_img > 0) { n0 = m_tile / prm.tiles_per_img; h0 = (m_tile % prm.tiles_per_img) * prm.rows_per_tile; }
  else { n0 = m_tile * prm.imgs_per_tile; h0 = 0; }
  const int k_total = prm.taps * prm.k_chunks;
  const int ks_rank = KS > 1 ? (int)blockIdx.z : 0;
  const int it_lo = KS > 1 ? k_total * ks_rank / KS : 0;
  const int k_iters = (KS > 1 ? k_total * (ks_rank + 1) / KS : k_total) - it_lo;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
    // a stage is written by all NC CTAs (multicast slices), so it is free only when all NC have consumed it
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], NC); }
    mbar_init(accum_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<BN>(tmem_slot);
  // everything above is private to this CTA and overlaps the previous kernel's tail (programmatic
  // dependent launch); global memory may only be touched after pdl_wait()
  pdl_launch_dependents();
  pdl_wait();
  if (!DGRAD) for (int i = threadIdx.x; i < BN; i += blockDim.x) bias_s[i] = bias ? ld_coh(&bias[n_tile * BN + i]) : 0.f;
  tc_fence_before();
  __syncthreads();
  if (NC > 1) cluster_barrier();      // every CTA's barriers exist before a peer's multicast can touch them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  constexpr uint16_t MC_MASK = (uint16_t)((1u << NC) - 1);

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      for (int it = 0; it < k_iters; ++it) {
        const int stage = it % STAGES, use = it / STAGES;
        if (use > 0) mbar_wait(&empty_bar[stage], (use - 1) & 1);
        const int tap = (it_lo + it) / prm.k_chunks, kc = (it_lo + it) % prm.k_chunks;
        const int r = prm.taps == 1 ? 1 : tap / 3, s = prm.taps == 1 ? 1 : tap % 3;
        uint8_t* sa = smem + stage * STAGE_BYTES;
        uint8_t* sb = sa + A_BYTES;
        mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
        // fprop reads X at p + (r-1, s-1); dgrad reads dY at p - (r-1, s-1)
        const int dh = DGRAD ? 1 - r : r - 1, dw = DGRAD ? 1 - s : s - 1;
        if (NC == 1) {
          tma_load_4d(sa, &tmA, &full_bar[stage], kc * KCHUNK, dw, h0 + dh, n0);
        } else {
          // tmA's box is one slice here; slice `cr` of the tile goes to every CTA of the cluster
          const int cr = (int)cluster_rank();
          const int hh = h0 + dh + (prm.slice_dim == 2 ? cr * prm.slice_extent : 0);
          const int nn = n0 + (prm.slice_dim == 3 ? cr * prm.slice_extent : 0);
          tma_load_4d_mc(sa + cr * (A_BYTES / NC), &tmA, &full_bar[stage], kc * KCHUNK, dw, hh, nn, MC_MASK);
        }
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
        // frees the stage when these MMAs have read it (in every CTA that multicasts into it)
        if (NC == 1) umma_commit(&empty_bar[stage]); else umma_commit_mc(&empty_bar[stage], MC_MASK);
      }
      umma_commit(accum_bar);                  // accumulator complete
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    // 1. TMEM -> registers -> (+bias) -> bf16 -> shared memory, in the 128B-swizzled layout a TMA
    //    store expects (the pipeline stages are free once the accumulator barrier has fired).
    // 2. one thread stores the tile with TMA (coalesced, clipped at the tensor edge, asynchronous).
    // 3. fprop: per-channel sum / sum-of-squares of the ROUNDED tile, read back from shared memory
    //    (8-byte column-group loads; no shuffles, no shared-memory atomics).
    const int q = warp & 3;                    // TMEM lane quarter this warp may access
    const int row = q * 32 + lane;             // pixel row inside the tile
    const int t = threadIdx.x - 64;            // 0..127 among the epilogue threads
    uint8_t* tile = smem;                      // [BN/64][128 rows][128 B]
    float* red = reinterpret_cast<float*>(smem + 32768);      // [parts][2*BN]
    mbar_wait(accum_bar, 0);
    tc_fence_after();
    if constexpr (KS == 1) {
#pragma unroll 1
    for (int c = 0; c < BN; c += 32) {
      float v[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + c, v);
      tmem_ld_wait();
      if (!DGRAD) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] += bias_s[c + j];
      }
      uint8_t* rbase = tile + (c >> 6) * (TILE_M * 128) + row * 128;
      const int j0 = (c & 63) >> 3;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint4 pk = make_uint4(pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                              pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
        *reinterpret_cast<uint4*>(rbase + (((j0 + j) ^ (row & 7)) << 4)) = pk;
      }
    }
    fence_async_proxy();                                       // generic-proxy writes -> visible to TMA
    asm volatile("bar.sync 1, 128;" ::: "memory");             // the four epilogue warps only
    if (t == 0) {
#pragma unroll
      for (int hblk = 0; hblk < BN / 64; ++hblk)
        tma_store_2d(&tmC, tile + hblk * (TILE_M * 128), n_tile * BN + hblk * 64, m_tile * TILE_M);
      tma_store_commit();
    }
    if (!DGRAD && stats) {
      constexpr int GROUPS = BN / 4, PARTS = 128 / GROUPS, ROWS = TILE_M / PARTS;
      const int g = t % GROUPS, part = t / GROUPS;
      const long long left = prm.total_pixels - (long long)m_tile * TILE_M;
      const int valid_rows = left < TILE_M ? (int)left : TILE_M;
      const uint8_t* cbase = tile + ((4 * g) >> 6) * (TILE_M * 128);
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
    if (t == 0) tma_store_wait_read();        // shared memory must stay valid until TMA has read it
    }  // KS == 1
  }
  if constexpr (KS > 1) {
    constexpr int CW = BN / KS;                 // columns this CTA finishes
    constexpr int RP = 128;                     // receive buffer: [sender][column][row] fp32 (column-major: conflict-free)
    float* recv = reinterpret_cast<float*>(smem);
    float* colbuf = recv + (KS - 1) * CW * RP;  // [CW][129] rounded outputs for the BN statistics
    float* red = colbuf + CW * 129;             // [128 / CW][2 * CW]
    const int q = warp & 3, row = q * 32 + lane, t = threadIdx.x - 64;
    const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
    __syncwarp();
    // #0: every CTA's accumulator is complete, i.e. every CTA's pipeline stages are free to be overwritten
    cluster_barrier();
    if (warp >= 2) {
#pragma unroll 1
      for (int d = 0; d < KS; ++d) {
        if (d == ks_rank) continue;
        const int slot = ks_rank < d ? ks_rank : ks_rank - 1;
        const uint32_t remote = cluster_map_shared(smem_u32(recv + slot * CW * RP), (uint32_t)d);
#pragma unroll 1
        for (int c = 0; c < CW; c += 16) {
          float v[16];
          tmem_ld_32x32_x16(lane_base + d * CW + c, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) st_cluster_f32(remote + (uint32_t)(((c + j) * RP + row) * 4), v[j]);
        }
      }
    }
    __syncwarp();
    cluster_barrier();                          // #1: all partial column blocks have been delivered
    if (warp >= 2) {
      const long long grow = (long long)m_tile * TILE_M + row;
      const bool valid = grow < prm.total_pixels;
      bf16* orow = prm.out_raw + grow * prm.n_out + n_tile * BN + ks_rank * CW;
      const bool want_stats = !DGRAD && stats != nullptr;
#pragma unroll 1
      for (int c = 0; c < CW; c += 16) {
        float v[16];
        tmem_ld_32x32_x16(lane_base + ks_rank * CW + c, v);
        tmem_ld_wait();
#pragma unroll
        for (int sl = 0; sl < KS - 1; ++sl)
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] += recv[(sl * CW + c + j) * RP + row];
        if (!DGRAD) {
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] += bias_s[ks_rank * CW + c + j];
        }
        uint32_t pk[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) pk[j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
        if (valid) {
          reinterpret_cast<uint4*>(orow + c)[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
          reinterpret_cast<uint4*>(orow + c)[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
        }
        if (want_stats) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float2 f = unpack_bf16x2(pk[j]);      // statistics of the ROUNDED tile, like KS == 1
            colbuf[(c + 2 * j) * 129 + row] = valid ? f.x : 0.f;
            colbuf[(c + 2 * j + 1) * 129 + row] = valid ? f.y : 0.f;
          }
        }
      }
      if (want_stats) {
        constexpr int PARTS = 128 / CW;           // each part sums CW rows of one column
        asm volatile("bar.sync 1, 128;" ::: "memory");
        const int col = t % CW, part = t / CW;
        float s1 = 0.f, s2 = 0.f;
#pragma unroll 8
        for (int r = part * CW; r < (part + 1) * CW; ++r) {
          const float f = colbuf[col * 129 + r];
          s1 += f; s2 = fmaf(f, f, s2);
        }
        red[part * 2 * CW + col] = s1;
        red[part * 2 * CW + CW + col] = s2;
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (t < 2 * CW) {
          float acc = 0.f;
#pragma unroll
          for (int pz = 0; pz < PARTS; ++pz) acc += red[pz * 2 * CW + t];
          const int cc = t < CW ? t : t - CW;
          atomicAdd(&stats[(t < CW ? 0 : prm.n_out) + n_tile * BN + ks_rank * CW + cc], acc);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (NC > 1) cluster_barrier();      // peers may still arrive on our barriers: do not exit before them
  if (warp == 1) tmem_dealloc<BN>(tmem_base);
}

template <int BN, int STAGES>
constexpr int conv_tc_smem() { return STAGES * (A_BYTES + BN * KCHUNK * 2) + (2 * STAGES + 1) * 8 + 16 + BN * 4 + 1024; }

// pixel box (w, h, n) holding `pixels` (128 or 64) consecutive NHWC pixels of an HxW image
void pixel_box(int H, int W, int pixels, int* bw, int* bh, int* bn) {
  if (H * W >= pixels) { *bw = W; *bh = pixels / W; *bn = 1; }
  else { *bw = W; *bh = H; *bn = pixels / (H * W); }
}

template <int BN, int STAGES, bool DGRAD, int NC = 1, int KS = 1>
static int launch(const bf16* in, const bf16* w, bf16* out, const float* bias, float* stats, int B, int H, int W,
                  int Cin, int Cout, cudaStream_t stream, int taps = 9) {
  const int Kc = DGRAD ? Cout : Cin;     // reduction channels
  const int Nc = DGRAD ? Cin : Cout;     // output channels
  int bw, bh, bn;
  pixel_box(H, W, TILE_M, &bw, &bh, &bn);
  ConvTcParams prm;
  prm.slice_dim = 0; prm.slice_extent = 0;
  int sbh = bh, sbn = bn;                // box of one multicast slice
  if (NC > 1) {
    if (bn % NC == 0) { sbn = bn / NC; prm.slice_dim = 3; prm.slice_extent = sbn; }
    else if (bn == 1 && bh % NC == 0) { sbh = bh / NC; prm.slice_dim = 2; prm.slice_extent = sbh; }
    else DDPB_REQUIRE(false, "conv_tc: cannot cut a %dx%dx%d pixel box into %d multicast slices", bw, bh, bn, NC);
    DDPB_REQUIRE((Nc / BN) % NC == 0, "conv_tc: %d channel tiles are not a multiple of the cluster size %d", Nc / BN, NC);
  }
  CUtensorMap tmA, tmB, tmC;
  int rc = make_tmap_nhwc_bf16(&tmA, in, B, H, W, Kc, KCHUNK, bw, sbh, sbn);
  if (rc) return rc;
  // output matrix [pixels][Nc]; the epilogue stores 64-column x 128-row boxes (clipped at the edge)
  rc = make_tmap_2d_bf16(&tmC, out, (long long)B * H * W, Nc, 64, TILE_M);
  if (rc) return rc;
  // weight matrix [Cout rows][9*Cin cols]; fprop box = 64 cols x BN rows, dgrad box = 64 cols x 64 rows
  rc = make_tmap_2d_bf16(&tmB, w, Cout, (long long)taps * Cin, KCHUNK, DGRAD ? 64 : BN);
  if (rc) return rc;
  prm.H = H; prm.W = W;
  prm.tiles_per_img = H * W >= TILE_M ? H * W / TILE_M : 0;
  prm.imgs_per_tile = H * W >= TILE_M ? 0 : TILE_M / (H * W);
  prm.rows_per_tile = bh;
  prm.total_pixels = (long long)B * H * W;
  prm.n_out = Nc;
  prm.k_chunks = Kc / KCHUNK;
  prm.k_total_cols = Cin;
  prm.taps = taps;
  prm.out_raw = out;
  static_assert(KS == 1 || NC == 1, "split-K and multicast clusters are separate variants");
  static_assert(KS == 1 || ((KS - 1) * (BN / KS) * 128 + (BN / KS) * 129 + 256) * 4 <= STAGES * (A_BYTES + BN * KCHUNK * 2),
                "split-K receive buffers must fit in the pipeline stages");
  int m_tiles = prm.tiles_per_img > 0 ? B * prm.tiles_per_img : ceil_div(B, prm.imgs_per_tile);
  dim3 grid(m_tiles, Nc / BN, KS);
  constexpr int smem = conv_tc_smem<BN, STAGES>();
  auto kern = conv3x3_tc_kernel<BN, STAGES, DGRAD, NC, KS>;
  static bool configured = false;
  if (!configured) {
    DDPB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  DDPB_CUDA_CHECK(launch_ex(kern, grid, dim3(192), smem, stream, dim3(1, NC, KS), 1, tmA, tmB, tmC, bias, stats, prm));
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}

static long long cta_count(int B, int H, int W, int n_channels, int tile_n) {
  long long m_tiles = H * W >= TILE_M ? (long long)B * (H * W / TILE_M) : ceil_div(B, TILE_M / (H * W));
  return m_tiles * (n_channels / tile_n);
}

// CTA-pair (cta_group::2) variant, conv_tc2.cu
int launch_conv_tc2(bool dgrad, int bn, const void* in, const void* w, void* out, const float* bias, float* stats,
                    int B, int H, int W, int Cin, int Cout, cudaStream_t stream);

// tile_n codes: 4264 / 4464 / 4228 / 4428 = cluster split-K (2 or 4 CTAs per 64- or 128-channel tile);
// 64 / 128 = one CTA per tile; 1864 / 1464 / 1264 = 64-channel tiles in clusters of 8 / 4 / 2
// CTAs that multicast the activation tile, 1228 = 128-channel tiles in clusters of 2; 2128 / 2256 = CTA
// pair (cta_group::2) with 128 / 256 channels per tile; 0 = automatic.  DDPB_CONV_2CTA=1 in the environment lets the automatic choice use pair tiles.
static bool pair_tiles_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("DDPB_CONV_2CTA");
    v = (e && e[0] == '1') ? 1 : 0;   // measured on B200: pair tiles are within +-7% of single-CTA tiles for these shapes
  }
  return v == 1;
}

static int auto_tile(int B, int H, int W, int n_channels, int k_channels) {
  const long long pixels = (long long)B * H * W;
  if (pair_tiles_enabled() && pixels >= 2 * TILE_M) {
    // a pair tile is worth it when it still leaves >= ~1 CTA per SM
    if (n_channels % 256 == 0 && (pixels / TILE_M) * (n_channels / 256) >= 120) return 2256;
    if (n_channels % 128 == 0 && (pixels / TILE_M) * (n_channels / 128) >= 120) return 2128;
  }
  // One CTA ingests only ~30 B/clk, two per SM ~45 (the chip-wide L2 limit): aim for >= 256 CTAs, with
  // the widest tile (least L2 traffic per FLOP) that gets there, splitting K across a cluster when the
  // layer has too few output tiles (measured per layer in profiles/kernel_bench_r1.log).
  const long long t128 = n_channels % 128 == 0 ? cta_count(B, H, W, n_channels, 128) : 0;
  const long long t64 = cta_count(B, H, W, n_channels, 64);
  const int k_total = 9 * (k_channels / KCHUNK);
  const int max_ks = k_total >= 36 ? 4 : (k_total >= 18 ? 2 : 1);     // keep >= 9 k-iterations per CTA
  if (t128 >= 256) return 128;
  if (max_ks >= 2 && t128 * 2 >= 256) return 4228;
  if (max_ks >= 4 && t128 * 4 >= 256) return 4428;
  if (t64 >= 256 || max_ks == 1) return t128 >= 148 ? 128 : 64;
  if (t64 * 2 >= 256 || max_ks == 2) return 4264;
  return 4464;
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
  if (tile_n == 0) tile_n = auto_tile(B, H, W, Cout, Cin);
  if (tile_n > 2000 && tile_n < 4000) {
    DDPB_REQUIRE(Cout % (tile_n - 2000) == 0, "conv_tc fwd: Cout %d not a multiple of pair tile %d", Cout, tile_n - 2000);
    return launch_conv_tc2(false, tile_n - 2000, x, w_bf16, y, bias, stats, B, H, W, Cin, Cout, stream);
  }
  if (tile_n > 4000) {          // 4<splits><tile>: cluster split-K
    const bf16 *xi = (const bf16*)x, *wi = (const bf16*)w_bf16;
    bf16* yo = (bf16*)y;
    DDPB_REQUIRE(Cout % (tile_n % 100 == 28 ? 128 : 64) == 0, "conv_tc fwd: Cout %d does not fit tile code %d", Cout, tile_n);
    switch (tile_n) {
      case 4264: return cta_count(B, H, W, Cout, 64) * 2 <= 148
                            ? launch<64, 8, false, 1, 2>(xi, wi, yo, bias, stats, B, H, W, Cin, Cout, stream)
                            : launch<64, 4, false, 1, 2>(xi, wi, yo, bias, stats, B, H, W, Cin, Cout, stream);
      case 4464: return launch<64, 4, false, 1, 4>(xi, wi, yo, bias, stats, B, H, W, Cin, Cout, stream);
      case 4228: return launch<128, 3, false, 1, 2>(xi, wi, yo, bias, stats, B, H, W, Cin, Cout, stream);
      case 4428: return launch<128, 3, false, 1, 4>(xi, wi, yo, bias, stats, B, H, W, Cin, Cout, stream);
      default: DDPB_REQUIRE(false, "conv_tc fwd: unknown split-K tile code %d", tile_n);
    }
  }
  if (tile_n > 1000) {          // 1<cluster><tile>: activation tile multicast across a cluster of n-tiles
    const bf16 *xi = (const bf16*)x, *wi = (const bf16*)w_bf16;
    bf16* yo = (bf16*)y;
    const bool deep = cta_count(B, H, W, Cout, 64) <= 148;
    switch (tile_n) {
      case 1864: return deep ? launch<64, 8, false, 8>(xi, wi, yo, bias, stats, B, H, W, Cin, Cout, stream)
                             : launch<64, 4, false, 8>(xi, wi, yo, bias, stats, B, H, W, Cin, Cout, stream);
      case 1464: return deep ? launch<64, 8, false, 4>(xi, wi, yo, bias, stats, B, H, W, Cin, Cout, stream)
                             : launch<64, 4, false, 4>(xi, wi, yo, bias, stats, B, H, W, Cin, Cout, stream);
      case 1264: return launch<64, 4, false, 2>(xi, wi, yo, bias, stats, B, H, W, Cin, Cout, stream);
      case 1228: return launch<128, 3, false, 2>(xi, wi, yo, bias, stats, B, H, W, Cin, Cout, stream);
      case 1428: return launch<128, 3, false, 4>(xi, wi, yo, bias, stats, B, H, W, Cin, Cout, stream);
      default: DDPB_REQUIRE(false, "conv_tc fwd: unknown multicast tile code %d", tile_n);
    }
  }
  DDPB_REQUIRE(Cout % tile_n == 0, "conv_tc fwd: Cout %d not a multiple of tile %d", Cout, tile_n);
  if (tile_n == 128)
    return launch<128, 3, false>((const bf16*)x, (const bf16*)w_bf16, (bf16*)y, bias, stats, B, H, W, Cin, Cout, stream);
  if (tile_n == 64) {
    // fewer CTAs than SMs: one CTA per SM with a deep pipeline (more bytes in flight per CTA)
    if (cta_count(B, H, W, Cout, 64) <= 148)
      return launch<64, 8, false>((const bf16*)x, (const bf16*)w_bf16, (bf16*)y, bias, stats, B, H, W, Cin, Cout, stream);
    return launch<64, 4, false>((const bf16*)x, (const bf16*)w_bf16, (bf16*)y, bias, stats, B, H, W, Cin, Cout, stream);
  }
  DDPB_REQUIRE(false, "conv_tc fwd: tile_n must be 64 or 128");
}

// dx[B,H,W,Cin] = conv3x3_transpose(dy[B,H,W,Cout], w[Cout,3,3,Cin]); bf16 in/out.
extern "C" int ddpb_conv3x3_dgrad_tc(const void* dy, const void* w_bf16, void* dx, int B, int H, int W, int Cin,
                                     int Cout, int tile_n, cudaStream_t stream) {
  int rc = check_shapes(B, H, W, Cin, Cout);
  if (rc) return rc;
  if (tile_n == 0) tile_n = auto_tile(B, H, W, Cin, Cout);
  if (tile_n > 2000 && tile_n < 4000) {
    DDPB_REQUIRE(Cin % (tile_n - 2000) == 0, "conv_tc dgrad: Cin %d not a multiple of pair tile %d", Cin, tile_n - 2000);
    return launch_conv_tc2(true, tile_n - 2000, dy, w_bf16, dx, nullptr, nullptr, B, H, W, Cin, Cout, stream);
  }
  if (tile_n > 4000) {
    const bf16 *di = (const bf16*)dy, *wi = (const bf16*)w_bf16;
    bf16* xo = (bf16*)dx;
    DDPB_REQUIRE(Cin % (tile_n % 100 == 28 ? 128 : 64) == 0, "conv_tc dgrad: Cin %d does not fit tile code %d", Cin, tile_n);
    switch (tile_n) {
      case 4264: return cta_count(B, H, W, Cin, 64) * 2 <= 148
                            ? launch<64, 8, true, 1, 2>(di, wi, xo, nullptr, nullptr, B, H, W, Cin, Cout, stream)
                            : launch<64, 4, true, 1, 2>(di, wi, xo, nullptr, nullptr, B, H, W, Cin, Cout, stream);
      case 4464: return launch<64, 4, true, 1, 4>(di, wi, xo, nullptr, nullptr, B, H, W, Cin, Cout, stream);
      case 4228: return launch<128, 3, true, 1, 2>(di, wi, xo, nullptr, nullptr, B, H, W, Cin, Cout, stream);
      case 4428: return launch<128, 3, true, 1, 4>(di, wi, xo, nullptr, nullptr, B, H, W, Cin, Cout, stream);
      default: DDPB_REQUIRE(false, "conv_tc dgrad: unknown split-K tile code %d", tile_n);
    }
  }
  if (tile_n > 1000) {
    const bf16 *di = (const bf16*)dy, *wi = (const bf16*)w_bf16;
    bf16* xo = (bf16*)dx;
    const bool deep = cta_count(B, H, W, Cin, 64) <= 148;
    switch (tile_n) {
      case 1864: return deep ? launch<64, 8, true, 8>(di, wi, xo, nullptr, nullptr, B, H, W, Cin, Cout, stream)
                             : launch<64, 4, true, 8>(di, wi, xo, nullptr, nullptr, B, H, W, Cin, Cout, stream);
      case 1464: return deep ? launch<64, 8, true, 4>(di, wi, xo, nullptr, nullptr, B, H, W, Cin, Cout, stream)
                             : launch<64, 4, true, 4>(di, wi, xo, nullptr, nullptr, B, H, W, Cin, Cout, stream);
      case 1264: return launch<64, 4, true, 2>(di, wi, xo, nullptr, nullptr, B, H, W, Cin, Cout, stream);
      case 1228: return launch<128, 3, true, 2>(di, wi, xo, nullptr, nullptr, B, H, W, Cin, Cout, stream);
      case 1428: return launch<128, 3, true, 4>(di, wi, xo, nullptr, nullptr, B, H, W, Cin, Cout, stream);
      default: DDPB_REQUIRE(false, "conv_tc dgrad: unknown multicast tile code %d", tile_n);
    }
  }
  DDPB_REQUIRE(Cin % tile_n == 0, "conv_tc dgrad: Cin %d not a multiple of tile %d", Cin, tile_n);
  if (tile_n == 128)
    return launch<128, 3, true>((const bf16*)dy, (const bf16*)w_bf16, (bf16*)dx, nullptr, nullptr, B, H, W, Cin, Cout, stream);
  if (tile_n == 64) {
    if (cta_count(B, H, W, Cin, 64) <= 148)
      return launch<64, 8, true>((const bf16*)dy, (const bf16*)w_bf16, (bf16*)dx, nullptr, nullptr, B, H, W, Cin, Cout, stream);
    return launch<64, 4, true>((const bf16*)dy, (const bf16*)w_bf16, (bf16*)dx, nullptr, nullptr, B, H, W, Cin, Cout, stream);
  }
  DDPB_REQUIRE(false, "conv_tc dgrad: tile_n must be 64 or 128");
}

// y[B,H,W,Cout] = x[B,H,W,Cin] * w[Cout,Cin]^T + bias: the same kernel with a single (centre) tap.
// Used for the first VGG layer, whose 3x3x3 patches are laid out as a [pixels][64] matrix first
// (ddpb_im2col3x3_c3) so that it runs on the tensor cores like every other layer.
extern "C" int ddpb_conv1x1_fwd_tc(const void* x, const void* w_bf16, const float* bias, void* y, float* stats,
                                   int B, int H, int W, int Cin, int Cout, cudaStream_t stream) {
  int rc = check_shapes(B, H, W, Cin, Cout);
  if (rc) return rc;
  if (Cout % 128 == 0 && cta_count(B, H, W, Cout, 128) >= 148)
    return launch<128, 3, false>((const bf16*)x, (const bf16*)w_bf16, (bf16*)y, bias, stats, B, H, W, Cin, Cout, stream, 1);
  return launch<64, 4, false>((const bf16*)x, (const bf16*)w_bf16, (bf16*)y, bias, stats, B, H, W, Cin, Cout, stream, 1);
}
