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
  int pdl;                 // 1: execute griddepcontrol.launch_dependents (captured launches only, see common.cuh)
  int taps;                // 9 = 3x3 filter, 1 = 1x1 (plain GEMM over the pixel rows; first-layer patch matrix)
  unsigned long long* trace;   // nullable: [ctas][8] %globaltimer stamps of this CTA's phases (ddpb_conv3x3_trace)
  bf16* out_raw;           // output matrix [total_pixels][n_out]
  int w_shift, h_shift;    // log2(W), log2(H)
  // dgrad only: BatchNorm-backward reduction of the previous block, fused into the epilogue (bn_y == null: off).
  // bn_y: per output pixel and channel, the previous block's conv output that won its ReLU / pool window
  // (NaN = blocked), [total_pixels][n_out].
  const bf16* bn_y;
  const float *bn_mean, *bn_invstd, *bn_gamma, *bn_beta;
  float* bn_sums;          // [2 * n_out], accumulated
  // fprop only: this block's BatchNorm + ReLU (+ 2x2 max-pool) applied in the epilogue after a grid-wide
  // barrier (bnf_out == null: off).  Needs every CTA of the grid resident at once (checked by the launcher).
  bf16* bnf_out;           // block output [B, H/(1+pool), W/(1+pool), n_out]
  bf16* bnf_winner;        // nullable: conv output that won each window (NaN: blocked by ReLU)
  const float *bnf_gamma, *bnf_beta;
  float *bnf_save_mean, *bnf_save_invstd, *bnf_running_mean, *bnf_running_var;
  long long* bnf_nbt;      // nullable
  unsigned int* bnf_barrier;   // zeroed by the caller before every launch
  unsigned int bnf_total_ctas;
  int bnf_pool;
  float bnf_eps, bnf_momentum;
};

// XOR swizzle of the 16-byte chunks of the epilogue's bf16 tile ([128 rows][CH chunks]): both the
// row-per-thread writes and the chunk-per-thread reads then spread over all shared-memory banks.
template <int CH>
__device__ __forceinline__ int tile_swizzle(int row) {
  if (CH >= 8) return row & 7;
  if (CH == 4) return (row >> 1) & 3;
  if (CH == 2) return (row >> 2) & 1;
  return 0;
}

// NC > 1: the NC CTAs of a cluster (consecutive n-tiles, same pixel tile) need the SAME activation
// tile.  Each loads 1/NC of it and multicasts that slice into all NC shared memories, so the tile is
// read from L2 once per cluster instead of once per CTA (the kernels are bound by L2 -> SM traffic).
//
// KS > 1 (cluster split-K): the deep layers have few output tiles (64 for the 2x2 layers) but a long
// reduction (K = 4608), and one CTA cannot pull more than ~30-50 B/clk through TMA.  KS CTAs of a
// cluster (blockIdx.z) each accumulate 1/KS of the K range of the SAME tile; then every CTA sends the
// other CTAs' column blocks of its fp32 partial tile through distributed shared memory and finishes
// BN/KS columns itself (reduce-scatter), so the epilogue work is split KS ways too.
//
// PAIR (cta_group::2): two CTAs with consecutive pixel tiles (cluster ranks 2k, 2k+1, the two SMs of a TPC)
// form ONE tcgen05.mma of M = 256: each loads its own 128-pixel A tile and only HALF of the weight tile,
// the leader issues the MMAs, which read both CTAs' shared memory and write each CTA's 128 x BN accumulator
// into that CTA's TMEM.  Per byte brought in from L2 a CTA performs 2x (BN = 256) the MMA work of the
// single-CTA 128 x 128 tile - and these kernels are bound by L2 -> SM traffic, not by the tensor pipe.
// PAIR combines with KS (cluster = 2 x KS CTAs) so that layers with few pixel tiles still fill the GPU.
template <int BN, int STAGES, bool DGRAD, int NC, int KS, bool PAIR>
__global__ void __launch_bounds__(192)
conv3x3_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const float* __restrict__ bias, float* __restrict__ stats, const ConvTcParams prm) {
  static_assert(!PAIR || NC == 1, "pair tiles do not combine with the activation-multicast clusters");
  constexpr int BH = PAIR ? BN / 2 : BN;             // weight-tile channels held by THIS CTA
  constexpr int B_BYTES = BH * KCHUNK * 2;
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
  unsigned long long* const trc =
      prm.trace ? prm.trace + 8ull * (blockIdx.x + gridDim.x * (blockIdx.y + (unsigned long long)gridDim.y * blockIdx.z)) : nullptr;
  if (trc && threadIdx.x == 0) trc[0] = global_timer_ns();            // CTA started

  // output tile -> (image, row) origin of the pixel box
  int n0, h0;
  if (prm.tiles_per_img > 0) { n0 = m_tile / prm.tiles_per_img; h0 = (m_tile % prm.tiles_per_img) * prm.rows_per_tile; }
  else { n0 = m_tile * prm.imgs_per_tile; h0 = 0; }
  const int k_total = prm.taps * prm.k_chunks;
  const int ks_rank = KS > 1 ? (int)blockIdx.z : 0;
  const int pr = PAIR ? (int)(cluster_rank() & 1u) : 0;        // rank inside the pair; 0 = leader
  // cluster ranks of this CTA's pair (bits of the commit multicast mask) and of its split-K partners
  const uint16_t pair_mask = PAIR ? (uint16_t)(3u << (2 * ks_rank)) : (uint16_t)0;
  const int it_lo = KS > 1 ? k_total * ks_rank / KS : 0;
  const int k_iters = (KS > 1 ? k_total * (ks_rank + 1) / KS : k_total) - it_lo;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    // a stage is written by all NC CTAs (multicast slices), so it is free only when all NC have consumed it
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], NC); }   // (pair: the leader's full barrier counts both CTAs' bytes)
    mbar_init(accum_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) { if (PAIR) tmem2_alloc<BN>(tmem_slot); else tmem_alloc<BN>(tmem_slot); }
  // everything above is private to this CTA and overlaps the previous kernel's tail (programmatic
  // dependent launch); global memory may only be touched after pdl_wait()
  if (prm.pdl & 1) pdl_launch_dependents();
  pdl_wait();
  ktrace_begin(prm.pdl);      // stamped once the predecessor has finished: the start of this kernel's own work
  if (trc && threadIdx.x == 0) trc[1] = global_timer_ns();            // prologue done, predecessor finished
  if (!DGRAD) for (int i = threadIdx.x; i < BN; i += blockDim.x) bias_s[i] = bias ? ld_coh(&bias[n_tile * BN + i]) : 0.f;
  tc_fence_before();
  __syncthreads();
  if (NC > 1 || PAIR) cluster_barrier();      // every CTA's barriers exist before a peer's loads / commits can touch them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  constexpr uint16_t MC_MASK = (uint16_t)((1u << NC) - 1);
  // dgrad + fused BN reduction only (epilogue warps): this thread's share of the winner tensor, fetched ahead
  // of the accumulator when it fits in registers (up to 16 x 16 bytes), else loaded where it is used
  constexpr bool WIN_PREFETCH = BN / KS / 8 <= 16;
  uint4 ywin[WIN_PREFETCH ? BN / KS / 8 : 1];
  float bn_mu[8], bn_is[8], bn_sc[8], bn_sh[8];   // its per-channel BatchNorm constants, fetched up front as well

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
        if (!PAIR) mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
        else if (pr == 0) mbar_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);
        // fprop reads X at p + (r-1, s-1); dgrad reads dY at p - (r-1, s-1)
        const int dh = DGRAD ? 1 - r : r - 1, dw = DGRAD ? 1 - s : s - 1;
        if (PAIR) {
          // both CTAs load into their own shared memory; the bytes are counted on the LEADER's barrier, which
          // is what the one MMA-issuing thread waits for
          const uint32_t lbar = pair_leader_bar(&full_bar[stage]);
          tma2_load_4d(sa, &tmA, lbar, kc * KCHUNK, dw, h0 + dh, n0);
          if (!DGRAD) {
            tma2_load_2d(sb, &tmB, lbar, tap * prm.k_total_cols + kc * KCHUNK, n_tile * BN + pr * BH);
          } else {
#pragma unroll
            for (int j = 0; j < BH / 64; ++j)
              tma2_load_2d(sb + j * (64 * 128), &tmB, lbar, tap * prm.k_total_cols + n_tile * BN + pr * BH + j * 64,
                           kc * KCHUNK);
          }
          continue;
        }
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
    if (lane == 0 && pr == 0) {         // (pair: the leader's thread drives both SMs)
      constexpr uint32_t idesc = umma_idesc_bf16(PAIR ? 2 * TILE_M : TILE_M, BN, false, DGRAD);
      for (int it = 0; it < k_iters; ++it) {
        const int stage = it % STAGES, use = it / STAGES;
        mbar_wait(&full_bar[stage], use & 1);
        if (trc && it == 0) trc[2] = global_timer_ns();               // first stage landed
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
          if (PAIR) umma2_bf16(tmem_base, da, db, idesc, (it | kk) != 0);
          else umma_bf16(tmem_base, da, db, idesc, (it | kk) != 0);
        }
        // frees the stage when these MMAs have read it (in every CTA that multicasts into it / in both CTAs of a pair)
        if (PAIR) umma2_commit_mc(&empty_bar[stage], pair_mask);
        else if (NC == 1) umma_commit(&empty_bar[stage]);
        else umma_commit_mc(&empty_bar[stage], MC_MASK);
      }
      if (PAIR) umma2_commit_mc(accum_bar, pair_mask);       // accumulator complete, in both CTAs
      else umma_commit(accum_bar);
      if (trc) trc[3] = global_timer_ns();                             // last MMA issued
    }
  } else {
    // ===================== epilogue warps (2..5) =====================
    // dgrad with fused BN-backward reduction: fetch this thread's share of the previous block's
    // "winner" tensor now, so the loads are long complete when the accumulator is
    if (DGRAD && prm.bn_y != nullptr) {
      const int t2 = threadIdx.x - 64;
      const int col = n_tile * BN + ks_rank * (BN / KS) + (t2 % (BN / KS / 8)) * 8;
      ld_coh_vec<8>(prm.bn_mean + col, bn_mu);
      ld_coh_vec<8>(prm.bn_invstd + col, bn_is);
      ld_coh_vec<8>(prm.bn_gamma + col, bn_sc);
      ld_coh_vec<8>(prm.bn_beta + col, bn_sh);
#pragma unroll
      for (int k = 0; k < (WIN_PREFETCH ? BN / KS / 8 : 0); ++k) {
        const long long grow = (long long)m_tile * TILE_M + t2 / (BN / KS / 8) + k * (128 / (BN / KS / 8));
        ywin[k] = grow < prm.total_pixels
                      ? DDPB_LD_COH(reinterpret_cast<const uint4*>(prm.bn_y + grow * prm.n_out + col))
                      : make_uint4(0x7fc07fc0u, 0x7fc07fc0u, 0x7fc07fc0u, 0x7fc07fc0u);
      }
    }
    mbar_wait(accum_bar, 0);
    tc_fence_after();
    if (trc && threadIdx.x == 64) trc[4] = global_timer_ns();         // accumulator complete
  }
  // late trigger (see kernel_flags): every thread is past its main loop - the successor's CTAs become resident
  // while this CTA runs its epilogue, not while it still needs the SM to itself and its neighbours
  if (prm.pdl & 2) pdl_launch_dependents();
  // Epilogue, common to all variants.  This CTA finishes CW = BN / KS columns of the 128-row tile:
  //   0. (KS > 1) reduce-scatter of the fp32 partial tiles through distributed shared memory
  //   1. TMEM (+ partials) (+ bias) -> bf16 -> shared-memory tile (16-byte chunks, XOR-swizzled)
  //   2. re-mapped threads (CH consecutive threads = one row) store the tile with coalesced 16-byte
  //      stores and, from the same ROUNDED values, accumulate per-channel sums in registers:
  //        fprop: BatchNorm statistics (sum, sum of squares) of this layer
  //        dgrad: the BatchNorm-backward reduction of the PREVIOUS layer (sum dz, sum dz * xhat),
  //               which would otherwise be a kernel of its own on the critical path
  //   3. cross-thread reduction through shared memory, one atomicAdd per channel and CTA
  constexpr int CW = BN / KS;                 // columns this CTA finishes
  constexpr int CH = CW / 8;                  // 16-byte chunks per tile row
  constexpr int PARTS = 128 / CH;             // row groups in step 2
  constexpr int RP = 128;                     // receive buffer: [sender][column][row] fp32 (column-major: conflict-free)
  float* recv = reinterpret_cast<float*>(smem);
  uint8_t* tile = smem + (KS - 1) * CW * RP * 4;                        // [128 rows][CW] bf16
  // [PARTS][2 * CW]; with split-K it reuses the receive area, which is dead once step 1 below has run
  float* red = KS > 1 ? recv : reinterpret_cast<float*>(tile + TILE_M * CW * 2);
  const int q = warp & 3, row = q * 32 + lane, t = threadIdx.x - 64;
  const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
  if constexpr (KS > 1) {
    __syncwarp();
    // #0: every CTA's accumulator is complete, i.e. every CTA's pipeline stages are free to be overwritten
    cluster_barrier();
    if (warp >= 2) {
#pragma unroll 1
      for (int d = 0; d < KS; ++d) {
        if (d == ks_rank) continue;
        const int slot = ks_rank < d ? ks_rank : ks_rank - 1;
        const uint32_t remote = cluster_map_shared(smem_u32(recv + slot * CW * RP), (uint32_t)(PAIR ? 2 * d + pr : d));
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
    if (trc && threadIdx.x == 64) trc[5] = global_timer_ns();         // split-K partials exchanged
  }
  if (warp >= 2) {
    // ---- 1. accumulator -> bf16 tile in shared memory
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
      uint8_t* rbase = tile + row * (CW * 2);
      const int j0 = c >> 3, sw = tile_swizzle<CH>(row);
      *reinterpret_cast<uint4*>(rbase + ((j0 ^ sw) << 4)) =
          make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
      *reinterpret_cast<uint4*>(rbase + (((j0 + 1) ^ sw) << 4)) =
          make_uint4(pack_bf16x2(v[8], v[9]), pack_bf16x2(v[10], v[11]), pack_bf16x2(v[12], v[13]), pack_bf16x2(v[14], v[15]));
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");             // the four epilogue warps only
    // ---- 2. coalesced store + per-channel sums
    const int ch = t % CH, part = t / CH;
    const int col0 = n_tile * BN + ks_rank * CW + ch * 8;      // first of this thread's 8 output channels
    const long long row0 = (long long)m_tile * TILE_M;
    const bool want_stats = !DGRAD && stats != nullptr;
    const bool bn_red = DGRAD && prm.bn_y != nullptr;
    float s1[8], s2[8], sc[8], sh[8], mu[8], is[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { s1[j] = 0.f; s2[j] = 0.f; sc[j] = 0.f; sh[j] = 0.f; mu[j] = 0.f; is[j] = 0.f; }
    if (bn_red) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        mu[j] = bn_mu[j]; is[j] = bn_is[j];
        sc[j] = bn_sc[j] * is[j];
        sh[j] = bn_sh[j] - mu[j] * sc[j];
      }
    }
#pragma unroll
    for (int k = 0; k < CH; ++k) {
      const int r = part + k * PARTS;
      const long long grow = row0 + r;
      if (grow >= prm.total_pixels) break;
      const uint4 u = *reinterpret_cast<const uint4*>(tile + r * (CW * 2) + ((ch ^ tile_swizzle<CH>(r)) << 4));
      *reinterpret_cast<uint4*>(prm.out_raw + grow * prm.n_out + col0) = u;
      float d[8];
      { const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c2 = unpack_bf16x2(u.z), e = unpack_bf16x2(u.w);
        d[0] = a.x; d[1] = a.y; d[2] = b.x; d[3] = b.y; d[4] = c2.x; d[5] = c2.y; d[6] = e.x; d[7] = e.y; }
      if (want_stats) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { s1[j] += d[j]; s2[j] = fmaf(d[j], d[j], s2[j]); }
      }
      if (bn_red) {
        // yw: the previous block's conv output behind this pixel (NaN: ReLU blocked it -> z > 0 is false)
        float v[8];
        { const uint4 w4 = WIN_PREFETCH ? ywin[WIN_PREFETCH ? k : 0]
                                        : DDPB_LD_COH(reinterpret_cast<const uint4*>(prm.bn_y + grow * prm.n_out + col0));
          const float2 a = unpack_bf16x2(w4.x), b = unpack_bf16x2(w4.y), c2 = unpack_bf16x2(w4.z), e = unpack_bf16x2(w4.w);
          v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c2.x; v[5] = c2.y; v[6] = e.x; v[7] = e.y; }
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (fmaf(v[j], sc[j], sh[j]) > 0.f) { s1[j] += d[j]; s2[j] = fmaf(d[j], (v[j] - mu[j]) * is[j], s2[j]); }
      }
    }
    if (trc && threadIdx.x == 64) trc[6] = global_timer_ns();         // tile stored
    // ---- 3. per-channel totals of the CTA
    if (want_stats || bn_red) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        red[part * 2 * CW + ch * 8 + j] = s1[j];
        red[part * 2 * CW + CW + ch * 8 + j] = s2[j];
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      float* acc_out = want_stats ? stats : prm.bn_sums;
      for (int i = t; i < 2 * CW; i += 128) {
        float acc = 0.f;
#pragma unroll 4
        for (int pz = 0; pz < PARTS; ++pz) acc += red[pz * 2 * CW + i];
        const int cc = i < CW ? i : i - CW;
        atomicAdd(&acc_out[(i < CW ? 0 : prm.n_out) + n_tile * BN + ks_rank * CW + cc], acc);
      }
    }
    // ---- 4. (fprop) fused BatchNorm + ReLU + max-pool of this block
    if (!DGRAD && prm.bnf_out != nullptr) {
      // grid-wide barrier: the statistics are complete when every CTA has added its share
      __threadfence();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (t == 0) {
        atomicAdd(prm.bnf_barrier, 1u);
        const unsigned long long t0 = global_timer_ns();
        while (ld_acquire_gpu(prm.bnf_barrier) < prm.bnf_total_ctas) {
          if (global_timer_ns() - t0 > 2000000000ull) {
            printf("ddpb: conv+BN grid barrier timed out (block %d,%d,%d)\n", blockIdx.x, blockIdx.y, blockIdx.z);
            __trap();
          }
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      const float inv_p = 1.f / (float)prm.total_pixels;
      float bsc[8], bsh[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float mean = __ldcg(&stats[col0 + j]) * inv_p;
        const float var = fmaxf(__ldcg(&stats[prm.n_out + col0 + j]) * inv_p - mean * mean, 0.f);
        const float invstd = rsqrtf(var + prm.bnf_eps);
        bsc[j] = __ldcg(&prm.bnf_gamma[col0 + j]) * invstd;
        bsh[j] = __ldcg(&prm.bnf_beta[col0 + j]) - mean * bsc[j];
        if (m_tile == 0 && part == 0) {      // one thread per channel in the whole grid
          prm.bnf_save_mean[col0 + j] = mean;
          prm.bnf_save_invstd[col0 + j] = invstd;
          const float n = (float)prm.total_pixels;
          const float unbiased = prm.total_pixels > 1 ? var * (n / (n - 1.f)) : var;
          prm.bnf_running_mean[col0 + j] = (1.f - prm.bnf_momentum) * __ldcg(&prm.bnf_running_mean[col0 + j]) + prm.bnf_momentum * mean;
          prm.bnf_running_var[col0 + j] = (1.f - prm.bnf_momentum) * __ldcg(&prm.bnf_running_var[col0 + j]) + prm.bnf_momentum * unbiased;
        }
      }
      if (m_tile == 0 && n_tile == 0 && ks_rank == 0 && t == 0 && prm.bnf_nbt) *prm.bnf_nbt += 1;
      // the tile is a box of `rows_per_tile` image rows (H*W >= 128) or of whole images (H*W < 128)
      const int W = prm.W, H = prm.H;
      const int pool = prm.bnf_pool;
      const int ow_n = pool ? W >> 1 : W;                                   // outputs per image row
      const int img_rows = prm.tiles_per_img > 0 ? prm.rows_per_tile : H;   // image rows per sub-box
      const int orow_n = pool ? img_rows >> 1 : img_rows;                   // output rows per sub-box
      const int sub_out = ow_n * orow_n;                                    // outputs per sub-box
      const int n_sub = prm.tiles_per_img > 0 ? 1 : prm.imgs_per_tile;
      const int out_per_img = pool ? (H >> 1) * (W >> 1) : H * W;
      const long long total_imgs = prm.total_pixels / ((long long)H * W);
      for (int o = part; o < sub_out * n_sub; o += PARTS) {
        const int sub = o / sub_out, oi = o - sub * sub_out;
        const int orow = oi / ow_n, ocol = oi - orow * ow_n;
        const long long img = n0 + sub;
        if (img >= total_imgs) break;
        const int r_base = sub * (H * W) + (pool ? (2 * orow) * W + 2 * ocol : orow * W + ocol);
        float best[8], by[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) { best[j] = 0.f; by[j] = __int_as_float(0x7fc00000); }
        const int nwin = pool ? 4 : 1;
        for (int wq = 0; wq < nwin; ++wq) {
          const int r = r_base + (wq & 1) + (wq >> 1) * W;
          const uint4 u = *reinterpret_cast<const uint4*>(tile + r * (CW * 2) + ((ch ^ tile_swizzle<CH>(r)) << 4));
          float v[8];
          { const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c2 = unpack_bf16x2(u.z), e = unpack_bf16x2(u.w);
            v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c2.x; v[5] = c2.y; v[6] = e.x; v[7] = e.y; }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float z = fmaf(v[j], bsc[j], bsh[j]);
            if (z > best[j]) { best[j] = z; by[j] = v[j]; }       // strict >: first maximum wins (matches the backward pass)
          }
        }
        const long long opix = img * out_per_img + (long long)((prm.tiles_per_img > 0 ? (pool ? h0 >> 1 : h0) : 0) + orow) * ow_n + ocol;
        const uint4 ov = make_uint4(pack_bf16x2(best[0], best[1]), pack_bf16x2(best[2], best[3]),
                                    pack_bf16x2(best[4], best[5]), pack_bf16x2(best[6], best[7]));
        *reinterpret_cast<uint4*>(prm.bnf_out + opix * prm.n_out + col0) = ov;
        if (prm.bnf_winner) {
          const uint4 wv = make_uint4(pack_bf16x2(by[0], by[1]), pack_bf16x2(by[2], by[3]),
                                      pack_bf16x2(by[4], by[5]), pack_bf16x2(by[6], by[7]));
          *reinterpret_cast<uint4*>(prm.bnf_winner + opix * prm.n_out + col0) = wv;
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (NC > 1 || PAIR) cluster_barrier();      // peers may still arrive on our barriers / read our shared memory through the pair MMA
  if (warp == 1) { if (PAIR) tmem2_dealloc<BN>(tmem_base); else tmem_dealloc<BN>(tmem_base); }
  ktrace_end(prm.pdl);
  if (trc && threadIdx.x == 0) trc[7] = global_timer_ns();            // CTA done
}

template <int BN, int STAGES, bool PAIR = false>
constexpr int conv_tc_smem() {
  return STAGES * (A_BYTES + (PAIR ? BN / 2 : BN) * KCHUNK * 2) + (2 * STAGES + 1) * 8 + 16 + BN * 4 + 1024;
}

// pixel box (w, h, n) holding `pixels` (128 or 64) consecutive NHWC pixels of an HxW image
void pixel_box(int H, int W, int pixels, int* bw, int* bh, int* bn) {
  if (H * W >= pixels) { *bw = W; *bh = pixels / W; *bn = 1; }
  else { *bw = W; *bh = H; *bn = pixels / (H * W); }
}

// Arguments of the fused BatchNorm-backward reduction (ddpb_conv3x3_dgrad_bnred_tc).
struct BnRedArgs {
  const void* y = nullptr;
  const float *mean = nullptr, *invstd = nullptr, *gamma = nullptr, *beta = nullptr;
  float* sums = nullptr;
};

// Arguments of the fused BatchNorm forward (ddpb_conv3x3_fwd_bn_tc).
struct BnFwdArgs {
  void *out = nullptr, *winner = nullptr;
  const float *gamma = nullptr, *beta = nullptr;
  float *save_mean = nullptr, *save_invstd = nullptr, *running_mean = nullptr, *running_var = nullptr;
  long long* nbt = nullptr;
  unsigned int* barrier = nullptr;
  int pool = 0;
  float eps = 1e-5f, momentum = 0.1f;
  bool applied = false;      // out: did the launch run with the fusion (all CTAs co-resident)?
};

// Everything beyond the plain convolution that an entry point may ask of a launch; handed down the tile
// dispatch explicitly (it used to travel in thread-local variables around the public API).
struct ConvExtra {
  const BnRedArgs* bnred = nullptr;       // dgrad: fuse the previous block's BatchNorm-backward reduction
  BnFwdArgs* bnfwd = nullptr;             // fprop: fuse this block's BatchNorm forward behind a grid barrier
  unsigned long long* trace = nullptr;    // per-CTA phase time stamps
};
static const ConvExtra kNoExtra;

template <int BN, int STAGES, bool DGRAD, int NC = 1, int KS = 1, bool PAIR = false>
static int launch(const bf16* in, const bf16* w, bf16* out, const float* bias, float* stats, int B, int H, int W,
                  int Cin, int Cout, cudaStream_t stream, const ConvExtra& ex = kNoExtra, int taps = 9) {
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
    // TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another
    // GPU's memory hung the GPU: the multicast variants take nothing but a plain device allocation of the
    // current device as their activation tensor (no peer / IPC / imported mapping).
    int cur = -1;
    cudaPointerAttributes pa;
    DDPB_CUDA_CHECK(cudaGetDevice(&cur));
    DDPB_CUDA_CHECK(cudaPointerGetAttributes(&pa, in));
    DDPB_REQUIRE(pa.type == cudaMemoryTypeDevice && pa.device == cur,
                 "conv_tc: the multicast tile variants need an activation tensor in the current GPU's own memory "
                 "(pointer type %d on device %d, current device %d)", (int)pa.type, pa.device, cur);
  }
  CUtensorMap tmA, tmB;
  int rc = make_tmap_nhwc_bf16(&tmA, in, B, H, W, Kc, KCHUNK, bw, sbh, sbn);
  if (rc) return rc;
  // weight matrix [Cout rows][9*Cin cols]; fprop box = 64 cols x BN rows, dgrad box = 64 cols x 64 rows
  rc = make_tmap_2d_bf16(&tmB, w, Cout, (long long)taps * Cin, KCHUNK, DGRAD ? 64 : (PAIR ? BN / 2 : BN));
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
  prm.w_shift = 0; prm.h_shift = 0;
  while ((1 << prm.w_shift) < W) ++prm.w_shift;
  while ((1 << prm.h_shift) < H) ++prm.h_shift;
  // fused BatchNorm-backward reduction requested by ddpb_conv3x3_dgrad_bnred_tc (dgrad only)
  static const BnRedArgs no_bnred;
  const BnRedArgs& br = ex.bnred ? *ex.bnred : no_bnred;
  prm.trace = ex.trace;
  prm.bn_y = DGRAD ? (const bf16*)br.y : nullptr;
  prm.bn_mean = br.mean; prm.bn_invstd = br.invstd; prm.bn_gamma = br.gamma; prm.bn_beta = br.beta;
  prm.bn_sums = br.sums;
  prm.bnf_out = nullptr;
  static_assert(KS == 1 || NC == 1, "split-K and multicast clusters are separate variants");
  // epilogue buffers live in the pipeline stages: split-K receive area + bf16 tile (+ the partial sums, which
  // reuse the receive area when there is one)
  static_assert((KS - 1) * (BN / KS) * 128 * 4 + TILE_M * (BN / KS) * 2 + (KS > 1 ? 0 : TILE_M * 2 * 8 * 4) <=
                    STAGES * (A_BYTES + (PAIR ? BN / 2 : BN) * KCHUNK * 2),
                "epilogue buffers (split-K receive area, bf16 tile, partial sums) must fit in the pipeline stages");
  int m_tiles = prm.tiles_per_img > 0 ? B * prm.tiles_per_img : ceil_div(B, prm.imgs_per_tile);
  if (PAIR) m_tiles = (m_tiles + 1) & ~1;        // pairs; a surplus tile lies wholly out of range (zero-filled loads, no stores)
  dim3 grid(m_tiles, Nc / BN, KS);
  constexpr int smem = conv_tc_smem<BN, STAGES, PAIR>();
  auto kern = conv3x3_tc_kernel<BN, STAGES, DGRAD, NC, KS, PAIR>;
  static unsigned long long configured = 0;
  if (first_on_device(&configured)) {
    DDPB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    // ask for the largest shared-memory carve-out, so that two ~100 KB CTAs share an SM (and the
    // occupancy query below reports that)
    DDPB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  }
  if (!DGRAD && !PAIR && ex.bnfwd && ex.bnfwd->out && stats) {
    BnFwdArgs& g_bnfwd = *ex.bnfwd;
    // the fused BatchNorm needs a grid-wide barrier: only when every CTA is resident at the same time
    static int capacity = -1;
    if (capacity < 0) {
      // CTAs that fit on the GPU at once, from the kernel's actual resource use.  (The runtime's occupancy
      // query answers 1 CTA per SM for these ~100 KB kernels although two are resident - the TMA probe's
      // single-wave timing of 296 such CTAs shows it - so the limits are applied by hand.)
      int dev = 0, sms = 0, smem_sm = 0, regs_sm = 0;
      DDPB_CUDA_CHECK(cudaGetDevice(&dev));
      DDPB_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
      DDPB_CUDA_CHECK(cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev));
      DDPB_CUDA_CHECK(cudaDeviceGetAttribute(&regs_sm, cudaDevAttrMaxRegistersPerMultiprocessor, dev));
      cudaFuncAttributes fa;
      DDPB_CUDA_CHECK(cudaFuncGetAttributes(&fa, kern));
      const int by_smem = smem_sm / (smem + (int)fa.sharedSizeBytes + 1024);
      const int regs_cta = ((fa.numRegs * 32 + 255) / 256 * 256) * (192 / 32);
      const int by_regs = regs_sm / (regs_cta > 0 ? regs_cta : 1);
      const int by_tmem = 512 / BN;
      int per_sm = by_smem < by_regs ? by_smem : by_regs;
      if (by_tmem < per_sm) per_sm = by_tmem;
      capacity = per_sm * sms;
      // clusters are placed inside one GPC: leave room for SMs a partially filled GPC cannot use
      if (NC * KS > 1) capacity = capacity / 8 * 7 / (NC * KS) * (NC * KS);
    }
    const long long total = (long long)grid.x * grid.y * grid.z;
    if (getenv("DDPB_DEBUG_OCC")) fprintf(stderr, "[ddpb] conv<%d,%d,NC%d,KS%d> grid %lld capacity %d smem %d\n", BN, STAGES, NC, KS, total, capacity, smem);
    if (total <= capacity) {
      prm.bnf_out = (bf16*)g_bnfwd.out; prm.bnf_winner = (bf16*)g_bnfwd.winner;
      prm.bnf_gamma = g_bnfwd.gamma; prm.bnf_beta = g_bnfwd.beta;
      prm.bnf_save_mean = g_bnfwd.save_mean; prm.bnf_save_invstd = g_bnfwd.save_invstd;
      prm.bnf_running_mean = g_bnfwd.running_mean; prm.bnf_running_var = g_bnfwd.running_var;
      prm.bnf_nbt = g_bnfwd.nbt; prm.bnf_barrier = g_bnfwd.barrier; prm.bnf_total_ctas = (unsigned)total;
      prm.bnf_pool = g_bnfwd.pool; prm.bnf_eps = g_bnfwd.eps; prm.bnf_momentum = g_bnfwd.momentum;
      g_bnfwd.applied = true;
    }
  }
  prm.pdl = kernel_flags(stream, 1, DGRAD ? 128 : 1);
  DDPB_CUDA_CHECK(launch_ex(kern, grid, dim3(192), smem, stream, dim3(PAIR ? 2 : 1, NC, KS), 1, tmA, tmB, bias, stats, prm));
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

// tile_n codes: 61128 / 61256 / 62128 / 62256 = CTA pairs (cta_group::2) with 128 / 256 channels per pair tile,
// without / with 2-way cluster split-K (this file); 4264 / 4464 / 4228 / 4428 = cluster split-K (2 or 4 CTAs per 64- or 128-channel tile);
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

// DDPB_CONV_PAIR: which layers the automatic choice gives to the CTA-pair tiles of THIS file (0 = none).
static int pair_mode() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("DDPB_CONV_PAIR"); v = e ? atoi(e) : 0; }
  return v;
}

static int auto_tile(int B, int H, int W, int n_channels, int k_channels) {
  const long long pixels = (long long)B * H * W;
  if (pair_mode() > 0 && pixels >= 2 * TILE_M) {
    const long long m_pairs = (pixels + 2 * TILE_M - 1) / (2 * TILE_M);
    const int k_total9 = 9 * (k_channels / KCHUNK);
    const int mode = pair_mode();
    const long long min_ctas = 128;
    // CTAs = 2 per pair x n-tiles x split-K
    if ((mode == 1 || mode == 4) && n_channels % 256 == 0 && k_total9 >= 18 && m_pairs * (n_channels / 256) * 4 >= min_ctas) return 62256;
    if ((mode == 2 || mode == 4) && n_channels % 128 == 0 && k_total9 >= 18 && m_pairs * (n_channels / 128) * 4 >= min_ctas) return 62128;
    if (mode == 3 && n_channels % 256 == 0 && m_pairs * (n_channels / 256) * 2 >= min_ctas) return 61256;
    if ((mode == 3 || mode == 5) && n_channels % 128 == 0 && m_pairs * (n_channels / 128) * 2 >= 2 * min_ctas) return 61128;
  }
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
  if (t64 >= 256 || max_ks == 1) return t128 >= sm_count() ? 128 : 64;
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

// One dispatcher for fprop and dgrad: n_ch is the GEMM N dimension (Cout for fprop, Cin for dgrad).
template <bool DGRAD>
static int dispatch(const void* in_v, const void* w_v, void* out_v, const float* bias, float* stats, int B, int H, int W,
                    int Cin, int Cout, int tile_n, cudaStream_t stream, const ConvExtra& ex) {
  int rc = check_shapes(B, H, W, Cin, Cout);
  if (rc) return rc;
  const char* what = DGRAD ? "conv_tc dgrad" : "conv_tc fwd";
  const int n_ch = DGRAD ? Cin : Cout, k_ch = DGRAD ? Cout : Cin;
  if (tile_n == 0) tile_n = auto_tile(B, H, W, n_ch, k_ch);
  const bf16 *in = (const bf16*)in_v, *w = (const bf16*)w_v;
  bf16* out = (bf16*)out_v;
#define DDPB_CONV(BN_, ST_, NC_, KS_) launch<BN_, ST_, DGRAD, NC_, KS_>(in, w, out, bias, stats, B, H, W, Cin, Cout, stream, ex)
  if (tile_n > 60000) {         // 6<splits><tile>: CTA pairs (cta_group::2), optionally with cluster split-K
    const int bn = tile_n % 1000, ks = (tile_n / 1000) % 10;
    DDPB_REQUIRE((bn == 128 || bn == 256) && (ks == 1 || ks == 2), "%s: unknown pair tile code %d", what, tile_n);
    DDPB_REQUIRE(n_ch % bn == 0, "%s: %d channels are not a multiple of pair tile %d", what, n_ch, bn);
    DDPB_REQUIRE(!(ex.bnfwd && ex.bnfwd->out), "%s: pair tiles have no fused BatchNorm forward", what);
#define DDPB_CONVP(BN_, ST_, KS_) launch<BN_, ST_, DGRAD, 1, KS_, true>(in, w, out, bias, stats, B, H, W, Cin, Cout, stream, ex)
    if (bn == 128) return ks == 1 ? DDPB_CONVP(128, 4, 1) : DDPB_CONVP(128, 4, 2);
    return ks == 1 ? DDPB_CONVP(256, 3, 1) : DDPB_CONVP(256, 3, 2);
#undef DDPB_CONVP
  }
  if (tile_n > 2000 && tile_n < 4000) {
    DDPB_REQUIRE(n_ch % (tile_n - 2000) == 0, "%s: %d channels are not a multiple of pair tile %d", what, n_ch, tile_n - 2000);
    DDPB_REQUIRE(!(ex.bnred && ex.bnred->y) && !(ex.bnfwd && ex.bnfwd->out),
                 "%s: pair tiles (DDPB_CONV_2CTA) have no fused BatchNorm epilogues", what);
    return launch_conv_tc2(DGRAD, tile_n - 2000, in, w, out, bias, stats, B, H, W, Cin, Cout, stream);
  }
  if (tile_n > 4000) {          // 4<splits><tile>: cluster split-K
    DDPB_REQUIRE(n_ch % (tile_n % 100 == 28 ? 128 : 64) == 0, "%s: %d channels do not fit tile code %d", what, n_ch, tile_n);
    switch (tile_n) {
      case 4264: return cta_count(B, H, W, n_ch, 64) * 2 <= sm_count() ? DDPB_CONV(64, 8, 1, 2) : DDPB_CONV(64, 4, 1, 2);
      case 4464: return DDPB_CONV(64, 4, 1, 4);
      case 4228: return DDPB_CONV(128, 3, 1, 2);
      case 4428: return DDPB_CONV(128, 3, 1, 4);
      default: DDPB_REQUIRE(false, "%s: unknown split-K tile code %d", what, tile_n);
    }
  }
  if (tile_n > 1000) {          // 1<cluster><tile>: activation tile multicast across a cluster of n-tiles
    const bool deep = cta_count(B, H, W, n_ch, 64) <= sm_count();
    switch (tile_n) {
      case 1864: return deep ? DDPB_CONV(64, 8, 8, 1) : DDPB_CONV(64, 4, 8, 1);
      case 1464: return deep ? DDPB_CONV(64, 8, 4, 1) : DDPB_CONV(64, 4, 4, 1);
      case 1264: return DDPB_CONV(64, 4, 2, 1);
      case 1228: return DDPB_CONV(128, 3, 2, 1);
      case 1428: return DDPB_CONV(128, 3, 4, 1);
      default: DDPB_REQUIRE(false, "%s: unknown multicast tile code %d", what, tile_n);
    }
  }
  DDPB_REQUIRE(n_ch % tile_n == 0, "%s: %d channels are not a multiple of tile %d", what, n_ch, tile_n);
  if (tile_n == 128) return DDPB_CONV(128, 3, 1, 1);
  if (tile_n == 64) {
    // fewer CTAs than SMs: one CTA per SM with a deep pipeline (more bytes in flight per CTA)
    return cta_count(B, H, W, n_ch, 64) <= sm_count() ? DDPB_CONV(64, 8, 1, 1) : DDPB_CONV(64, 4, 1, 1);
  }
#undef DDPB_CONV
  DDPB_REQUIRE(false, "%s: tile_n must be 64 or 128", what);
}

}  // namespace ddpb
DDPB_KTRACE_SETTER(conv_tc)

using namespace ddpb;

// y[B,H,W,Cout] = conv3x3(x[B,H,W,Cin], w[Cout,3,3,Cin]) + bias; bf16 in/out, fp32 accumulate.
// `stats` (fp32 [2*Cout], may be null) ACCUMULATES per-channel sum and sum of squares of y.
extern "C" int ddpb_conv3x3_fwd_tc(const void* x, const void* w_bf16, const float* bias, void* y, float* stats,
                                   int B, int H, int W, int Cin, int Cout, int tile_n, cudaStream_t stream) {
  return dispatch<false>(x, w_bf16, y, bias, stats, B, H, W, Cin, Cout, tile_n, stream, kNoExtra);
}

// dx[B,H,W,Cin] = conv3x3_transpose(dy[B,H,W,Cout], w[Cout,3,3,Cin]); bf16 in/out.
extern "C" int ddpb_conv3x3_dgrad_tc(const void* dy, const void* w_bf16, void* dx, int B, int H, int W, int Cin,
                                     int Cout, int tile_n, cudaStream_t stream) {
  return dispatch<true>(dy, w_bf16, dx, nullptr, nullptr, B, H, W, Cin, Cout, tile_n, stream, kNoExtra);
}

// Either of the two with per-CTA phase time stamps (a tuning aid; no reference counterpart):
// trace[cta][8] = %globaltimer (ns) at CTA start, prologue done (predecessor finished), first stage landed,
// last MMA issued, accumulator complete, split-K partials exchanged, tile stored, CTA done.  Unused slots
// keep what the caller put there.
extern "C" int ddpb_conv3x3_trace(int dgrad, const void* in, const void* w_bf16, const float* bias, void* out,
                                  float* stats, int B, int H, int W, int Cin, int Cout, int tile_n,
                                  unsigned long long* trace, cudaStream_t stream) {
  ConvExtra ex;
  ex.trace = trace;
  return dgrad ? dispatch<true>(in, w_bf16, out, nullptr, nullptr, B, H, W, Cin, Cout, tile_n, stream, ex)
               : dispatch<false>(in, w_bf16, out, bias, stats, B, H, W, Cin, Cout, tile_n, stream, ex);
}

// y[B,H,W,Cout] = x[B,H,W,Cin] * w[Cout,Cin]^T + bias: the same kernel with a single (centre) tap.
// Used for the first VGG layer, whose 3x3x3 patches are laid out as a [pixels][64] matrix first
// (ddpb_im2col3x3_c3) so that it runs on the tensor cores like every other layer.
extern "C" int ddpb_conv1x1_fwd_tc(const void* x, const void* w_bf16, const float* bias, void* y, float* stats,
                                   int B, int H, int W, int Cin, int Cout, cudaStream_t stream) {
  int rc = check_shapes(B, H, W, Cin, Cout);
  if (rc) return rc;
  if (Cout % 128 == 0 && cta_count(B, H, W, Cout, 128) >= sm_count())
    return launch<128, 3, false>((const bf16*)x, (const bf16*)w_bf16, (bf16*)y, bias, stats, B, H, W, Cin, Cout, stream, kNoExtra, 1);
  return launch<64, 4, false>((const bf16*)x, (const bf16*)w_bf16, (bf16*)y, bias, stats, B, H, W, Cin, Cout, stream, kNoExtra, 1);
}

// dgrad as above, and in its epilogue the BatchNorm-backward reduction of the block whose OUTPUT dx is the
// gradient of: sums[0:Cin] += sum dz, sums[Cin:2Cin] += sum dz * xhat, where dz is dx masked by that
// block's ReLU / 2x2 max-pool.  y_win[B,H,W,Cin] (bf16) holds, per element of dx, the conv output of that
// block that won the window (ddpb_bn_relu_pool_fwd_winner; the conv output itself for a block without
// pooling); NaN marks "blocked by ReLU".  Afterwards only ddpb_bn_relu_pool_bwd_apply is needed.
extern "C" int ddpb_conv3x3_dgrad_bnred_tc(const void* dy, const void* w_bf16, void* dx, int B, int H, int W, int Cin,
                                           int Cout, int tile_n, const void* y_win, const float* mean,
                                           const float* invstd, const float* gamma, const float* beta, float* sums,
                                           cudaStream_t stream) {
  DDPB_REQUIRE(y_win && mean && invstd && gamma && beta && sums, "dgrad_bnred: null BatchNorm argument");
  DDPB_REQUIRE(tile_n < 2000 || tile_n > 4000, "dgrad_bnred: the conv_tc2.cu pair tiles (2128 / 2256) have no fused reduction");
  BnRedArgs br;
  br.y = y_win; br.mean = mean; br.invstd = invstd; br.gamma = gamma; br.beta = beta; br.sums = sums;
  ConvExtra ex;
  ex.bnred = &br;
  return dispatch<true>(dy, w_bf16, dx, nullptr, nullptr, B, H, W, Cin, Cout, tile_n, stream, ex);
}

// Forward conv as above and, when the whole grid fits on the GPU at once, this block's BatchNorm (batch
// statistics) + ReLU + optional 2x2 max-pool in the same kernel: after a grid-wide barrier every CTA
// normalises the tile it still holds in shared memory and writes `act` (and `winner`, nullable).
// *fused = 1 when that happened; 0 means only y and stats were produced and the caller must run
// ddpb_bn_relu_pool_fwd(_winner) itself.  `barrier` is one zeroed uint32 per launch; stats must be zeroed.
extern "C" int ddpb_conv3x3_fwd_bn_tc(const void* x, const void* w_bf16, const float* bias, void* y, float* stats,
                                      int B, int H, int W, int Cin, int Cout, int tile_n, const float* gamma,
                                      const float* beta, void* act, void* winner, float* save_mean,
                                      float* save_invstd, float* running_mean, float* running_var,
                                      long long* num_batches_tracked, int pool, float eps, float momentum,
                                      unsigned int* barrier, int* fused, cudaStream_t stream) {
  DDPB_REQUIRE(stats && gamma && beta && act && save_mean && save_invstd && running_mean && running_var && barrier &&
                   fused, "conv_fwd_bn: null argument");
  DDPB_REQUIRE(!pool || (H % 2 == 0 && W % 2 == 0), "conv_fwd_bn: pooling needs even H, W");
  if (tile_n == 0) tile_n = auto_tile(B, H, W, Cout, Cin);
  *fused = 0;
  if (tile_n > 2000 && tile_n < 4000)       // pair tiles have their own epilogue: no fusion
    return ddpb_conv3x3_fwd_tc(x, w_bf16, bias, y, stats, B, H, W, Cin, Cout, tile_n, stream);
  BnFwdArgs bf;
  bf.out = act; bf.winner = winner; bf.gamma = gamma; bf.beta = beta;
  bf.save_mean = save_mean; bf.save_invstd = save_invstd;
  bf.running_mean = running_mean; bf.running_var = running_var; bf.nbt = num_batches_tracked;
  bf.barrier = barrier; bf.pool = pool ? 1 : 0; bf.eps = eps; bf.momentum = momentum;
  ConvExtra ex;
  ex.bnfwd = &bf;
  const int rc = dispatch<false>(x, w_bf16, y, bias, stats, B, H, W, Cin, Cout, tile_n, stream, ex);
  *fused = bf.applied ? 1 : 0;
  return rc;
}
