// SIMT 3x3/stride-1/pad-1 convolutions on NHWC activations with KRSC fp32 weights.
//
// Three roles (ref Part 1/model.py:18-23 and its autograd; SURVEY K-C1..K-C3):
//   1. the fp32 parity path on GPUs (exact fp32 FMA accumulation, no tensor cores);
//   2. the numerical oracle the tcgen05 kernels are tested against on the box;
//   3. the first layer (Cin = 3, K = 27): a poor tensor-core fit, handled by dedicated kernels
//      (``first_layer_*``) that keep the 64x27 filter bank in shared memory and fuse the
//      bias and the BN sum / sum-of-squares accumulation.
#include "../common.cuh"
#include "../../include/ddp_b200.h"

namespace ddpb {

// ------------------------------------------------------------------------------------------
// generic kernels: one thread per output element, inner loop over the contiguous channel dim
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
conv_fwd_direct(const T* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                T* __restrict__ y, int B, int H, int W, int Cin, int Cout) {
  long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  long long total = (long long)B * H * W * Cout;
  if (idx >= total) return;
  int co = idx % Cout;
  long long pix = idx / Cout;
  int ww = pix % W, hh = (pix / W) % H;
  long long n = pix / ((long long)W * H);
  float acc = bias ? bias[co] : 0.f;
  for (int r = 0; r < 3; ++r) {
    int ih = hh + r - 1;
    if (ih < 0 || ih >= H) continue;
    for (int s = 0; s < 3; ++s) {
      int iw = ww + s - 1;
      if (iw < 0 || iw >= W) continue;
      const T* xp = x + ((n * H + ih) * W + iw) * Cin;
      const float* wp = w + ((long long)(co * 3 + r) * 3 + s) * Cin;
      for (int ci = 0; ci < Cin; ++ci) acc = fmaf(to_f32(xp[ci]), wp[ci], acc);
    }
  }
  y[idx] = from_f32<T>(acc);
}

// dX[n,h,w,ci] = sum_{r,s,co} dY[n,h+1-r,w+1-s,co] * W[co,r,s,ci]
template <typename T>
__global__ void __launch_bounds__(256)
conv_dgrad_direct(const T* __restrict__ dy, const float* __restrict__ w, T* __restrict__ dx, int B, int H,
                  int W, int Cin, int Cout) {
  long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  long long total = (long long)B * H * W * Cin;
  if (idx >= total) return;
  int ci = idx % Cin;
  long long pix = idx / Cin;
  int ww = pix % W, hh = (pix / W) % H;
  long long n = pix / ((long long)W * H);
  float acc = 0.f;
  for (int r = 0; r < 3; ++r) {
    int oh = hh + 1 - r;
    if (oh < 0 || oh >= H) continue;
    for (int s = 0; s < 3; ++s) {
      int ow = ww + 1 - s;
      if (ow < 0 || ow >= W) continue;
      const T* dp = dy + ((n * H + oh) * W + ow) * Cout;
      const float* wp = w + (long long)(r * 3 + s) * Cin + ci;
      for (int co = 0; co < Cout; ++co) acc = fmaf(to_f32(dp[co]), wp[(long long)co * 9 * Cin], acc);
    }
  }
  dx[idx] = from_f32<T>(acc);
}

// dW[co,r,s,ci] += sum over a pixel chunk of dY[p,co] * X[p shifted by (r-1,s-1), ci]
template <typename T>
__global__ void __launch_bounds__(256)
conv_wgrad_direct(const T* __restrict__ dy, const T* __restrict__ x, float* __restrict__ dw, int B, int H,
                  int W, int Cin, int Cout, int chunk) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  int total = Cout * 9 * Cin;
  if (idx >= total) return;
  int ci = idx % Cin;
  int tap = (idx / Cin) % 9;
  int co = idx / (9 * Cin);
  int r = tap / 3, s = tap % 3;
  long long P = (long long)B * H * W;
  long long p0 = (long long)blockIdx.y * chunk, p1 = p0 + chunk < P ? p0 + chunk : P;
  float acc = 0.f;
  for (long long p = p0; p < p1; ++p) {
    int ww = p % W, hh = (p / W) % H;
    int ih = hh + r - 1, iw = ww + s - 1;
    if (ih < 0 || ih >= H || iw < 0 || iw >= W) continue;
    long long q = p + (long long)(r - 1) * W + (s - 1);
    acc = fmaf(to_f32(dy[p * Cout + co]), to_f32(x[q * Cin + ci]), acc);
  }
  atomicAdd(&dw[idx], acc);
}

// ------------------------------------------------------------------------------------------
// first layer, forward: Cin = 3, Cout = 64.  256 threads = 64 pixels x 4 channel groups of 16;
// all lanes of a warp share the channel group, so filter reads are warp-broadcast LDS.128.
// Fuses: bias add, rounding to the activation dtype, per-channel sum / sum-of-squares.
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
first_layer_fwd(const T* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                T* __restrict__ y, float* __restrict__ stats, int B, int H, int W) {
  constexpr int CO = 64, K = 27, CG = 16;
  __shared__ __align__(16) float wt[K][CO];     // transposed filter bank: [tap*3+ci][co]
  __shared__ float sb[CO];
  __shared__ float sred[2 * CO];
  for (int i = threadIdx.x; i < K * CO; i += blockDim.x) {
    int co = i / K, k = i % K;
    wt[k][co] = w[i];
  }
  if (threadIdx.x < CO) sb[threadIdx.x] = bias ? bias[threadIdx.x] : 0.f;
  if (threadIdx.x < 2 * CO) sred[threadIdx.x] = 0.f;
  __syncthreads();
  const int pl = threadIdx.x & 63, grp = threadIdx.x >> 6, c0 = grp * CG;
  const long long P = (long long)B * H * W;
  float s1[CG], s2[CG];
#pragma unroll
  for (int j = 0; j < CG; ++j) { s1[j] = 0.f; s2[j] = 0.f; }
  for (long long p = (long long)blockIdx.x * 64 + pl; p < P; p += (long long)gridDim.x * 64) {
    int ww = p % W, hh = (p / W) % H;
    float acc[CG];
#pragma unroll
    for (int j = 0; j < CG; ++j) acc[j] = sb[c0 + j];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      int ih = hh + r - 1;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        int iw = ww + s - 1;
        bool ok = ih >= 0 && ih < H && iw >= 0 && iw < W;
        const T* xp = x + (p + (long long)(r - 1) * W + (s - 1)) * 3;
#pragma unroll
        for (int ci = 0; ci < 3; ++ci) {
          float xv = ok ? to_f32(xp[ci]) : 0.f;
          const float4* wr = reinterpret_cast<const float4*>(&wt[(r * 3 + s) * 3 + ci][c0]);
#pragma unroll
          for (int q = 0; q < CG / 4; ++q) {
            float4 wv = wr[q];
            acc[4 * q + 0] = fmaf(xv, wv.x, acc[4 * q + 0]);
            acc[4 * q + 1] = fmaf(xv, wv.y, acc[4 * q + 1]);
            acc[4 * q + 2] = fmaf(xv, wv.z, acc[4 * q + 2]);
            acc[4 * q + 3] = fmaf(xv, wv.w, acc[4 * q + 3]);
          }
        }
      }
    }
    // round to the storage dtype first: the statistics must describe what BN will read back
    T* yp = y + p * CO + c0;
    constexpr int V = 16 / sizeof(T);
#pragma unroll
    for (int q = 0; q < CG / V; ++q) {
      float tmp[V];
#pragma unroll
      for (int j = 0; j < V; ++j) {
        tmp[j] = to_f32(from_f32<T>(acc[q * V + j]));
        s1[q * V + j] += tmp[j];
        s2[q * V + j] += tmp[j] * tmp[j];
      }
      store_vec<T>(yp + q * V, tmp);
    }
  }
  if (stats) {
#pragma unroll
    for (int j = 0; j < CG; ++j) {
      float a = warp_sum(s1[j]), b = warp_sum(s2[j]);
      if ((threadIdx.x & 31) == 0) { atomicAdd(&sred[c0 + j], a); atomicAdd(&sred[CO + c0 + j], b); }
    }
    __syncthreads();
    if (threadIdx.x < 2 * CO) atomicAdd(&stats[threadIdx.x], sred[threadIdx.x]);
  }
}

// ------------------------------------------------------------------------------------------
// first layer, wgrad: dW[64][27] += dY^T (64 x P) * im2col(X) (P x 27).
// Per iteration a block stages 64 pixels: dY tile [64][64] and the im2col patch [64][28] in
// shared memory; thread = (4 output channels) x (7 taps) x (1 of 4 pixel streams).
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
first_layer_wgrad(const T* __restrict__ dy, const T* __restrict__ x, float* __restrict__ dw, int B, int H,
                  int W) {
  constexpr int CO = 64, K = 27, KP = 28, TP = 64;
  __shared__ __align__(16) float sdy[TP][CO];
  __shared__ float sx[TP][KP];
  __shared__ float sacc[CO * KP];
  for (int i = threadIdx.x; i < CO * KP; i += blockDim.x) sacc[i] = 0.f;
  const int co4 = (threadIdx.x & 15) * 4, tg = (threadIdx.x >> 4) & 3, stream = threadIdx.x >> 6;
  const long long P = (long long)B * H * W;
  float acc[4][7];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 7; ++b) acc[a][b] = 0.f;
  for (long long base = (long long)blockIdx.x * TP; base < P; base += (long long)gridDim.x * TP) {
    __syncthreads();
    for (int i = threadIdx.x; i < TP * CO; i += blockDim.x) {
      long long p = base + i / CO;
      sdy[i / CO][i % CO] = p < P ? to_f32(dy[p * CO + (i % CO)]) : 0.f;
    }
    for (int i = threadIdx.x; i < TP * KP; i += blockDim.x) {
      int pl = i / KP, k = i % KP;
      long long p = base + pl;
      float v = 0.f;
      if (p < P && k < K) {
        int tap = k / 3, ci = k % 3, r = tap / 3, s = tap % 3;
        int ww = p % W, hh = (p / W) % H;
        int ih = hh + r - 1, iw = ww + s - 1;
        if (ih >= 0 && ih < H && iw >= 0 && iw < W)
          v = to_f32(x[(p + (long long)(r - 1) * W + (s - 1)) * 3 + ci]);
      }
      sx[pl][k] = v;
    }
    __syncthreads();
#pragma unroll 4
    for (int i = 0; i < TP / 4; ++i) {
      int pl = stream * (TP / 4) + i;
      float4 d = *reinterpret_cast<const float4*>(&sdy[pl][co4]);
#pragma unroll
      for (int b = 0; b < 7; ++b) {
        float xv = sx[pl][tg * 7 + b];
        acc[0][b] = fmaf(d.x, xv, acc[0][b]);
        acc[1][b] = fmaf(d.y, xv, acc[1][b]);
        acc[2][b] = fmaf(d.z, xv, acc[2][b]);
        acc[3][b] = fmaf(d.w, xv, acc[3][b]);
      }
    }
  }
  __syncthreads();
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 7; ++b) atomicAdd(&sacc[(co4 + a) * KP + tg * 7 + b], acc[a][b]);
  __syncthreads();
  for (int i = threadIdx.x; i < CO * K; i += blockDim.x) {
    int co = i / K, k = i % K;
    atomicAdd(&dw[i], sacc[co * KP + k]);
  }
}

// ------------------------------------------------------------------------------------------
// first layer on 32x32 images (the CIFAR case), tuned: a tile is 4 image rows (128 pixels); the
// 6 input rows it needs are staged once in shared memory as fp32 with a zero halo, so no thread
// touches global memory for x again and the conv's padding needs no branches.
// ------------------------------------------------------------------------------------------
constexpr int FL_W = 32, FL_ROWS = 4, FL_TILE = FL_W * FL_ROWS;     // 128 pixels
constexpr int FL_XR = FL_ROWS + 2, FL_XW = FL_W + 2;                // staged rows x (cols + halo)

// Staging is split in two so the global loads of tile t+1 are in flight while tile t is computed:
// fl_prefetch_x issues the loads into registers (threads 0..203 own one (row, col) position each),
// fl_commit_x writes them to the other shared-memory buffer once the compute is done.
struct FlPix { float v0, v1, v2; };

template <typename T>
__device__ __forceinline__ FlPix fl_prefetch_x(const T* __restrict__ x, long long tile, long long tiles, int H) {
  // tile -> image n, first row h0; position (r, c) <- x[n, h0-1+r, c-1, :] or 0 outside the image
  FlPix px{0.f, 0.f, 0.f};
  const int i = threadIdx.x;
  if (tile < tiles && i < FL_XR * FL_XW) {
    const int tiles_per_img = H / FL_ROWS;
    const long long n = tile / tiles_per_img;
    const int h0 = (int)(tile % tiles_per_img) * FL_ROWS;
    const int r = i / FL_XW, c = i % FL_XW;
    const int ih = h0 - 1 + r, iw = c - 1;
    if (ih >= 0 && ih < H && iw >= 0 && iw < FL_W) {
      const T* p = x + ((n * H + ih) * FL_W + iw) * 3;
      px.v0 = to_f32(p[0]); px.v1 = to_f32(p[1]); px.v2 = to_f32(p[2]);
    }
  }
  return px;
}

__device__ __forceinline__ void fl_commit_x(float (*xs)[FL_XW][3], const FlPix& px) {
  const int i = threadIdx.x;
  if (i < FL_XR * FL_XW) {
    const int r = i / FL_XW, c = i % FL_XW;
    xs[r][c][0] = px.v0; xs[r][c][1] = px.v1; xs[r][c][2] = px.v2;
  }
}

// forward: thread = 2 vertically adjacent pixels x 16 output channels (32 accumulators);
// warps share the channel quarter, so weight reads are broadcast LDS.128.
template <typename T>
__global__ void __launch_bounds__(256, 1)      // 32 acc + 32 stat registers per thread: no spills at 1 CTA/SM
first_layer_fwd_32(const T* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                   T* __restrict__ y, float* __restrict__ stats, int B, int H) {
  constexpr int CO = 64, K = 27, CG = 16;
  __shared__ __align__(16) float wt[K][CO];
  __shared__ float sb[CO];
  __shared__ float xs[2][FL_XR][FL_XW][3];
  __shared__ float sred[8][2 * CG];
  for (int i = threadIdx.x; i < K * CO; i += blockDim.x) wt[i % K][i / K] = w[i];
  if (threadIdx.x < CO) sb[threadIdx.x] = bias ? bias[threadIdx.x] : 0.f;
  const int pp = threadIdx.x & 63, cq = threadIdx.x >> 6, c0 = cq * CG;
  const int wcol = pp & 31, rp = pp >> 5;               // pixel pair: rows 2*rp and 2*rp+1, column wcol
  const long long tiles = (long long)B * (H / FL_ROWS);
  float s1[CG], s2[CG];
#pragma unroll
  for (int j = 0; j < CG; ++j) { s1[j] = 0.f; s2[j] = 0.f; }
  int buf = 0;
  fl_commit_x(xs[0], fl_prefetch_x<T>(x, blockIdx.x, tiles, H));
  __syncthreads();
  for (long long tile = blockIdx.x; tile < tiles; tile += gridDim.x, buf ^= 1) {
    const FlPix nxt = fl_prefetch_x<T>(x, tile + gridDim.x, tiles, H);      // in flight during the FMAs below
    float acc[2][CG];
#pragma unroll
    for (int j = 0; j < CG; ++j) { acc[0][j] = sb[c0 + j]; acc[1][j] = acc[0][j]; }
#pragma unroll
    for (int r = 0; r < 4; ++r) {            // input rows 2*rp + r feed output row 0 (r<3) and 1 (r>0)
#pragma unroll
      for (int s = 0; s < 3; ++s) {
#pragma unroll
        for (int ci = 0; ci < 3; ++ci) {
          const float xv = xs[buf][2 * rp + r][wcol + s][ci];
          if (r < 3) {
            const float4* wr = reinterpret_cast<const float4*>(&wt[(r * 3 + s) * 3 + ci][c0]);
#pragma unroll
            for (int q = 0; q < CG / 4; ++q) {
              const float4 wv = wr[q];
              acc[0][4 * q + 0] = fmaf(xv, wv.x, acc[0][4 * q + 0]);
              acc[0][4 * q + 1] = fmaf(xv, wv.y, acc[0][4 * q + 1]);
              acc[0][4 * q + 2] = fmaf(xv, wv.z, acc[0][4 * q + 2]);
              acc[0][4 * q + 3] = fmaf(xv, wv.w, acc[0][4 * q + 3]);
            }
          }
          if (r > 0) {
            const float4* wr = reinterpret_cast<const float4*>(&wt[((r - 1) * 3 + s) * 3 + ci][c0]);
#pragma unroll
            for (int q = 0; q < CG / 4; ++q) {
              const float4 wv = wr[q];
              acc[1][4 * q + 0] = fmaf(xv, wv.x, acc[1][4 * q + 0]);
              acc[1][4 * q + 1] = fmaf(xv, wv.y, acc[1][4 * q + 1]);
              acc[1][4 * q + 2] = fmaf(xv, wv.z, acc[1][4 * q + 2]);
              acc[1][4 * q + 3] = fmaf(xv, wv.w, acc[1][4 * q + 3]);
            }
          }
        }
      }
    }
    constexpr int V = 16 / sizeof(T);
#pragma unroll
    for (int o = 0; o < 2; ++o) {
      const long long p = tile * FL_TILE + (2 * rp + o) * FL_W + wcol;
      T* yp = y + p * CO + c0;
#pragma unroll
      for (int q = 0; q < CG / V; ++q) {
        float tmp[V];
#pragma unroll
        for (int j = 0; j < V; ++j) {
          tmp[j] = to_f32(from_f32<T>(acc[o][q * V + j]));      // statistics of the stored values
          s1[q * V + j] += tmp[j];
          s2[q * V + j] += tmp[j] * tmp[j];
        }
        store_vec<T>(yp + q * V, tmp);
      }
    }
    fl_commit_x(xs[buf ^ 1], nxt);
    __syncthreads();           // next tile staged; everyone is done reading xs[buf]
  }
  if (stats) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int j = 0; j < CG; ++j) {
      const float a = warp_sum(s1[j]), b = warp_sum(s2[j]);
      if (lane == 0) { sred[warp][j] = a; sred[warp][CG + j] = b; }
    }
    __syncthreads();
    if (threadIdx.x < 2 * CO) {          // channel c lives in warps 2*(c/16) and 2*(c/16)+1
      const int c = threadIdx.x % CO, which = threadIdx.x / CO;
      const int q = c / CG, j = c % CG;
      atomicAdd(&stats[which * CO + c], sred[2 * q][which * CG + j] + sred[2 * q + 1][which * CG + j]);
    }
  }
}

// wgrad: dW[64][27] += dY^T * im2col(X).  Per tile: dY [128][64] (kept in its storage dtype, copied
// with cp.async into a double buffer) and the 6 input rows are staged in shared memory while the
// previous tile is being multiplied; thread = 8 output channels x 7 taps (56 accumulators) over 1
// of 8 pixel streams.
__device__ __forceinline__ void cp_async_16(void* smem, const void* gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

template <typename T>
__device__ __forceinline__ void fl_prefetch_dy(T* sdst, const T* __restrict__ dy, long long tile, long long tiles) {
  constexpr int V = 16 / sizeof(T), CO = 64;
  if (tile < tiles)
    for (int i = threadIdx.x; i < FL_TILE * CO / V; i += blockDim.x)
      cp_async_16(sdst + (size_t)i * V, dy + tile * FL_TILE * CO + (long long)i * V);
  cp_async_commit();
}

template <typename T>
__global__ void __launch_bounds__(256)
first_layer_wgrad_32(const T* __restrict__ dy, const T* __restrict__ x, float* __restrict__ dw, int B, int H) {
  constexpr int CO = 64, K = 27, KP = 28, V = 16 / sizeof(T);
  extern __shared__ __align__(16) uint8_t fl_smem[];
  T* sdy = reinterpret_cast<T*>(fl_smem);                                   // [2][FL_TILE][CO]
  float (*xs)[FL_XR][FL_XW][3] =
      reinterpret_cast<float (*)[FL_XR][FL_XW][3]>(fl_smem + 2 * FL_TILE * CO * sizeof(T));   // [2][6][34][3]
  const int lane = threadIdx.x & 31, stream = threadIdx.x >> 5;     // one warp per pixel stream
  const int cg = lane & 7, tg = lane >> 3;                          // 8 channel groups x 4 tap groups
  // offsets (relative to the pixel's top-left patch corner) of this thread's 7 patch elements
  int koff[7];
#pragma unroll
  for (int b = 0; b < 7; ++b) {
    const int k = tg * 7 + b;
    const int kk = k < K ? k : 0;                       // k = 27 is padding: its column is never copied out
    const int tap = kk / 3, ci = kk % 3, r = tap / 3, s = tap % 3;
    koff[b] = (r * FL_XW + s) * 3 + ci;
  }
  float acc[8][7];
#pragma unroll
  for (int a = 0; a < 8; ++a)
#pragma unroll
    for (int b = 0; b < 7; ++b) acc[a][b] = 0.f;
  const long long tiles = (long long)B * (H / FL_ROWS);
  int buf = 0;
  fl_prefetch_dy<T>(sdy, dy, blockIdx.x, tiles);
  fl_commit_x(xs[0], fl_prefetch_x<T>(x, blockIdx.x, tiles, H));
  cp_async_wait_all();
  __syncthreads();
  for (long long tile = blockIdx.x; tile < tiles; tile += gridDim.x, buf ^= 1) {
    fl_prefetch_dy<T>(sdy + (buf ^ 1) * FL_TILE * CO, dy, tile + gridDim.x, tiles);
    const FlPix nxt = fl_prefetch_x<T>(x, tile + gridDim.x, tiles, H);
    const float* xflat = &xs[buf][0][0][0];
    const T* dtile = sdy + buf * FL_TILE * CO;
#pragma unroll 2
    for (int i = 0; i < FL_TILE / 8; ++i) {
      const int pl = stream * (FL_TILE / 8) + i;            // pixel inside the tile
      const int base = ((pl / FL_W) * FL_XW + (pl % FL_W)) * 3;
      float d[8];
#pragma unroll
      for (int q = 0; q < 8 / V; ++q) load_vec<T>(dtile + pl * CO + cg * 8 + q * V, d + q * V);
#pragma unroll
      for (int b = 0; b < 7; ++b) {
        const float xv = xflat[base + koff[b]];
#pragma unroll
        for (int a = 0; a < 8; ++a) acc[a][b] = fmaf(d[a], xv, acc[a][b]);
      }
    }
    fl_commit_x(xs[buf ^ 1], nxt);
    cp_async_wait_all();
    __syncthreads();
  }
  // combine the 8 pixel streams (deterministic order), then one atomic per output element
  float* sacc = reinterpret_cast<float*>(fl_smem);        // reuse: [64][28]
  for (int s = 0; s < 8; ++s) {
    if (stream == s) {
#pragma unroll
      for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 7; ++b) {
          float* dst = &sacc[(cg * 8 + a) * KP + tg * 7 + b];
          *dst = (s == 0 ? 0.f : *dst) + acc[a][b];
        }
    }
    __syncthreads();
  }
  for (int i = threadIdx.x; i < CO * K; i += blockDim.x) atomicAdd(&dw[i], sacc[(i / K) * KP + (i % K)]);
}

template <typename T>
static int launch_first_layer_wgrad_32(const void* dy, const void* x, float* dw, int B, int H, cudaStream_t stream) {
  const int smem = 2 * FL_TILE * 64 * (int)sizeof(T) + 2 * FL_XR * FL_XW * 3 * (int)sizeof(float);
  static unsigned long long configured = 0;
  if (first_on_device(&configured)) {
    DDPB_CUDA_CHECK(cudaFuncSetAttribute(first_layer_wgrad_32<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  }
  long long tiles = (long long)B * (H / FL_ROWS);
  int grid = (int)(tiles < sm_count() * 2 ? tiles : sm_count() * 2);
  first_layer_wgrad_32<T><<<grid, 256, smem, stream>>>((const T*)dy, (const T*)x, dw, B, H);
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}

}  // namespace ddpb

using namespace ddpb;

// y = conv3x3(x, w) + bias.  x: [B,H,W,Cin], w: fp32 [Cout,3,3,Cin], y: [B,H,W,Cout].
// If `stats` is non-null, per-channel sum / sum-of-squares of y are ACCUMULATED into it.
extern "C" int ddpb_conv3x3_fwd_simt(const void* x, const float* w, const float* bias, void* y, float* stats,
                                     int B, int H, int W, int Cin, int Cout, int dtype,
                                     cudaStream_t stream) {
  long long total = (long long)B * H * W * Cout;
  if (Cin == 3 && Cout == 64 && W == FL_W && H % FL_ROWS == 0) {
    long long tiles = (long long)B * (H / FL_ROWS);
    int grid = (int)(tiles < sm_count() * 2 ? tiles : sm_count() * 2);
    if (dtype == DDPB_BF16)
      first_layer_fwd_32<bf16><<<grid, 256, 0, stream>>>((const bf16*)x, w, bias, (bf16*)y, stats, B, H);
    else
      first_layer_fwd_32<float><<<grid, 256, 0, stream>>>((const float*)x, w, bias, (float*)y, stats, B, H);
    DDPB_LAUNCH_CHECK();
    return DDPB_OK;
  }
  if (Cin == 3 && Cout == 64) {
    int grid = ceil_div((long long)B * H * W, 64);
    if (grid > sm_count() * 8) grid = sm_count() * 8;
    if (dtype == DDPB_BF16)
      first_layer_fwd<bf16><<<grid, 256, 0, stream>>>((const bf16*)x, w, bias, (bf16*)y, stats, B, H, W);
    else
      first_layer_fwd<float><<<grid, 256, 0, stream>>>((const float*)x, w, bias, (float*)y, stats, B, H, W);
    DDPB_LAUNCH_CHECK();
    return DDPB_OK;
  }
  int blocks = ceil_div(total, 256);
  if (dtype == DDPB_BF16)
    conv_fwd_direct<bf16><<<blocks, 256, 0, stream>>>((const bf16*)x, w, bias, (bf16*)y, B, H, W, Cin, Cout);
  else
    conv_fwd_direct<float><<<blocks, 256, 0, stream>>>((const float*)x, w, bias, (float*)y, B, H, W, Cin, Cout);
  DDPB_LAUNCH_CHECK();
  if (stats) return ddpb_channel_stats(y, stats, (long long)B * H * W, Cout, dtype, stream);
  return DDPB_OK;
}

extern "C" int ddpb_conv3x3_dgrad_simt(const void* dy, const float* w, void* dx, int B, int H, int W, int Cin,
                                       int Cout, int dtype, cudaStream_t stream) {
  long long total = (long long)B * H * W * Cin;
  int blocks = ceil_div(total, 256);
  if (dtype == DDPB_BF16)
    conv_dgrad_direct<bf16><<<blocks, 256, 0, stream>>>((const bf16*)dy, w, (bf16*)dx, B, H, W, Cin, Cout);
  else
    conv_dgrad_direct<float><<<blocks, 256, 0, stream>>>((const float*)dy, w, (float*)dx, B, H, W, Cin, Cout);
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}

// dw (fp32 [Cout,3,3,Cin]) is ACCUMULATED into: zero it first.
extern "C" int ddpb_conv3x3_wgrad_simt(const void* dy, const void* x, float* dw, int B, int H, int W, int Cin,
                                       int Cout, int dtype, cudaStream_t stream) {
  if (Cin == 3 && Cout == 64 && W == FL_W && H % FL_ROWS == 0) {
    return dtype == DDPB_BF16 ? launch_first_layer_wgrad_32<bf16>(dy, x, dw, B, H, stream)
                              : launch_first_layer_wgrad_32<float>(dy, x, dw, B, H, stream);
  }
  if (Cin == 3 && Cout == 64) {
    int grid = ceil_div((long long)B * H * W, 64);
    if (grid > sm_count() * 2) grid = sm_count() * 2;
    if (dtype == DDPB_BF16)
      first_layer_wgrad<bf16><<<grid, 256, 0, stream>>>((const bf16*)dy, (const bf16*)x, dw, B, H, W);
    else
      first_layer_wgrad<float><<<grid, 256, 0, stream>>>((const float*)dy, (const float*)x, dw, B, H, W);
    DDPB_LAUNCH_CHECK();
    return DDPB_OK;
  }
  long long P = (long long)B * H * W;
  int chunk = 512;
  dim3 grid(ceil_div((long long)Cout * 9 * Cin, 256), ceil_div(P, chunk));
  if (dtype == DDPB_BF16)
    conv_wgrad_direct<bf16><<<grid, 256, 0, stream>>>((const bf16*)dy, (const bf16*)x, dw, B, H, W, Cin, Cout, chunk);
  else
    conv_wgrad_direct<float><<<grid, 256, 0, stream>>>((const float*)dy, (const float*)x, dw, B, H, W, Cin, Cout, chunk);
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}
