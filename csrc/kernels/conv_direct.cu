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

}  // namespace ddpb

using namespace ddpb;

// y = conv3x3(x, w) + bias.  x: [B,H,W,Cin], w: fp32 [Cout,3,3,Cin], y: [B,H,W,Cout].
// If `stats` is non-null, per-channel sum / sum-of-squares of y are ACCUMULATED into it.
extern "C" int ddpb_conv3x3_fwd_simt(const void* x, const float* w, const float* bias, void* y, float* stats,
                                     int B, int H, int W, int Cin, int Cout, int dtype,
                                     cudaStream_t stream) {
  long long total = (long long)B * H * W * Cout;
  if (Cin == 3 && Cout == 64) {
    int grid = ceil_div((long long)B * H * W, 64);
    if (grid > 148 * 8) grid = 148 * 8;
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
  if (Cin == 3 && Cout == 64) {
    int grid = ceil_div((long long)B * H * W, 64);
    if (grid > 148 * 2) grid = 148 * 2;
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
