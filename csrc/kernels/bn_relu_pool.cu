// BatchNorm2d (+ReLU, + optional 2x2/2 max-pool) forward and backward on NHWC tensors.
//
// Replaces the reference's BatchNorm2d -> ReLU(inplace) [-> MaxPool2d(2,2)] module triplets
// (ref Part 1/model.py:16,24-25; SURVEY K-N1/K-N2/K-N3).  The per-channel batch statistics
// (sum, sum of squares) arrive already reduced: the convolution epilogue accumulates them
// (csrc/kernels/conv_tc.cu), so forward is ONE pass over the conv output:
//   y -> normalise + affine + ReLU (+ pool) -> next layer's input,
// plus the running-stat update (momentum 0.1, unbiased variance) done by block 0.
// Backward is two passes (the two per-channel reductions BN needs, then dX); the pool routing
// and ReLU mask are recomputed from the saved conv output instead of being stored.
#include <mutex>

#include "../common.cuh"
#include "../../include/ddp_b200.h"

namespace ddpb {

template <typename T> struct VecOf { static constexpr int N = 16 / sizeof(T); };

// 16 raw bytes (8 bf16 or 4 fp32) fetched with a coherent load; unpacked to floats when used.  Keeping the
// packed form lets the kernels issue the NEXT iteration's loads (and the first iteration's, before the
// per-channel constants are even read) without doubling their register count.
struct Raw16 { uint4 u; };
template <typename T> __device__ __forceinline__ Raw16 ld_raw(const T* p) {
  Raw16 r; r.u = DDPB_LD_COH(reinterpret_cast<const uint4*>(p)); return r;
}
template <typename T> __device__ __forceinline__ void unpack_raw(const Raw16& r, float* out);
template <> __device__ __forceinline__ void unpack_raw<float>(const Raw16& r, float* out) {
  out[0] = __uint_as_float(r.u.x); out[1] = __uint_as_float(r.u.y);
  out[2] = __uint_as_float(r.u.z); out[3] = __uint_as_float(r.u.w);
}
template <> __device__ __forceinline__ void unpack_raw<bf16>(const Raw16& r, float* out) {
  const float2 a = unpack_bf16x2(r.u.x), b = unpack_bf16x2(r.u.y), c = unpack_bf16x2(r.u.z), d = unpack_bf16x2(r.u.w);
  out[0] = a.x; out[1] = a.y; out[2] = b.x; out[3] = b.y; out[4] = c.x; out[5] = c.y; out[6] = d.x; out[7] = d.y;
}

// decode output pixel -> the (up to) four input pixel indices.  32-bit arithmetic on purpose: the
// 64-bit divisions this replaced cost more instructions than the rest of the loop body.  H and W
// are powers of two for every layer of the VGG family (shift path); other sizes take the divide.
struct PoolGeom {
  int H, W, ow_shift, oh_shift;     // shifts are -1 when OW / OH are not powers of two
};

__host__ __device__ inline int log2_exact(int v) {
  int s = 0;
  while ((1 << s) < v) ++s;
  return (1 << s) == v ? s : -1;
}

template <bool POOL>
__device__ __forceinline__ void window(int opix, const PoolGeom& gm, int* in) {
  if (!POOL) { in[0] = opix; return; }
  int ow, oh, n;
  if (gm.ow_shift >= 0 && gm.oh_shift >= 0) {
    ow = opix & ((1 << gm.ow_shift) - 1);
    const int r = opix >> gm.ow_shift;
    oh = r & ((1 << gm.oh_shift) - 1);
    n = r >> gm.oh_shift;
  } else {
    const unsigned OW = gm.W >> 1, OH = gm.H >> 1;
    ow = (unsigned)opix % OW;
    const unsigned r = (unsigned)opix / OW;
    oh = r % OH;
    n = r / OH;
  }
  const int base = (n * gm.H + 2 * oh) * gm.W + 2 * ow;
  in[0] = base; in[1] = base + 1; in[2] = base + gm.W; in[3] = base + gm.W + 1;
}

// Sum per-thread channel partials over the block's pixel lanes and add the result to gout[0..C).
// Plain shared-memory stores + a strided read instead of shared float atomics (those compile to
// compare-and-swap loops and serialise badly when 32 lanes hit one channel).
// sm: [ppb * C] floats.  All threads of the block must call this.
template <int V>
__device__ __forceinline__ void block_reduce_channels(float* sm, const float* a, int c0, int pl, int ppb, int C,
                                                      float* __restrict__ gout) {
#pragma unroll
  for (int j = 0; j < V; ++j) sm[pl * C + c0 + j] = a[j];
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s = 0.f;
    for (int q = 0; q < ppb; ++q) s += sm[q * C + c];
    if (gout) atomicAdd(&gout[c], s);
  }
  __syncthreads();
}

// ------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------
template <typename T, bool POOL>
__global__ void __launch_bounds__(256)
bn_fwd_kernel(const T* __restrict__ y, const float* __restrict__ stats, const float* __restrict__ gamma,
              const float* __restrict__ beta, T* __restrict__ out, float* __restrict__ save_mean,
              float* __restrict__ save_invstd, float* __restrict__ running_mean,
              float* __restrict__ running_var, long long* __restrict__ nbt, T* __restrict__ winner,
              long long P, long long Pout, int C, int H, int W, float eps, float momentum, int training, int pdl) {
  constexpr int V = VecOf<T>::N;
  const int groups = C / V;
  const int g = threadIdx.x % groups;
  const int pl = threadIdx.x / groups;
  const int ppb = blockDim.x / groups;
  const int c0 = g * V;
  const PoolGeom gm{H, W, log2_exact(W >> 1), log2_exact(H >> 1)};
  if (pdl & 1) pdl_launch_dependents();     // programmatic dependent launch: see common.cuh
  pdl_wait();
  ktrace_begin(pdl);      // stamped once the predecessor has finished: the start of this kernel's own work

  // software pipeline: the first window is requested before the per-channel constants, every later one
  // while the previous window is being processed (small layers are one dependent-latency chain long)
  constexpr int NQ = POOL ? 4 : 1;
  const int stride = gridDim.x * ppb;
  int op = blockIdx.x * ppb + pl;
  Raw16 nxt[NQ];
  if (pl < ppb && op < (int)Pout) {
    int in[4];
    window<POOL>(op, gm, in);
#pragma unroll
    for (int q = 0; q < NQ; ++q) nxt[q] = ld_raw<T>(y + (size_t)in[q] * C + c0);
  }

  float sc[V], sh[V];
  {
    float a[V], b[V], ga[V], be[V];
    if (training) { ld_coh_vec<V>(stats + c0, a); ld_coh_vec<V>(stats + C + c0, b); }
    else { ld_coh_vec<V>(running_mean + c0, a); ld_coh_vec<V>(running_var + c0, b); }
    ld_coh_vec<V>(gamma + c0, ga);
    ld_coh_vec<V>(beta + c0, be);
#pragma unroll
    for (int j = 0; j < V; ++j) {
      float mean, var;
      if (training) {
        mean = a[j] / (float)P;
        var = fmaxf(b[j] / (float)P - mean * mean, 0.f);
      } else {
        mean = a[j];
        var = b[j];
      }
      float invstd = rsqrtf(var + eps);
      sc[j] = ga[j] * invstd;
      sh[j] = be[j] - mean * sc[j];
      if (training && blockIdx.x == 0 && pl == 0) {
        save_mean[c0 + j] = mean;
        save_invstd[c0 + j] = invstd;
        float unbiased = P > 1 ? var * ((float)P / (float)(P - 1)) : var;
        running_mean[c0 + j] = (1.f - momentum) * ld_coh(&running_mean[c0 + j]) + momentum * mean;
        running_var[c0 + j] = (1.f - momentum) * ld_coh(&running_var[c0 + j]) + momentum * unbiased;
      }
    }
  }
  if (training && blockIdx.x == 0 && threadIdx.x == 0 && nbt) *nbt += 1;
  if (pl >= ppb) return;

  for (; op < (int)Pout; op += stride) {
    Raw16 cur[NQ];
#pragma unroll
    for (int q = 0; q < NQ; ++q) cur[q] = nxt[q];
    if (op + stride < (int)Pout) {
      int in[4];
      window<POOL>(op + stride, gm, in);
#pragma unroll
      for (int q = 0; q < NQ; ++q) nxt[q] = ld_raw<T>(y + (size_t)in[q] * C + c0);
    }
    float best[V], by[V];
#pragma unroll
    for (int j = 0; j < V; ++j) { best[j] = 0.f; by[j] = __int_as_float(0x7fc00000); }   // ReLU floor; NaN = "no winner"
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
      float v[V];
      unpack_raw<T>(cur[q], v);
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const float z = fmaf(v[j], sc[j], sh[j]);
        if (z > best[j]) { best[j] = z; by[j] = v[j]; }      // strict >: the first maximum wins, like the backward pass
      }
    }
    store_vec<T>(out + (size_t)op * C + c0, best);
    // the conv output that won the window (NaN when ReLU blocked all of them): lets the next layer's
    // dgrad epilogue do this block's BN-backward reduction from one value per output pixel
    if (winner) store_vec<T>(winner + (size_t)op * C + c0, by);
  }
  if (pdl & 2) pdl_launch_dependents();     // late trigger (see kernel_flags)
  ktrace_end(pdl);
}

// ------------------------------------------------------------------------------------------
// backward pass 1: s1[c] = sum dz, s2[c] = sum dz * xhat      (dz = dOut routed through pool+ReLU)
// ------------------------------------------------------------------------------------------
template <typename T, bool POOL>
__global__ void __launch_bounds__(256)
bn_bwd_reduce_kernel(const T* __restrict__ dout, const T* __restrict__ y, const float* __restrict__ mean,
                     const float* __restrict__ invstd, const float* __restrict__ gamma,
                     const float* __restrict__ beta, float* __restrict__ sums, long long Pout, int C, int H,
                     int W, int pdl) {
  constexpr int V = VecOf<T>::N;
  extern __shared__ float sm[];      // [ppb*C]
  const int groups = C / V;
  const int g = threadIdx.x % groups;
  const int pl = threadIdx.x / groups;
  const int ppb = blockDim.x / groups;
  const int c0 = g * V;
  const PoolGeom gm{H, W, log2_exact(W >> 1), log2_exact(H >> 1)};
  if (pdl & 1) pdl_launch_dependents();
  pdl_wait();
  ktrace_begin(pdl);      // stamped once the predecessor has finished: the start of this kernel's own work
  float sc[V], sh[V], mu[V], is[V], a1[V], a2[V];
  ld_coh_vec<V>(mean + c0, mu);
  ld_coh_vec<V>(invstd + c0, is);
  ld_coh_vec<V>(gamma + c0, sc);
  ld_coh_vec<V>(beta + c0, sh);
#pragma unroll
  for (int j = 0; j < V; ++j) {
    sc[j] *= is[j];
    sh[j] -= mu[j] * sc[j];
    a1[j] = 0.f; a2[j] = 0.f;
  }
  if (pl < ppb) {
    for (int op = blockIdx.x * ppb + pl; op < (int)Pout; op += gridDim.x * ppb) {
      int in[4];
      window<POOL>(op, gm, in);
      float d[V];
      load_vec_coh<T>(dout + (size_t)op * C + c0, d);
      float bz[V], by[V];
#pragma unroll
      for (int j = 0; j < V; ++j) { bz[j] = 0.f; by[j] = 0.f; }
#pragma unroll
      for (int q = 0; q < (POOL ? 4 : 1); ++q) {
        float v[V];
        load_vec_coh<T>(y + (size_t)in[q] * C + c0, v);
#pragma unroll
        for (int j = 0; j < V; ++j) {
          float z = fmaf(v[j], sc[j], sh[j]);
          if (z > bz[j]) { bz[j] = z; by[j] = v[j]; }     // strict >: first maximum wins, z<=0 never does
        }
      }
#pragma unroll
      for (int j = 0; j < V; ++j) {
        if (bz[j] > 0.f) {
          a1[j] += d[j];
          a2[j] += d[j] * (by[j] - mu[j]) * is[j];
        }
      }
    }
  }
  if (pdl & 2) pdl_launch_dependents();     // late trigger (see kernel_flags)
  block_reduce_channels<V>(sm, a1, c0, pl, ppb, C, sums);
  block_reduce_channels<V>(sm, a2, c0, pl, ppb, C, sums + C);
  ktrace_end(pdl);
}

// ------------------------------------------------------------------------------------------
// backward pass 2: dy = gamma*invstd * (dz - s1/P - xhat * s2/P); also dgamma, dbeta, conv dbias
// ------------------------------------------------------------------------------------------
template <typename T, bool POOL>
__global__ void __launch_bounds__(256, 2)
bn_bwd_apply_kernel(const T* __restrict__ dout, const T* __restrict__ y, const float* __restrict__ mean,
                    const float* __restrict__ invstd, const float* __restrict__ gamma,
                    const float* __restrict__ beta, const float* __restrict__ sums, T* __restrict__ dy,
                    float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ dbias,
                    long long P, long long Pout, int C, int H, int W, int pdl) {
  constexpr int V = VecOf<T>::N;
  extern __shared__ float sm[];      // [ppb*C]
  const int groups = C / V;
  const int g = threadIdx.x % groups;
  const int pl = threadIdx.x / groups;
  const int ppb = blockDim.x / groups;
  const int c0 = g * V;
  const PoolGeom gm{H, W, log2_exact(W >> 1), log2_exact(H >> 1)};
  if (pdl & 1) pdl_launch_dependents();
  pdl_wait();
  ktrace_begin(pdl);      // stamped once the predecessor has finished: the start of this kernel's own work
  // software pipeline as in the forward kernel: first window before the constants, next one during the current
  constexpr int NQ = POOL ? 4 : 1;
  const int stride = gridDim.x * ppb;
  int op = blockIdx.x * ppb + pl;
  Raw16 nd, nv[NQ];
  if (pl < ppb && op < (int)Pout) {
    int in[4];
    window<POOL>(op, gm, in);
    nd = ld_raw<T>(dout + (size_t)op * C + c0);
#pragma unroll
    for (int q = 0; q < NQ; ++q) nv[q] = ld_raw<T>(y + (size_t)in[q] * C + c0);
  }
  // dy = sc * (dz - k1 - xhat * k2) with xhat = (y - mu) * is, written as  sc * dz + cb - cy * y:
  // four per-channel constants instead of seven keep the pooled variant at two blocks per SM
  float sc[V], sh[V], cb[V], cy[V], ab[V];
  const float invP = 1.f / (float)P;
  ld_coh_vec<V>(gamma + c0, sc);
  ld_coh_vec<V>(beta + c0, sh);
  ld_coh_vec<V>(mean + c0, cb);          // cb / cy / ab double as staging for mean / invstd / sums
  ld_coh_vec<V>(invstd + c0, cy);
  float s1v[V], s2v[V];
  ld_coh_vec<V>(sums + c0, s1v);
  ld_coh_vec<V>(sums + C + c0, s2v);
#pragma unroll
  for (int j = 0; j < V; ++j) {
    const float mu = cb[j], is = cy[j];
    sc[j] *= is;
    sh[j] -= mu * sc[j];
    const float s1 = s1v[j], s2 = s2v[j];
    cy[j] = sc[j] * is * (s2 * invP);
    cb[j] = cy[j] * mu - sc[j] * (s1 * invP);
    ab[j] = 0.f;
    if (blockIdx.x == 0 && pl == 0) { dbeta[c0 + j] = s1; dgamma[c0 + j] = s2; }
  }
  if (pl < ppb) {
    for (; op < (int)Pout; op += stride) {
      int in[4];
      window<POOL>(op, gm, in);
      float d[V];
      unpack_raw<T>(nd, d);
      Raw16 cv[NQ];
#pragma unroll
      for (int q = 0; q < NQ; ++q) cv[q] = nv[q];
      if (op + stride < (int)Pout) {
        int in2[4];
        window<POOL>(op + stride, gm, in2);
        nd = ld_raw<T>(dout + (size_t)(op + stride) * C + c0);
#pragma unroll
        for (int q = 0; q < NQ; ++q) nv[q] = ld_raw<T>(y + (size_t)in2[q] * C + c0);
      }
      float bz[V];
      int bq[V];
#pragma unroll
      for (int j = 0; j < V; ++j) { bz[j] = 0.f; bq[j] = -1; }
#pragma unroll
      for (int q = 0; q < NQ; ++q) {
        float v[V];
        unpack_raw<T>(cv[q], v);
#pragma unroll
        for (int j = 0; j < V; ++j) {
          float z = fmaf(v[j], sc[j], sh[j]);
          if (z > bz[j]) { bz[j] = z; bq[j] = q; }
        }
      }
#pragma unroll
      for (int q = 0; q < NQ; ++q) {
        float v[V], o[V];
        unpack_raw<T>(cv[q], v);
#pragma unroll
        for (int j = 0; j < V; ++j) {
          const float dz = (bq[j] == q) ? d[j] : 0.f;
          o[j] = fmaf(sc[j], dz, fmaf(-cy[j], v[j], cb[j]));
          ab[j] += o[j];
        }
        store_vec<T>(dy + (size_t)in[q] * C + c0, o);
      }
    }
  }
  if (pdl & 2) pdl_launch_dependents();     // late trigger (see kernel_flags): the main loop is done
  if (dbias) block_reduce_channels<V>(sm, ab, c0, pl, ppb, C, dbias);
  ktrace_end(pdl);
}

// ------------------------------------------------------------------------------------------
// standalone per-channel statistics (used when the conv kernel did not produce them:
// SIMT fallback convolutions, unit tests) and standalone 2x2 max-pool fwd/bwd (K-N3, oracle).
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
channel_stats_kernel(const T* __restrict__ y, float* __restrict__ stats, long long P, int C) {
  constexpr int V = VecOf<T>::N;
  extern __shared__ float sm[];      // [ppb*C]
  const int groups = C / V;
  const int g = threadIdx.x % groups, pl = threadIdx.x / groups, ppb = blockDim.x / groups, c0 = g * V;
  float a1[V], a2[V];
#pragma unroll
  for (int j = 0; j < V; ++j) { a1[j] = 0.f; a2[j] = 0.f; }
  if (pl < ppb) {
    for (long long p = (long long)blockIdx.x * ppb + pl; p < P; p += (long long)gridDim.x * ppb) {
      float v[V];
      load_vec<T>(y + p * C + c0, v);
#pragma unroll
      for (int j = 0; j < V; ++j) { a1[j] += v[j]; a2[j] += v[j] * v[j]; }
    }
  }
  block_reduce_channels<V>(sm, a1, c0, pl, ppb, C, stats);
  block_reduce_channels<V>(sm, a2, c0, pl, ppb, C, stats + C);
}

template <typename T>
__global__ void maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ out, long long Pout, int C, int H,
                                   int W) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Pout * C) return;
  int c = i % C;
  int in[4];
  const PoolGeom gm{H, W, log2_exact(W >> 1), log2_exact(H >> 1)};
  window<true>((int)(i / C), gm, in);
  float m = to_f32(x[(size_t)in[0] * C + c]);
#pragma unroll
  for (int q = 1; q < 4; ++q) m = fmaxf(m, to_f32(x[(size_t)in[q] * C + c]));
  out[i] = from_f32<T>(m);
}

template <typename T>
__global__ void maxpool_bwd_kernel(const T* __restrict__ x, const T* __restrict__ dout, T* __restrict__ dx,
                                   long long Pout, int C, int H, int W) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Pout * C) return;
  int c = i % C;
  int in[4];
  const PoolGeom gm{H, W, log2_exact(W >> 1), log2_exact(H >> 1)};
  window<true>((int)(i / C), gm, in);
  float m = to_f32(x[(size_t)in[0] * C + c]);
  int bq = 0;
#pragma unroll
  for (int q = 1; q < 4; ++q) {
    float v = to_f32(x[(size_t)in[q] * C + c]);
    if (v > m) { m = v; bq = q; }
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) dx[(size_t)in[q] * C + c] = (q == bq) ? dout[i] : from_f32<T>(0.f);
}

// Persistent grid: at most one wave of resident blocks (SM count x blocks that fit per SM); the
// kernels stride over pixels, so a tail wave of partially filled SMs never happens.
template <typename K>
static int bn_grid(K kernel, long long Pout, int ppb, int smem_bytes) {
  int occ = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, 256, smem_bytes) != cudaSuccess || occ < 1) {
    (void)cudaGetLastError();
    occ = 2;
  }
  long long g = (Pout + ppb - 1) / ppb;
  const long long cap = (long long)sm_count() * occ;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

// Channel layout of the BN kernels: a thread owns V = 16 bytes of channels (8 bf16 / 4 fp32), the C / V
// channel groups of a pixel sit side by side in a 256-thread block.  Hence C / V must divide 256:
// bf16 C in {8, 16, ..., 2048}, fp32 C in {4, 8, ..., 1024}, C / V a power of two.
template <typename T>
static int bn_check(int C, int H, int W, bool pool) {
  constexpr int V = VecOf<T>::N;
  DDPB_REQUIRE(C > 0 && C % V == 0, "bn: channel count %d is not a multiple of %d (16 bytes of %s)", C, V,
               sizeof(T) == 2 ? "bf16" : "fp32");
  DDPB_REQUIRE(C / V <= 256, "bn: channel count %d exceeds the %s limit of %d (one 256-thread block spans a pixel)", C,
               sizeof(T) == 2 ? "bf16" : "fp32", 256 * V);
  DDPB_REQUIRE(256 % (C / V) == 0, "bn: channel count %d / %d must be a power of two", C, V);
  DDPB_REQUIRE(H > 0 && W > 0, "bn: bad spatial size %dx%d", H, W);
  DDPB_REQUIRE(!pool || (H % 2 == 0 && W % 2 == 0), "bn: pooling needs even H, W (got %dx%d)", H, W);
  return DDPB_OK;
}

}  // namespace ddpb
DDPB_KTRACE_SETTER(bn_relu_pool)

using namespace ddpb;

namespace {

template <typename T, bool POOL>
int fwd_launch(const void* y, const float* stats, const float* gamma, const float* beta, void* out,
               float* save_mean, float* save_invstd, float* rm, float* rv, long long* nbt, void* winner, int batch,
               int H, int W, int C, int training, float eps, float momentum, cudaStream_t stream) {
  int rc = bn_check<T>(C, H, W, POOL);
  if (rc) return rc;
  long long P = (long long)batch * H * W, Pout = POOL ? P / 4 : P;
  int ppb = 256 / (C / VecOf<T>::N);
  DDPB_CUDA_CHECK(launch_ex(bn_fwd_kernel<T, POOL>, dim3(bn_grid(bn_fwd_kernel<T, POOL>, Pout, ppb, 0)), dim3(256), 0,
                            stream, dim3(1, 1, 1), 4, (const T*)y, stats, gamma, beta, (T*)out, save_mean,
                            save_invstd, rm, rv, nbt, (T*)winner, P, Pout, C, H, W, eps, momentum, training,
                            kernel_flags(stream, 4)));
  return DDPB_OK;
}

template <typename T, bool POOL>
int bwd_launch(const void* dout, const void* y, const float* mean, const float* invstd, const float* gamma,
               const float* beta, float* sums, void* dy, float* dgamma, float* dbeta, float* dbias, int batch,
               int H, int W, int C, cudaStream_t stream, bool reduce = true) {
  int rc = bn_check<T>(C, H, W, POOL);
  if (rc) return rc;
  long long P = (long long)batch * H * W, Pout = POOL ? P / 4 : P;
  int ppb = 256 / (C / VecOf<T>::N);
  const int smem = 256 * VecOf<T>::N * sizeof(float);      // [ppb][C] partials
  if (reduce)
    DDPB_CUDA_CHECK(launch_ex(bn_bwd_reduce_kernel<T, POOL>, dim3(bn_grid(bn_bwd_reduce_kernel<T, POOL>, Pout, ppb, smem)),
                              dim3(256), smem, stream, dim3(1, 1, 1), 8, (const T*)dout, (const T*)y, mean, invstd,
                              gamma, beta, sums, Pout, C, H, W, kernel_flags(stream, 8)));
  DDPB_CUDA_CHECK(launch_ex(bn_bwd_apply_kernel<T, POOL>, dim3(bn_grid(bn_bwd_apply_kernel<T, POOL>, Pout, ppb, smem)),
                            dim3(256), smem, stream, dim3(1, 1, 1), 16, (const T*)dout, (const T*)y, mean, invstd,
                            gamma, beta, (const float*)sums, (T*)dy, dgamma, dbeta, dbias, P, Pout, C, H, W,
                            kernel_flags(stream, 16)));
  return DDPB_OK;
}

}  // namespace

// winner (may be null; same shape and dtype as `out`): for every output element the conv output y that
// won the ReLU + max-pool window, NaN where ReLU blocked the whole window (see ddpb_conv3x3_dgrad_bnred_tc).
extern "C" int ddpb_bn_relu_pool_fwd_winner(const void* y, const float* stats, const float* gamma, const float* beta,
                                            void* out, float* save_mean, float* save_invstd, float* running_mean,
                                            float* running_var, long long* num_batches_tracked, void* winner,
                                            int batch, int H, int W, int C, int pool, int training, float eps,
                                            float momentum, int dtype, cudaStream_t stream) {
#define ARGS y, stats, gamma, beta, out, save_mean, save_invstd, running_mean, running_var, \
             num_batches_tracked, winner, batch, H, W, C, training, eps, momentum, stream
  if (dtype == DDPB_BF16) return pool ? fwd_launch<bf16, true>(ARGS) : fwd_launch<bf16, false>(ARGS);
  return pool ? fwd_launch<float, true>(ARGS) : fwd_launch<float, false>(ARGS);
#undef ARGS
}

extern "C" int ddpb_bn_relu_pool_fwd(const void* y, const float* stats, const float* gamma, const float* beta,
                                     void* out, float* save_mean, float* save_invstd, float* running_mean,
                                     float* running_var, long long* num_batches_tracked, int batch, int H,
                                     int W, int C, int pool, int training, float eps, float momentum,
                                     int dtype, cudaStream_t stream) {
  return ddpb_bn_relu_pool_fwd_winner(y, stats, gamma, beta, out, save_mean, save_invstd, running_mean, running_var,
                                      num_batches_tracked, nullptr, batch, H, W, C, pool, training, eps, momentum,
                                      dtype, stream);
}

// `sums` is a zeroed [2*C] scratch; dgamma/dbeta are overwritten, dbias (may be null) accumulated.
extern "C" int ddpb_bn_relu_pool_bwd(const void* dout, const void* y, const float* save_mean,
                                     const float* save_invstd, const float* gamma, const float* beta,
                                     float* sums, void* dy, float* dgamma, float* dbeta, float* dbias,
                                     int batch, int H, int W, int C, int pool, int dtype,
                                     cudaStream_t stream) {
#define ARGS dout, y, save_mean, save_invstd, gamma, beta, sums, dy, dgamma, dbeta, dbias, batch, H, W, C, stream
  if (dtype == DDPB_BF16) return pool ? bwd_launch<bf16, true>(ARGS) : bwd_launch<bf16, false>(ARGS);
  return pool ? bwd_launch<float, true>(ARGS) : bwd_launch<float, false>(ARGS);
#undef ARGS
}

// Second pass only: `sums` already holds sum dz and sum dz * xhat (produced by the epilogue of the
// following layer's data-gradient kernel, ddpb_conv3x3_dgrad_bnred_tc).
extern "C" int ddpb_bn_relu_pool_bwd_apply(const void* dout, const void* y, const float* save_mean,
                                           const float* save_invstd, const float* gamma, const float* beta,
                                           float* sums, void* dy, float* dgamma, float* dbeta, float* dbias,
                                           int batch, int H, int W, int C, int pool, int dtype,
                                           cudaStream_t stream) {
#define ARGS dout, y, save_mean, save_invstd, gamma, beta, sums, dy, dgamma, dbeta, dbias, batch, H, W, C, stream, false
  if (dtype == DDPB_BF16) return pool ? bwd_launch<bf16, true>(ARGS) : bwd_launch<bf16, false>(ARGS);
  return pool ? bwd_launch<float, true>(ARGS) : bwd_launch<float, false>(ARGS);
#undef ARGS
}

// ------------------------------------------------------------------------------------------
// BN backward for callers that have no workspace to give.
//
// ddpb_bn_relu_pool_bwd needs 2*C floats of zeroed scratch for its two per-channel reductions
// and expects the caller to own that buffer.  Some callers cannot: a host program that handed the
// whole GPU to its own memory pool at start-up and forbids further cudaMalloc calls, or a plug-in
// slot whose signature carries only the op's inputs and outputs and no workspace argument.  For
// them the scratch lives inside the library image, as a fixed device array:
//
//  * capacity: DDPB_BN_NOALLOC_MAX_C channels (2 * 2048 floats = 16 KB).  Every GPU that loads the
//    library carries these 16 KB until the process exits, whether or not this entry point is used.
//    The BN kernels themselves add a tighter bound for fp32 tensors: a thread owns 16 bytes of
//    channels and one 256-thread block spans a pixel, so bf16 reaches C = 2048 but fp32 only
//    C = DDPB_BN_NOALLOC_MAX_C_F32 = 1024; in both cases C / (16 bytes' worth) must be a power of
//    two.  Sizes outside these limits are rejected with a message that names the limit, before the
//    shared array is touched;
//  * the array is one per device and shared by all calls: it still holds the previous call's sums,
//    so each call clears the 2*C floats it is about to use (a memset enqueued on `stream`) before
//    the kernels run.  Two calls whose work overlaps on the GPU (different streams, or different
//    host threads) would clear and accumulate into the same floats and corrupt each other:
//    serialise them;
//  * the first call on a device looks up that device's address of the array (the host-side name
//    of a __device__ variable is not a device pointer) and caches it; that look-up is a host-side
//    runtime query, so make the first call on each device before any stream capture begins.
//
// No cudaMalloc / cudaFree and no allocating library routine runs here.  Results are the same as
// ddpb_bn_relu_pool_bwd's up to the order of the floating-point atomic additions.
// ------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_bn_bwd_static_sums[2 * DDPB_BN_NOALLOC_MAX_C];

namespace {
std::mutex g_static_sums_mutex;
float* g_static_sums_addr[64] = {nullptr};     // per device ordinal

// Device address of g_bn_bwd_static_sums on the CURRENT device (looked up once per device).
int static_sums_for_current_device(float** out) {
  int dev = -1;
  DDPB_CUDA_CHECK(cudaGetDevice(&dev));
  DDPB_REQUIRE(dev >= 0 && dev < 64, "bn_bwd_noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_sums_mutex);
  if (!g_static_sums_addr[dev]) {
    float* p = nullptr;
    DDPB_CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_bn_bwd_static_sums));
    DDPB_REQUIRE(p != nullptr, "bn_bwd_noalloc: cudaGetSymbolAddress returned a null pointer");
    g_static_sums_addr[dev] = p;
  }
  *out = g_static_sums_addr[dev];
  return DDPB_OK;
}
}  // namespace

extern "C" int ddpb_bn_relu_pool_bwd_noalloc(const void* dout, const void* y, const float* save_mean,
                                             const float* save_invstd, const float* gamma, const float* beta,
                                             void* dy, float* dgamma, float* dbeta, float* dbias, int batch,
                                             int H, int W, int C, int pool, int dtype, cudaStream_t stream) {
  // Every argument is checked BEFORE the shared array is touched: a rejected call must not clear scratch
  // that an earlier, correctly serialised call may still be using on another stream.
  DDPB_REQUIRE(dtype == DDPB_BF16 || dtype == DDPB_F32, "bn_bwd_noalloc: unknown dtype code %d", dtype);
  DDPB_REQUIRE(C > 0 && C <= DDPB_BN_NOALLOC_MAX_C,
               "bn_bwd_noalloc: C = %d exceeds the built-in scratch capacity of %d channels; use "
               "ddpb_bn_relu_pool_bwd with a caller-owned [2*C] buffer instead",
               C, DDPB_BN_NOALLOC_MAX_C);
  DDPB_REQUIRE(dtype == DDPB_BF16 || C <= DDPB_BN_NOALLOC_MAX_C_F32,
               "bn_bwd_noalloc: C = %d exceeds the fp32 limit of %d channels (the BN kernels give a thread 16 "
               "bytes of channels, so fp32 reaches half as far as bf16's %d)",
               C, DDPB_BN_NOALLOC_MAX_C_F32, DDPB_BN_NOALLOC_MAX_C);
  DDPB_REQUIRE(batch > 0 && dout && y && save_mean && save_invstd && gamma && beta && dy && dgamma && dbeta,
               "bn_bwd_noalloc: null pointer or empty batch");
  int rc = dtype == DDPB_BF16 ? bn_check<bf16>(C, H, W, pool != 0) : bn_check<float>(C, H, W, pool != 0);
  if (rc) return rc;
  float* sums = nullptr;
  rc = static_sums_for_current_device(&sums);
  if (rc) return rc;
  DDPB_CUDA_CHECK(cudaMemsetAsync(sums, 0, sizeof(float) * 2 * (size_t)C, stream));
  return ddpb_bn_relu_pool_bwd(dout, y, save_mean, save_invstd, gamma, beta, sums, dy, dgamma, dbeta, dbias, batch,
                               H, W, C, pool, dtype, stream);
}

// stats[0:C] += sum over pixels, stats[C:2C] += sum of squares (stats must be zeroed by the caller)
extern "C" int ddpb_channel_stats(const void* y, float* stats, long long pixels, int C, int dtype,
                                  cudaStream_t stream) {
  if (dtype == DDPB_BF16) {
    int rc = bn_check<bf16>(C, 2, 2, false);
    if (rc) return rc;
    int ppb = 256 / (C / 8);
    channel_stats_kernel<bf16><<<bn_grid(channel_stats_kernel<bf16>, pixels, ppb, 256 * 8 * 4), 256, 256 * 8 * 4, stream>>>(
        (const bf16*)y, stats, pixels, C);
  } else {
    int rc = bn_check<float>(C, 2, 2, false);
    if (rc) return rc;
    int ppb = 256 / (C / 4);
    channel_stats_kernel<float><<<bn_grid(channel_stats_kernel<float>, pixels, ppb, 256 * 4 * 4), 256, 256 * 4 * 4, stream>>>(
        (const float*)y, stats, pixels, C);
  }
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}

extern "C" int ddpb_maxpool2x2_fwd(const void* x, void* out, int batch, int H, int W, int C, int dtype,
                                   cudaStream_t stream) {
  long long Pout = (long long)batch * (H / 2) * (W / 2);
  int blocks = ceil_div(Pout * C, 256);
  if (dtype == DDPB_BF16)
    maxpool_fwd_kernel<bf16><<<blocks, 256, 0, stream>>>((const bf16*)x, (bf16*)out, Pout, C, H, W);
  else
    maxpool_fwd_kernel<float><<<blocks, 256, 0, stream>>>((const float*)x, (float*)out, Pout, C, H, W);
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}

extern "C" int ddpb_maxpool2x2_bwd(const void* x, const void* dout, void* dx, int batch, int H, int W, int C,
                                   int dtype, cudaStream_t stream) {
  long long Pout = (long long)batch * (H / 2) * (W / 2);
  int blocks = ceil_div(Pout * C, 256);
  if (dtype == DDPB_BF16)
    maxpool_bwd_kernel<bf16><<<blocks, 256, 0, stream>>>((const bf16*)x, (const bf16*)dout, (bf16*)dx, Pout, C, H, W);
  else
    maxpool_bwd_kernel<float><<<blocks, 256, 0, stream>>>((const float*)x, (const float*)dout, (float*)dx, Pout, C, H, W);
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}
