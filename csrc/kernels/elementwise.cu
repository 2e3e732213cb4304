// Memory-bound helpers: data augmentation, fused SGD, casts, layout changes.
//
//  * ddpb_augment         K-D1  uint8 dataset gather + crop(pad 4) + flip + /255 + normalise
//                               (ref Part 1/main.py:82-93 transforms) -> NHWC fp32/bf16
//  * ddpb_sgd_step        K-O3  SGD momentum + weight decay over a flat arena
//                               (ref Part 1/main.py:114-115; torch.optim.SGD semantics)
//  * ddpb_cast_f32_bf16 / ddpb_fill_f32 / ddpb_nchw_to_nhwc
#include "../common.cuh"
#include "../../include/ddp_b200.h"

namespace ddpb {

// ------------------------------------------------------------------------------------------
// K-D1 augmentation. One thread per output pixel (3 channels).
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void augment_kernel(const uint8_t* __restrict__ images,   // [N,32,32,3]
                               const int64_t* __restrict__ index,    // [B] rows of `images`, or null
                               const int32_t* __restrict__ crop_xy,  // [B,2] in [0,8], or null
                               const uint8_t* __restrict__ flip,     // [B], or null
                               T* __restrict__ out,                  // [B,32,32,3]
                               const int64_t* __restrict__ labels_src, int64_t* __restrict__ labels_dst,   // [B] each, or null
                               int batch, float m0, float m1, float m2, float s0, float s1, float s2) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= batch * 1024) return;
  if (labels_dst && t < batch) labels_dst[t] = labels_src[t];       // the batch's targets ride along: one launch less per step
  int b = t >> 10, h = (t >> 5) & 31, w = t & 31;
  long long src = index ? index[b] : b;
  int ox = crop_xy ? crop_xy[2 * b] - 4 : 0;
  int oy = crop_xy ? crop_xy[2 * b + 1] - 4 : 0;
  int ws = (flip && flip[b]) ? 31 - w : w;     // flip happens after the crop
  int sh = h + oy, sw = ws + ox;
  float v0 = 0.f, v1 = 0.f, v2 = 0.f;          // zero padding is applied BEFORE normalisation
  if (sh >= 0 && sh < 32 && sw >= 0 && sw < 32) {
    const uint8_t* p = images + ((src * 32 + sh) * 32 + sw) * 3;
    v0 = p[0] / 255.f; v1 = p[1] / 255.f; v2 = p[2] / 255.f;
  }
  T* o = out + (size_t)t * 3;
  o[0] = from_f32<T>((v0 - m0) / s0);
  o[1] = from_f32<T>((v1 - m1) / s1);
  o[2] = from_f32<T>((v2 - m2) / s2);
}

// ------------------------------------------------------------------------------------------
// K-O3 fused SGD:  g = grad*gscale + wd*p ; buf = mom*buf + g ; p -= lr*buf
// (momentum buffer starts at zero, which reproduces torch's "buf = g" first step exactly).
// Optionally refreshes the bf16 compute copy in the same pass.
// ------------------------------------------------------------------------------------------
__global__ void sgd_kernel(float* __restrict__ param, float* __restrict__ mom,
                           const float* __restrict__ grad, bf16* __restrict__ shadow, long long n4,
                           float lr, float momentum, float wd, float gscale, int flags) {
  ktrace_begin(flags);
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n4; i += stride) {
    float4 p = reinterpret_cast<float4*>(param)[i];
    float4 m = reinterpret_cast<float4*>(mom)[i];
    float4 g = reinterpret_cast<const float4*>(grad)[i];
    g.x = g.x * gscale + wd * p.x; g.y = g.y * gscale + wd * p.y;
    g.z = g.z * gscale + wd * p.z; g.w = g.w * gscale + wd * p.w;
    m.x = momentum * m.x + g.x; m.y = momentum * m.y + g.y;
    m.z = momentum * m.z + g.z; m.w = momentum * m.w + g.w;
    p.x -= lr * m.x; p.y -= lr * m.y; p.z -= lr * m.z; p.w -= lr * m.w;
    reinterpret_cast<float4*>(param)[i] = p;
    reinterpret_cast<float4*>(mom)[i] = m;
    if (shadow) {
      uint2 s;
      s.x = pack_bf16x2(p.x, p.y); s.y = pack_bf16x2(p.z, p.w);
      reinterpret_cast<uint2*>(shadow)[i] = s;
    }
  }
  ktrace_end(flags);
}

__global__ void cast_kernel(const float* __restrict__ src, bf16* __restrict__ dst, long long n4) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n4; i += stride) {
    float4 p = reinterpret_cast<const float4*>(src)[i];
    uint2 s;
    s.x = pack_bf16x2(p.x, p.y); s.y = pack_bf16x2(p.z, p.w);
    reinterpret_cast<uint2*>(dst)[i] = s;
  }
}

__global__ void fill_kernel(float* __restrict__ dst, float v, long long n) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) dst[i] = v;
}

template <typename TI, typename TO>
__global__ void nchw_to_nhwc_kernel(const TI* __restrict__ src, TO* __restrict__ dst, int n, int c, int hw) {
  long long total = (long long)n * c * hw;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  int ch = i % c;
  long long r = i / c;
  int p = r % hw;
  int b = r / hw;
  dst[i] = from_f32<TO>(to_f32(src[((long long)b * c + ch) * hw + p]));
}

// First layer on the tensor cores: its 3x3x3 patches become rows of a [pixels][64] bf16 matrix
// (columns (r*3+s)*3+c, 27 real + 37 zero), zero outside the image = the convolution's padding.
// One thread writes one 16-byte segment, so a warp writes four whole 128-byte rows.  The first
// Cout*8 threads also lay the [Cout][27] bf16 weights out as [Cout][64] (zero padded).
__global__ void __launch_bounds__(256)
im2col3x3_c3_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w, bf16* __restrict__ cols,
                    bf16* __restrict__ w_pad, long long n_seg, int H, int W, int h_shift, int w_shift, int Cout, int pdl) {
  ktrace_begin(pdl);
  if (pdl & 1) pdl_launch_dependents();     // launched normally itself (it follows the scratch memset)
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned short* xs = reinterpret_cast<const unsigned short*>(x);
  if (i < (long long)Cout * 8) {
    const int co = (int)(i >> 3), seg = (int)(i & 7);
    const unsigned short* ws = reinterpret_cast<const unsigned short*>(w) + co * 27;
    unsigned short v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { const int k = seg * 8 + j; v[j] = k < 27 ? ws[k] : (unsigned short)0; }
    uint4 pk = make_uint4(v[0] | (v[1] << 16), v[2] | (v[3] << 16), v[4] | (v[5] << 16), v[6] | (v[7] << 16));
    reinterpret_cast<uint4*>(w_pad)[i] = pk;
  }
  if (i >= n_seg) return;
  const int seg = (int)(i & 7);
  uint4 pk = make_uint4(0u, 0u, 0u, 0u);
  if (seg < 4) {
    // 32-bit index math (B*H*W*8 < 2^31 is checked by the launcher); H, W are powers of two
    const unsigned p = (unsigned)(i >> 3);
    const int wx = (int)(p & (unsigned)(W - 1));
    const int hy = (int)((p >> w_shift) & (unsigned)(H - 1));
    const unsigned img = p >> (w_shift + h_shift);
    unsigned short v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int k = seg * 8 + j;
      const int tap = k / 3, c = k - tap * 3;
      const int r = tap / 3, s2 = tap - r * 3;
      const int yy = hy + r - 1, xx = wx + s2 - 1;
      const bool ok = k < 27 && yy >= 0 && yy < H && xx >= 0 && xx < W;
      v[j] = ok ? xs[((((img << h_shift) + (unsigned)yy) << w_shift) + (unsigned)xx) * 3u + (unsigned)c] : (unsigned short)0;
    }
    pk = make_uint4(v[0] | (v[1] << 16), v[2] | (v[3] << 16), v[4] | (v[5] << 16), v[6] | (v[7] << 16));
  }
  reinterpret_cast<uint4*>(cols)[i] = pk;
  ktrace_end(pdl);
}

// W == 32 fast path: one block = one image row.  The three input rows it needs are staged in shared
// memory with coalesced loads (the generic kernel's eight scattered 2-byte global loads per thread made
// it LSU-bound: 23 us for a 33 MB write), then every thread assembles its 16-byte segment from there.
__global__ void __launch_bounds__(256)
im2col3x3_c3_row32_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w, bf16* __restrict__ cols,
                          bf16* __restrict__ w_pad, int H, int h_shift, int Cout, int zero_pad, int pdl) {
  ktrace_begin(pdl);
  if (pdl & 1) pdl_launch_dependents();     // launched normally itself (it follows the scratch memset)
  __shared__ unsigned short rows[3][34 * 3];       // pixels -1..32 of rows h-1..h+1, zero outside the image
  const unsigned short* xs = reinterpret_cast<const unsigned short*>(x);
  const int tid = threadIdx.x;
  const long long gi = (long long)blockIdx.x * 256 + tid;
  if (gi < (long long)Cout * 8) {
    const int co = (int)(gi >> 3), seg = (int)(gi & 7);
    const unsigned short* ws = reinterpret_cast<const unsigned short*>(w) + co * 27;
    unsigned short v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { const int k = seg * 8 + j; v[j] = k < 27 ? ws[k] : (unsigned short)0; }
    reinterpret_cast<uint4*>(w_pad)[gi] = make_uint4(v[0] | (v[1] << 16), v[2] | (v[3] << 16), v[4] | (v[5] << 16), v[6] | (v[7] << 16));
  }
  const unsigned row = blockIdx.x;                 // (n, h)
  const int h = (int)(row & (unsigned)(H - 1));
  for (int i = tid; i < 3 * 96; i += 256) {
    const int r = i / 96, e = i - r * 96;
    const int yy = h + r - 1;
    rows[r][3 + e] = (yy >= 0 && yy < H) ? xs[((size_t)(row + r - 1) * 32) * 3 + e] : (unsigned short)0;
  }
  if (tid < 18) { const int r = tid / 6, e = tid - r * 6; rows[r][e < 3 ? e : 96 + e] = 0; }
  __syncthreads();
  const int wx = tid >> 3, seg = tid & 7;
  uint4 pk = make_uint4(0u, 0u, 0u, 0u);
  if (seg >= 4 && !zero_pad) return;        // columns 32..63 are zero for good: the caller cleared them once
  if (seg < 4) {
    unsigned short v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int k = seg * 8 + j;
      const int tap = k / 3, c = k - tap * 3;
      const int r = tap / 3, s2 = tap - r * 3;
      v[j] = k < 27 ? rows[r][(wx + s2) * 3 + c] : (unsigned short)0;
    }
    pk = make_uint4(v[0] | (v[1] << 16), v[2] | (v[3] << 16), v[4] | (v[5] << 16), v[6] | (v[7] << 16));
  }
  reinterpret_cast<uint4*>(cols)[(size_t)row * 256 + tid] = pk;
  ktrace_end(pdl);
}

// W == 32, H % 8 == 0, columns 32..63 already zero: one block = EIGHT image rows.  Ten input rows are staged with
// 4-byte loads, then all 256 threads write four 16-byte segments each (the one-row kernel above starts 8192 CTAs of
// which half the threads exit at once; per-CTA staging and the barrier dominated its 15 us).
constexpr int IM2_ROWS = 8;
constexpr int IM2_PITCH = 104;          // 1 pad + pixel -1 (3) + 32 pixels (96) + pixel 32 (3) + 1: rows stay 4-byte aligned
__global__ void __launch_bounds__(256)
im2col3x3_c3_rows8_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w, bf16* __restrict__ cols,
                          bf16* __restrict__ w_pad, int H, int Cout, int pdl) {
  ktrace_begin(pdl);
  if (pdl & 1) pdl_launch_dependents();     // launched normally itself (it follows the scratch memset)
  __shared__ __align__(16) unsigned short rows[IM2_ROWS + 2][IM2_PITCH];
  const int tid = threadIdx.x;
  const long long gi = (long long)blockIdx.x * 256 + tid;
  if (gi < (long long)Cout * 8) {
    const int co = (int)(gi >> 3), seg = (int)(gi & 7);
    const unsigned short* ws = reinterpret_cast<const unsigned short*>(w) + co * 27;
    unsigned short v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { const int k = seg * 8 + j; v[j] = k < 27 ? ws[k] : (unsigned short)0; }
    reinterpret_cast<uint4*>(w_pad)[gi] = make_uint4(v[0] | (v[1] << 16), v[2] | (v[3] << 16), v[4] | (v[5] << 16), v[6] | (v[7] << 16));
  }
  const long long row0 = (long long)blockIdx.x * IM2_ROWS;          // (n, h0): first of this block's image rows
  const int h0 = (int)(row0 & (long long)(H - 1));
  const uint32_t* x32 = reinterpret_cast<const uint32_t*>(x);        // one image row = 96 bf16 = 48 words
  for (int i = tid; i < (IM2_ROWS + 2) * 48; i += 256) {
    const int r = i / 48, e = i - r * 48;
    const int yy = h0 + r - 1;
    const uint32_t v = (yy >= 0 && yy < H) ? x32[(row0 + r - 1) * 48 + e] : 0u;
    *reinterpret_cast<uint32_t*>(&rows[r][4 + 2 * e]) = v;           // pixel p, channel c at [4 + 3 p + c]
  }
  if (tid < (IM2_ROWS + 2) * 6) { const int r = tid / 6, e = tid - r * 6; rows[r][e < 3 ? 1 + e : 100 + (e - 3)] = 0; }   // pixels -1 and 32
  __syncthreads();
#pragma unroll
  for (int it = 0; it < IM2_ROWS * 32 * 4 / 256; ++it) {
    const int item = it * 256 + tid;
    const int seg = item & 3, wx = (item >> 2) & 31, rl = item >> 7;
    unsigned short v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int k = seg * 8 + j;
      const int tap = k / 3, c = k - tap * 3;
      const int r = tap / 3, s2 = tap - r * 3;
      v[j] = k < 27 ? rows[rl + r][1 + (wx + s2) * 3 + c] : (unsigned short)0;
    }
    reinterpret_cast<uint4*>(cols)[((row0 + rl) * 32 + wx) * 8 + seg] =
        make_uint4(v[0] | (v[1] << 16), v[2] | (v[3] << 16), v[4] | (v[5] << 16), v[6] | (v[7] << 16));
  }
  ktrace_end(pdl);
}

static int grid_for(long long n, int block, int cap = 0) {
  if (cap <= 0) cap = sm_count() * 16;
  long long g = (n + block - 1) / block;
  if (g < 1) g = 1;
  return (int)(g > cap ? cap : g);
}

}  // namespace ddpb
DDPB_KTRACE_SETTER(elementwise)

using namespace ddpb;

// labels_src / labels_dst (int64 [batch], nullable): the batch's targets are copied by the same kernel, so a staged
// batch reaches the engine's input and target buffers in ONE launch (ddpb_augment = without labels).
extern "C" int ddpb_augment_labels(const void* images, const void* index, const void* crop_xy, const void* flip,
                                   void* out, const void* labels_src, void* labels_dst, int batch, int out_dtype,
                                   const float* mean3, const float* std3, cudaStream_t stream) {
  DDPB_REQUIRE(batch > 0, "augment: batch must be positive");
  DDPB_REQUIRE((labels_src == nullptr) == (labels_dst == nullptr), "augment: labels_src and labels_dst go together");
  int threads = 256, blocks = ceil_div((long long)batch * 1024, threads);
  if (out_dtype == DDPB_BF16) {
    augment_kernel<bf16><<<blocks, threads, 0, stream>>>(
        (const uint8_t*)images, (const int64_t*)index, (const int32_t*)crop_xy, (const uint8_t*)flip,
        (bf16*)out, (const int64_t*)labels_src, (int64_t*)labels_dst, batch, mean3[0], mean3[1], mean3[2], std3[0],
        std3[1], std3[2]);
  } else {
    augment_kernel<float><<<blocks, threads, 0, stream>>>(
        (const uint8_t*)images, (const int64_t*)index, (const int32_t*)crop_xy, (const uint8_t*)flip,
        (float*)out, (const int64_t*)labels_src, (int64_t*)labels_dst, batch, mean3[0], mean3[1], mean3[2], std3[0],
        std3[1], std3[2]);
  }
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}

extern "C" int ddpb_augment(const void* images, const void* index, const void* crop_xy, const void* flip,
                            void* out, int batch, int out_dtype, const float* mean3, const float* std3,
                            cudaStream_t stream) {
  return ddpb_augment_labels(images, index, crop_xy, flip, out, nullptr, nullptr, batch, out_dtype, mean3, std3, stream);
}

extern "C" int ddpb_sgd_step(float* param, float* mom, const float* grad, void* shadow_bf16, long long n,
                             float lr, float momentum, float wd, float grad_scale, cudaStream_t stream) {
  DDPB_REQUIRE(n % 4 == 0, "sgd: element count must be a multiple of 4 (arena slots are 64-aligned)");
  long long n4 = n / 4;
  sgd_kernel<<<grid_for(n4, 256), 256, 0, stream>>>(param, mom, grad, (bf16*)shadow_bf16, n4, lr, momentum,
                                                    wd, grad_scale, kernel_flags(stream, 0));
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}

extern "C" int ddpb_cast_f32_bf16(const float* src, void* dst, long long n, cudaStream_t stream) {
  DDPB_REQUIRE(n % 4 == 0, "cast: element count must be a multiple of 4");
  cast_kernel<<<grid_for(n / 4, 256), 256, 0, stream>>>(src, (bf16*)dst, n / 4);
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}

extern "C" int ddpb_fill_f32(float* dst, float value, long long n, cudaStream_t stream) {
  fill_kernel<<<grid_for(n, 256), 256, 0, stream>>>(dst, value, n);
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}

extern "C" int ddpb_nchw_to_nhwc(const void* src, void* dst, int n, int c, int hw, int src_dtype,
                                 int dst_dtype, cudaStream_t stream) {
  long long total = (long long)n * c * hw;
  int blocks = ceil_div(total, 256);
  if (src_dtype == DDPB_F32 && dst_dtype == DDPB_F32)
    nchw_to_nhwc_kernel<float, float><<<blocks, 256, 0, stream>>>((const float*)src, (float*)dst, n, c, hw);
  else if (src_dtype == DDPB_F32 && dst_dtype == DDPB_BF16)
    nchw_to_nhwc_kernel<float, bf16><<<blocks, 256, 0, stream>>>((const float*)src, (bf16*)dst, n, c, hw);
  else if (src_dtype == DDPB_BF16 && dst_dtype == DDPB_BF16)
    nchw_to_nhwc_kernel<bf16, bf16><<<blocks, 256, 0, stream>>>((const bf16*)src, (bf16*)dst, n, c, hw);
  else
    DDPB_REQUIRE(false, "nchw_to_nhwc: unsupported dtype pair");
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}

// x bf16 [B,H,W,3] -> cols bf16 [B*H*W][64] (3x3 patches, zero padded), w bf16 [Cout][27] -> w_pad [Cout][64].
// zero_pad = 0: columns 32..63 of `cols` are not touched (the caller zeroed them once; they never change).
extern "C" int ddpb_im2col3x3_c3(const void* x, const void* w_bf16, void* cols, void* w_pad, int B, int H, int W,
                                 int Cout, int zero_pad, cudaStream_t stream) {
  DDPB_REQUIRE(B > 0 && H > 0 && W > 0 && Cout > 0 && (H & (H - 1)) == 0 && (W & (W - 1)) == 0,
               "im2col3x3_c3: H and W must be powers of two");
  const long long n_seg = (long long)B * H * W * 8;
  DDPB_REQUIRE(n_seg < (1LL << 31), "im2col3x3_c3: batch too large");
  int h_shift = 0, w_shift = 0;
  while ((1 << h_shift) < H) ++h_shift;
  while ((1 << w_shift) < W) ++w_shift;
  DDPB_REQUIRE(n_seg >= (long long)Cout * 8, "im2col3x3_c3: fewer pixels than output channels");
  const long long blocks = (n_seg + 255) / 256;
  static const bool rows8 = !(getenv("DDPB_IM2COL_ROWS8") && atoi(getenv("DDPB_IM2COL_ROWS8")) == 0);
  if (W == 32 && !zero_pad && rows8 && H % IM2_ROWS == 0 && (long long)B * H / IM2_ROWS * 256 >= (long long)Cout * 8) {
    im2col3x3_c3_rows8_kernel<<<(unsigned)((long long)B * H / IM2_ROWS), 256, 0, stream>>>(
        (const bf16*)x, (const bf16*)w_bf16, (bf16*)cols, (bf16*)w_pad, H, Cout, kernel_flags(stream, 64));
    DDPB_LAUNCH_CHECK();
    return DDPB_OK;
  }
  if (W == 32) {
    im2col3x3_c3_row32_kernel<<<(unsigned)((long long)B * H), 256, 0, stream>>>((const bf16*)x, (const bf16*)w_bf16, (bf16*)cols,
                                                                               (bf16*)w_pad, H, h_shift, Cout, zero_pad, kernel_flags(stream, 64));
    DDPB_LAUNCH_CHECK();
    return DDPB_OK;
  }
  im2col3x3_c3_kernel<<<(unsigned)blocks, 256, 0, stream>>>((const bf16*)x, (const bf16*)w_bf16, (bf16*)cols,
                                                            (bf16*)w_pad, n_seg, H, W, h_shift, w_shift, Cout, kernel_flags(stream, 64));
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}
