// Classifier (Linear 512 -> 10) fused with softmax cross-entropy, forward and backward.
//
// Replaces ``self.fc1`` + ``CrossEntropyLoss()`` (ref Part 1/model.py:40,45; Part 1/main.py:39,110;
// SURVEY K-O1/K-O2).  N = 10 is far too narrow for a tensor-core tile, and the whole problem is
// B x 512 x 10, so this is a warp-per-sample SIMT kernel:
//   forward : logits = a W^T + b ; loss += -log softmax[target] / B ; dlogits = (softmax - onehot)/B ;
//             correct += (argmax == target)          (the eval loop's accuracy count, Part 1/main.py:70-71)
//   backward: dA = dlogits W ; dW = dlogits^T A ; db = sum_b dlogits
#include "../common.cuh"
#include "../../include/ddp_b200.h"

namespace ddpb {

constexpr int NCLS = 10;
constexpr int FEAT = 512;
constexpr int FC_SLICE = 16;

// One warp per sample: lane l owns features [16 l, 16 l + 16) (two 16-byte activation loads).  The 20 KB weight
// matrix is staged in shared memory by the whole CTA BEFORE griddepcontrol.wait - it does not depend on the
// preceding kernel, so the staging overlaps that kernel's tail - in a lane-interleaved layout (float4 index
// (class * 4 + quarter) * 32 + lane: conflict-free 16-byte reads).  Round 1 read the weights with ten dependent
// groups of global loads per lane (42 registers: the compiler serialised them): 8.1 us, of which ~5 us were those
// round trips.
template <typename T>
__global__ void __launch_bounds__(128)
fc_ce_fwd_kernel(const T* __restrict__ act, const float* __restrict__ w, const float* __restrict__ bias,
                 const int64_t* __restrict__ target, float* __restrict__ logits, float* __restrict__ dlogits,
                 float* __restrict__ loss, int* __restrict__ correct, T* __restrict__ dact, int batch, float inv_batch,
                 int pdl) {
  __shared__ float4 ws[NCLS * 4 * 32];
  __shared__ float bs[NCLS];
  if (pdl & 1) pdl_launch_dependents();     // programmatic dependent launch: see common.cuh
  for (int e = threadIdx.x; e < NCLS * FEAT / 4; e += blockDim.x) {
    const int j = e / (FEAT / 4), rem = e - j * (FEAT / 4);
    ws[(j * 4 + (rem & 3)) * 32 + (rem >> 2)] = DDPB_LD_COH(reinterpret_cast<const float4*>(w) + e);
  }
  if (threadIdx.x < NCLS) bs[threadIdx.x] = ld_coh(&bias[threadIdx.x]);
  pdl_wait();
  ktrace_begin(pdl);      // stamped once the predecessor has finished: the start of this kernel's own work
  __syncthreads();
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int b = blockIdx.x * (blockDim.x >> 5) + warp;
  if (b >= batch) return;
  constexpr int V = Vec16<T>::N;
  float av[16];
#pragma unroll
  for (int i = 0; i < 16; i += V) load_vec_coh<T>(act + (size_t)b * FEAT + lane * 16 + i, av + i);
  float acc[NCLS];
#pragma unroll
  for (int j = 0; j < NCLS; ++j) {
    float a = 0.f;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 wv = ws[(j * 4 + q) * 32 + lane];
      a = fmaf(av[4 * q], wv.x, a); a = fmaf(av[4 * q + 1], wv.y, a);
      a = fmaf(av[4 * q + 2], wv.z, a); a = fmaf(av[4 * q + 3], wv.w, a);
    }
    acc[j] = a;
  }
#pragma unroll
  for (int j = 0; j < NCLS; ++j) acc[j] = warp_sum(acc[j]) + bs[j];
  // every lane now holds all ten logits
  float mx = acc[0];
  int arg = 0;
#pragma unroll
  for (int j = 1; j < NCLS; ++j)
    if (acc[j] > mx) { mx = acc[j]; arg = j; }
  float se = 0.f;
#pragma unroll
  for (int j = 0; j < NCLS; ++j) se += expf(acc[j] - mx);
  float lse = mx + logf(se);
  int t = (int)DDPB_LD_COH(reinterpret_cast<const long long*>(&target[b]));
  float dl[NCLS];                 // every lane: the sample's logit gradient (identical values in all lanes)
#pragma unroll
  for (int j = 0; j < NCLS; ++j) dl[j] = (expf(acc[j] - lse) - (j == t ? 1.f : 0.f)) * inv_batch;
  if (lane == 0) {
    float lt = 0.f;
#pragma unroll
    for (int j = 0; j < NCLS; ++j) {
      if (logits) logits[b * NCLS + j] = acc[j];
      if (dlogits) dlogits[b * NCLS + j] = dl[j];
      if (j == t) lt = acc[j];
    }
    atomicAdd(loss, (lse - lt) * inv_batch);
    if (correct && arg == t) atomicAdd(correct, 1);
  }
  if (dact) {
    // the gradient of the classifier's input, dA[b, k] = sum_j dlogits[b, j] W[j, k], from the staged weights: the
    // backward pass then starts with the last block's BatchNorm backward and the classifier's weight gradient
    // leaves the critical path (same summation order as fc_bwd_kernel: bit-identical)
    float o[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) o[i] = 0.f;
#pragma unroll
    for (int j = 0; j < NCLS; ++j)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 wv = ws[(j * 4 + q) * 32 + lane];
        o[4 * q] = fmaf(dl[j], wv.x, o[4 * q]); o[4 * q + 1] = fmaf(dl[j], wv.y, o[4 * q + 1]);
        o[4 * q + 2] = fmaf(dl[j], wv.z, o[4 * q + 2]); o[4 * q + 3] = fmaf(dl[j], wv.w, o[4 * q + 3]);
      }
#pragma unroll
    for (int i = 0; i < 16; i += V) store_vec<T>(dact + (size_t)b * FEAT + lane * 16 + i, o + i);
  }
  ktrace_end(pdl);
}

// blocks [0, batch*FEAT/256): dA ; remaining blocks: dW and db (one thread per (class, feature))
template <typename T>
__global__ void __launch_bounds__(256)
fc_bwd_kernel(const T* __restrict__ act, const float* __restrict__ w, const float* __restrict__ dlogits,
              T* __restrict__ dact, float* __restrict__ dw, float* __restrict__ db, int batch, int blocks_da, int pdl) {
  ktrace_begin(pdl);
  if (pdl & 1) pdl_launch_dependents();     // launched normally itself (it follows the gradient memset)
  if ((int)blockIdx.x < blocks_da) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= batch * FEAT) return;
    int b = i / FEAT, k = i % FEAT;
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < NCLS; ++j) s = fmaf(dlogits[b * NCLS + j], w[j * FEAT + k], s);
    dact[i] = from_f32<T>(s);
  } else {
    // dW/db: the batch is cut into slices of FC_SLICE samples, partial sums are accumulated
    // atomically into the (pre-zeroed) gradient arena.
    int rel = blockIdx.x - blocks_da;
    int per = (NCLS * FEAT + NCLS + 255) / 256;
    int slice = rel / per;
    int i = (rel % per) * blockDim.x + threadIdx.x;
    int b0 = slice * FC_SLICE, b1 = min(batch, b0 + FC_SLICE);
    if (i < NCLS * FEAT) {
      int j = i / FEAT, k = i % FEAT;
      float s = 0.f;
      for (int b = b0; b < b1; ++b) s = fmaf(dlogits[b * NCLS + j], to_f32(act[(size_t)b * FEAT + k]), s);
      atomicAdd(&dw[i], s);
    } else if (i < NCLS * FEAT + NCLS) {
      int j = i - NCLS * FEAT;
      float s = 0.f;
      for (int b = b0; b < b1; ++b) s += dlogits[b * NCLS + j];
      atomicAdd(&db[j], s);
    }
  }
  ktrace_end(pdl);
}

}  // namespace ddpb
DDPB_KTRACE_SETTER(fc_ce)

using namespace ddpb;

// `loss` (1 float) and `correct` (1 int, may be null) are ACCUMULATED into: zero them first.
// dact (nullable, same dtype as act, [batch, 512]): also write the gradient of the classifier's input, which
// ddpb_fc_bwd would otherwise compute (call that with dact = null then: weight and bias gradients only).
extern "C" int ddpb_fc_ce_fwd_dact(const void* act, const float* w, const float* bias, const void* target_i64,
                                   float* logits, float* dlogits, float* loss, int* correct, void* dact, int batch,
                                   int loss_divisor, int dtype, cudaStream_t stream) {
  DDPB_REQUIRE(batch > 0 && loss_divisor > 0, "fc_ce_fwd: bad batch");
  int blocks = ceil_div(batch, 4);
  float inv = 1.f / (float)loss_divisor;
  if (dtype == DDPB_BF16)
    DDPB_CUDA_CHECK(launch_ex(fc_ce_fwd_kernel<bf16>, dim3(blocks), dim3(128), 0, stream, dim3(1, 1, 1), 32,
                              (const bf16*)act, w, bias, (const int64_t*)target_i64, logits, dlogits, loss, correct,
                              (bf16*)dact, batch, inv, kernel_flags(stream, 32)));
  else
    DDPB_CUDA_CHECK(launch_ex(fc_ce_fwd_kernel<float>, dim3(blocks), dim3(128), 0, stream, dim3(1, 1, 1), 32,
                              (const float*)act, w, bias, (const int64_t*)target_i64, logits, dlogits, loss, correct,
                              (float*)dact, batch, inv, kernel_flags(stream, 32)));
  return DDPB_OK;
}

extern "C" int ddpb_fc_ce_fwd(const void* act, const float* w, const float* bias, const void* target_i64,
                              float* logits, float* dlogits, float* loss, int* correct, int batch,
                              int loss_divisor, int dtype, cudaStream_t stream) {
  return ddpb_fc_ce_fwd_dact(act, w, bias, target_i64, logits, dlogits, loss, correct, nullptr, batch, loss_divisor, dtype,
                             stream);
}

// dw [10*512] and db [10] are ACCUMULATED into (they live in the zeroed gradient arena).
extern "C" int ddpb_fc_bwd(const void* act, const float* w, const float* dlogits, void* dact, float* dw,
                           float* db, int batch, int dtype, cudaStream_t stream) {
  int blocks_da = dact ? ceil_div((long long)batch * FEAT, 256) : 0;     // dact = null: weight / bias gradients only
  int blocks_dw = ceil_div(NCLS * FEAT + NCLS, 256) * ceil_div(batch, FC_SLICE);
  if (dtype == DDPB_BF16)
    fc_bwd_kernel<bf16><<<blocks_da + blocks_dw, 256, 0, stream>>>((const bf16*)act, w, dlogits, (bf16*)dact, dw,
                                                                  db, batch, blocks_da, kernel_flags(stream, 64));
  else
    fc_bwd_kernel<float><<<blocks_da + blocks_dw, 256, 0, stream>>>((const float*)act, w, dlogits, (float*)dact,
                                                                   dw, db, batch, blocks_da, kernel_flags(stream, 64));
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}
