// Shared device/host helpers for the sm_100a kernels: error handling, bf16 packing,
// mbarrier / TMA / tcgen05 PTX wrappers and shared-memory (UMMA) descriptors.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#define DDPB_OK 0
#define DDPB_ERR_CUDA 1
#define DDPB_ERR_ARG 2
#define DDPB_ERR_UNSUPPORTED 3

namespace ddpb {

void set_last_error(const char* fmt, ...);

#define DDPB_CUDA_CHECK(expr)                                                              \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess) {                                                               \
      ::ddpb::set_last_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,                 \
                             cudaGetErrorString(_e));                                      \
      (void)cudaGetLastError(); /* do not leave a non-sticky error behind for the next call */ \
      return DDPB_ERR_CUDA;                                                                \
    }                                                                                      \
  } while (0)

#define DDPB_LAUNCH_CHECK() DDPB_CUDA_CHECK(cudaGetLastError())

#define DDPB_REQUIRE(cond, ...)                                                            \
  do {                                                                                     \
    if (!(cond)) {                                                                         \
      ::ddpb::set_last_error(__VA_ARGS__);                                                 \
      return DDPB_ERR_ARG;                                                                 \
    }                                                                                      \
  } while (0)

static inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

// Number of SMs of the CURRENT device (queried once per device; 148 on B200) - grids are sized from this,
// never from a literal.
int sm_count();
// One-time per-DEVICE setup (cudaFuncSetAttribute is per device, a process may drive several):
//   static unsigned long long done = 0;  if (first_on_device(&done)) { ...set attributes... }
bool first_on_device(unsigned long long* mask);

// Programmatic dependent launch: a kernel launched through launch_ex(..., pdl = true) may be scheduled
// while its predecessor on the stream is still running, so its prologue (barrier init, TMEM
// allocation, descriptor prefetch, index math) and the launch latency overlap the predecessor's tail
// (applied while a CUDA graph is being captured only, see launch_ex).
// Such a kernel MUST execute pdl_wait() before its first access to global memory; while capturing, every
// kernel calls pdl_launch_dependents() early so that ITS successor can be scheduled.  DDPB_PDL=0 disables it.
bool pdl_enabled();
int pdl_mask();         // DDPB_PDL_MASK: bit per kernel family (debugging aid), default all ones

// Is programmatic dependent launch in effect for a kernel of `family` launched on `stream` right now?
// Only inside a stream capture (the product path), where every dependency - the events between the
// engine's streams included - is an explicit graph edge.  Eager runs keep ordinary launches and do not
// execute the trigger either: kernels take the answer as a parameter (`if (pdl) pdl_launch_dependents()`).
static inline int pdl_active(cudaStream_t stream, int family) {
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  return family && pdl_enabled() && (pdl_mask() & family) && cudaStreamIsCapturing(stream, &cap) == cudaSuccess &&
         cap == cudaStreamCaptureStatusActive;
}

// ---- in-kernel step tracer --------------------------------------------------------------------
// Every kernel of the training step takes an int of launch flags: bits 0-1 = where to execute
// griddepcontrol.launch_dependents (see kernel_flags), bits 2.. = trace id (0: not traced).  A traced kernel stamps
// %globaltimer into slot 2*id (earliest CTA start) and 2*id+1 (latest warp end) of the trace buffer,
// which gives the step's real timeline INSIDE the production CUDA graph, programmatic launches
// included (ddpb_engine_trace_step, mode 2).  The buffer pointer lives in a per-translation-unit
// __device__ variable (no relocatable device code here), set through DDPB_KTRACE_SETTER's function.
// kernel_flags: bit 0 = trigger the successor at the START of every CTA (it becomes resident at once and waits
// in griddepcontrol.wait), bit 1 = trigger it LATE, when the CTA's main loop is done.  An early-triggered
// successor parks its CTAs on the SMs for the whole duration of this kernel; that is free while the step has
// the GPU to itself, but in the backward pass those parked CTAs hold the registers a bucket kernel of the
// communication stream (or a weight-gradient CTA) needs, and the bucket kernel then starts one layer late
// (seen with the step tracer).  DDPB_PDL_LATE is the mask of kernel families that trigger late.
#ifndef DDPB_PDL_LATE_DEFAULT
#define DDPB_PDL_LATE_DEFAULT 0
#endif
int pdl_late_mask();
// (late_family: the bit DDPB_PDL_LATE tests when it differs from the DDPB_PDL_MASK family; 128 = dgrad)
int kernel_flags(cudaStream_t stream, int family, int late_family = 0);     // early | late << 1 | next trace id << 2
void ktrace_enable(int on);                             // host: start / stop handing out trace ids (restarts at 1)
int ktrace_last_id();                                   // host: the last id handed out

template <typename... KArgs, typename... Args>
static inline cudaError_t launch_ex(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                    dim3 cluster, int pdl, Args... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int n = 0;
  if (cluster.x * cluster.y * cluster.z > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = cluster.x; attr[n].val.clusterDim.y = cluster.y; attr[n].val.clusterDim.z = cluster.z;
    ++n;
  }
  if (pdl_active(stream, pdl)) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = attr; cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

typedef __nv_bfloat16 bf16;

// cache operator of the "coherent" loads (see load_vec_coh): ld.global.ca; ld.global.cg was measured slower and no more correct
#ifndef DDPB_LD_COH
#define DDPB_LD_COH __ldca
#endif

// ------------------------------------------------------------------------------------------
// scalar / vector conversion helpers
// ------------------------------------------------------------------------------------------
template <typename T> struct Vec16;   // 16-byte vector of T
template <> struct Vec16<float> { static constexpr int N = 4; };
template <> struct Vec16<bf16> { static constexpr int N = 8; };

__device__ __forceinline__ float to_f32(float v) { return v; }
__device__ __forceinline__ float to_f32(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f32<bf16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}

// Load/store VEC contiguous elements (16 bytes) as floats.
template <typename T> __device__ __forceinline__ void load_vec(const T* p, float* out);
template <> __device__ __forceinline__ void load_vec<float>(const float* p, float* out) {
  float4 v = *reinterpret_cast<const float4*>(p);
  out[0] = v.x; out[1] = v.y; out[2] = v.z; out[3] = v.w;
}
template <> __device__ __forceinline__ void load_vec<bf16>(const bf16* p, float* out) {
  uint4 v = *reinterpret_cast<const uint4*>(p);
  float2 a = unpack_bf16x2(v.x), b = unpack_bf16x2(v.y), c = unpack_bf16x2(v.z), d = unpack_bf16x2(v.w);
  out[0] = a.x; out[1] = a.y; out[2] = b.x; out[3] = b.y;
  out[4] = c.x; out[5] = c.y; out[6] = d.x; out[7] = d.y;
}
// Coherent variants (ld.global.ca, never ld.global.nc).  Kernels launched as programmatic dependents
// must use these for everything an earlier kernel wrote: their lifetime overlaps the producer's, so the
// read-only path (LDG.CONSTANT) the compiler picks for const __restrict__ pointers returns stale lines
// (measured: BatchNorm statistics of the PREVIOUS step).  ld.global.cg would also be correct but
// compiles to LDG.STRONG.GPU, which made the BatchNorm kernels ~2x slower.
template <typename T> __device__ __forceinline__ void load_vec_coh(const T* p, float* out);
template <> __device__ __forceinline__ void load_vec_coh<float>(const float* p, float* out) {
  float4 v = DDPB_LD_COH(reinterpret_cast<const float4*>(p));
  out[0] = v.x; out[1] = v.y; out[2] = v.z; out[3] = v.w;
}
template <> __device__ __forceinline__ void load_vec_coh<bf16>(const bf16* p, float* out) {
  uint4 v = DDPB_LD_COH(reinterpret_cast<const uint4*>(p));
  float2 a = unpack_bf16x2(v.x), b = unpack_bf16x2(v.y), c = unpack_bf16x2(v.z), d = unpack_bf16x2(v.w);
  out[0] = a.x; out[1] = a.y; out[2] = b.x; out[3] = b.y;
  out[4] = c.x; out[5] = c.y; out[6] = d.x; out[7] = d.y;
}
__device__ __forceinline__ float ld_coh(const float* p) { return DDPB_LD_COH(p); }
// N (multiple of 4) consecutive floats with 16-byte loads: per-channel parameters of a thread's channel
// group (scalar loads made the BatchNorm kernels load-instruction bound: 32-48 of them per thread)
template <int N>
__device__ __forceinline__ void ld_coh_vec(const float* p, float* out) {
#pragma unroll
  for (int i = 0; i < N; i += 4) {
    const float4 v = DDPB_LD_COH(reinterpret_cast<const float4*>(p + i));
    out[i] = v.x; out[i + 1] = v.y; out[i + 2] = v.z; out[i + 3] = v.w;
  }
}
__device__ __forceinline__ float ld_coh_any(const float* p) { return DDPB_LD_COH(p); }
__device__ __forceinline__ bf16 ld_coh_any(const bf16* p) {
  unsigned short u = DDPB_LD_COH(reinterpret_cast<const unsigned short*>(p));
  return *reinterpret_cast<bf16*>(&u);
}
template <typename T> __device__ __forceinline__ void store_vec(T* p, const float* in);
template <> __device__ __forceinline__ void store_vec<float>(float* p, const float* in) {
  *reinterpret_cast<float4*>(p) = make_float4(in[0], in[1], in[2], in[3]);
}
template <> __device__ __forceinline__ void store_vec<bf16>(bf16* p, const float* in) {
  uint4 v;
  v.x = pack_bf16x2(in[0], in[1]); v.y = pack_bf16x2(in[2], in[3]);
  v.z = pack_bf16x2(in[4], in[5]); v.w = pack_bf16x2(in[6], in[7]);
  *reinterpret_cast<uint4*>(p) = v;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ------------------------------------------------------------------------------------------
// sm_100a PTX wrappers (device only)
// ------------------------------------------------------------------------------------------
#ifdef __CUDACC__
static __device__ unsigned long long* g_ktrace;
#define DDPB_KTRACE_SETTER(tu)                                                                   \
  extern "C" int ddpb_ktrace_set_##tu(void* p) {                                                 \
    return cudaMemcpyToSymbol(::ddpb::g_ktrace, &p, sizeof(p)) == cudaSuccess ? 0 : 1;                  \
  }
__device__ __forceinline__ unsigned long long ktrace_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// Slot layout: [id][32 lanes][begin, end]; a CTA uses lane (linear block index % 32) so that a grid of
// thousands of CTAs does not serialise on one address.  One stamp per CTA, taken by its thread 0 (in every
// kernel here thread 0 leaves last or together with the rest: final barriers, uniform loops).
constexpr int KTRACE_LANES = 32;
__device__ __forceinline__ unsigned long long* ktrace_slot(int id) {
  const unsigned b = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
  return g_ktrace + ((size_t)id * KTRACE_LANES + (b % KTRACE_LANES)) * 2;
}
__device__ __forceinline__ void ktrace_begin(int flags) {
  const int id = flags >> 2;
  if (id && threadIdx.x == 0 && threadIdx.y == 0 && threadIdx.z == 0) atomicMin(ktrace_slot(id), ktrace_now());
}
__device__ __forceinline__ void ktrace_end(int flags) {      // call where thread 0 leaves the kernel
  const int id = flags >> 2;
  if (id && threadIdx.x == 0 && threadIdx.y == 0 && threadIdx.z == 0) atomicMax(ktrace_slot(id) + 1, ktrace_now());
}
#endif

__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .pred P;\n\t.reg .b32 R;\n\t"
      "elect.sync R|P, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, P;\n\t}\n"
      : "=r"(pred));
  return pred != 0;
}

// ---- mbarrier -----------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_async_proxy() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, P;\n\t}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// Bounded wait (2 s of wall clock): a protocol bug becomes a trap (launch failure) instead of a
// hung GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const unsigned long long t0 = global_timer_ns();
  while (!mbar_try_wait(bar, parity)) {
    if (global_timer_ns() - t0 > 2000000000ull) {
      printf("ddpb: mbarrier wait timed out (block %d,%d,%d thread %d)\n", blockIdx.x, blockIdx.y,
             blockIdx.z, threadIdx.x);
      __trap();
    }
  }
}

// ---- TMA -----------------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem, const CUtensorMap* m, uint64_t* bar, int c0,
                                            int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(void* smem, const CUtensorMap* m, uint64_t* bar, int c0,
                                            int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1),
      "r"(c2), "r"(c3)
      : "memory");
}

// multicast variant: the box lands at the same shared-memory offset in every CTA of `mask`, and each
// destination CTA's mbarrier (same offset) is credited with the bytes written into ITS memory
__device__ __forceinline__ void tma_load_4d_mc(void* smem, const CUtensorMap* m, uint64_t* bar, int c0, int c1,
                                               int c2, int c3, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], %7;"
      ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2),
      "r"(c3), "h"(mask)
      : "memory");
}
__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_barrier() {      // all threads of all CTAs of the cluster
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, const void* smem, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(smem_u32(smem)), "r"(c0), "r"(c1)
               : "memory");
}
// global[tile] += smem[tile] (fp32 add performed by the memory system, whole 128-byte lines)
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* m, const void* smem, int c0, int c1) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(smem_u32(smem)), "r"(c0), "r"(c1)
               : "memory");
}
// global[0 .. bytes) += smem[0 .. bytes) as ONE bulk operation (fp32 adds performed in L2); both addresses 16-byte
// aligned, bytes a multiple of 16.  Replaces bytes / 4 separate RED instructions.
__device__ __forceinline__ void bulk_reduce_add_f32(float* gmem, const void* smem, uint32_t bytes) {
  asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(gmem), "r"(smem_u32(smem)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// wait until the bulk stores / reductions of this thread are complete (not only their shared-memory reads)
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// wait until the bulk stores of this thread have finished READING shared memory
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }

// ---- tcgen05 / TMEM -------------------------------------------------------------------------
template <int COLS>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot) {   // whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)),
               "n"(COLS));
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {      // whole warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS));
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem]; bf16 inputs, fp32 accumulate; issued by ONE thread.
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]; fp32 storage read as TF32 (10-bit mantissa), fp32 accumulate; K = 8 per instruction.
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// mbarrier arrive once all previously issued MMAs of this thread have completed.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// same, arriving on the barrier at this offset in every CTA of `mask` (stages shared through multicast)
__device__ __forceinline__ void umma_commit_mc(uint64_t* bar, uint16_t mask) {
  asm volatile(
      "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(smem_u32(bar)), "h"(mask)
      : "memory");
}
// ---- CTA pairs (cta_group::2): two CTAs on the two SMs of a TPC form one MMA of M = 256 --------------------
// shared::cluster address of the LEADER CTA's copy of a barrier: pair = cluster ranks (2k, 2k+1), the rank sits
// above bit 24 of a shared::cluster address, so clearing bit 24 selects the even rank
__device__ __forceinline__ uint32_t pair_leader_bar(const uint64_t* bar) { return smem_u32(bar) & 0xFEFFFFFFu; }
// TMA loads whose bytes are accounted on the leader's barrier (each CTA loads into ITS OWN shared memory)
__device__ __forceinline__ void tma2_load_2d(void* smem, const CUtensorMap* m, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma2_load_4d(void* smem, const CUtensorMap* m, uint32_t bar, int c0, int c1, int c2,
                                             int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem2_alloc(uint32_t* smem_slot) {   // one warp in EACH CTA of the pair
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "n"(COLS));
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
}
template <int COLS>
__device__ __forceinline__ void tmem2_dealloc(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS));
}
// D[tmem of both CTAs] (+)= A[256 rows: 128 from each CTA] * B[N columns: N/2 from each CTA]; leader thread only
__device__ __forceinline__ void umma2_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at this shared-memory offset in the CTAs of `mask` once the pair's MMAs have retired
__device__ __forceinline__ void umma2_commit_mc(uint64_t* bar, uint16_t mask) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(smem_u32(bar)), "h"(mask)
      : "memory");
}

// 32 lanes x 32 columns of fp32: lane i of the warp gets TMEM lane (base_lane + i), 32 columns.
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]),
        "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]),
        "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]),
        "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
}
// 16 consecutive fp32 columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld_32x32_x16(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);      // valid only after tmem_ld_wait()
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
}
// distributed shared memory: address of `local_smem_addr` in CTA `cta` of this cluster, and a store to it
__device__ __forceinline__ uint32_t cluster_map_shared(uint32_t local_smem_addr, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(cta));
  return r;
}
// mbarrier of ANOTHER CTA of the cluster: arrive with release semantics (after st_cluster_* stores into that
// CTA), and the matching wait with acquire semantics at cluster scope
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t remote_bar_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote_bar_addr) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  const unsigned long long t0 = global_timer_ns();
  for (;;) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred P;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 P, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (ok) return;
    if (global_timer_ns() - t0 > 2000000000ull) {
      printf("ddpb: cluster mbarrier wait timed out (block %d,%d,%d thread %d)\n", blockIdx.x, blockIdx.y, blockIdx.z,
             threadIdx.x);
      __trap();
    }
  }
}
__device__ __forceinline__ void st_cluster_f32(uint32_t addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- UMMA shared-memory descriptors (SWIZZLE_128B, as written by TMA with 128-byte rows) ------
// Field layout (cute/arch/mma_sm100_desc.hpp): start>>4 [0,14), LBO>>4 [16,30), SBO>>4 [32,46),
// version=1 [46,48), layout_type [61,64) with SWIZZLE_128B = 2.
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// MN-major 32-bit (tf32) operands: layout type 1 = SWIZZLE_128B_BASE32B, written by TMA with
// CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B.  Atom = 4 K-rows x 128 bytes; SBO = distance between 4-row groups.
__device__ __forceinline__ uint64_t umma_desc_base32(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)1 << 61;
  return d;
}
// Instruction descriptor for kind::f16 with bf16 A/B and fp32 D (InstrDescriptor bitfield).
__device__ __forceinline__ constexpr uint32_t umma_idesc_bf16(int M, int N, bool a_mn_major, bool b_mn_major) {
  return (1u << 4)                         // D format: F32
         | (1u << 7)                       // A format: BF16
         | (1u << 10)                      // B format: BF16
         | ((a_mn_major ? 1u : 0u) << 15)  // A major
         | ((b_mn_major ? 1u : 0u) << 16)  // B major
         | ((uint32_t)(N >> 3) << 17)      // N / 8
         | ((uint32_t)(M >> 4) << 24);     // M / 16
}

// Instruction descriptor for kind::tf32 (fp32 operands in shared memory, fp32 D).
__device__ __forceinline__ constexpr uint32_t umma_idesc_tf32(int M, int N, bool a_mn_major, bool b_mn_major) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((a_mn_major ? 1u : 0u) << 15) | ((b_mn_major ? 1u : 0u) << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

}  // namespace ddpb
