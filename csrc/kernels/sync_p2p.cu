// Peer-to-peer gradient synchronisation over NVLink, fused with averaging, the bf16 cast and the
// SGD momentum / weight-decay update.  No NCCL call and no separate elementwise kernel runs on
// these paths.
//
// Every rank maps every peer's symmetric arena (gradients, fp32 master weights, bf16 compute
// copy, a signal pad) into its own address space (runtime/symm_mem.cpp), so the kernels below
// simply dereference peer pointers; NVSwitch gives each GPU full bandwidth to every peer.
//
//   MODE_ONESHOT   K-S1  every rank reads all W copies of the range, reduces in rank order,
//                        updates its own replica.                     (small ranges: 1 hop)
//   MODE_TWOSHOT   K-S2  rank r reduces slice r (reduce-scatter), applies SGD to the slice it
//                        owns (its momentum shard), then writes the new fp32 weights and bf16
//                        copy into every peer (all-gather).  2(W-1)/W*S bytes per direction.
//   MODE_ROOT      K-S4  Part 2a semantics (ref Part 2a/main.py:117-127): rank 0 gathers every
//                        rank's gradient, takes the mean, updates, scatters the result.
//   MODE_BCAST     K-S5  every rank copies rank 0's range (DDP's init / BN-buffer broadcast,
//                        ref Part 3/main.py:61).
//   MODE_ALLGATHER       rank r publishes its two-shot slice of a buffer (momentum shards at
//                        checkpoint time).
//
// K-S7 barriers: each CTA b owns flags[slot][b][rank] in every rank's signal pad.  A barrier is
// "store my epoch into everybody's pad (release.sys), spin until everybody's epoch shows up in
// mine (acquire.sys)".  Epochs are per-CTA counters kept in device memory, so a captured CUDA
// graph can replay the kernels unchanged.  Spins are bounded: a dead peer traps the kernel
// instead of hanging the GPU.
//
// The reduction order is fixed (rank 0..W-1) so every rank computes bit-identical sums.
#include <stdlib.h>

#include "../common.cuh"
#include "../../include/ddp_b200.h"

namespace ddpb {

constexpr int P2P_MAX_RANKS = DDPB_MAX_RANKS;
constexpr int P2P_MAX_BLOCKS = DDPB_P2P_MAX_BLOCKS;

struct P2PArgs {
  float* grad[P2P_MAX_RANKS];
  float* master[P2P_MAX_RANKS];
  bf16* shadow[P2P_MAX_RANKS];
  uint32_t* flags[P2P_MAX_RANKS];     // [slots][P2P_MAX_BLOCKS][P2P_MAX_RANKS]
  float* mom;                          // local momentum (same offsets as the arenas)
  uint32_t* epochs;                    // local [slots][P2P_MAX_BLOCKS]
  int rank, world, slot;
  int skip_exit_barrier;               // 1: a later launch of the same step ends with the barrier instead
  long long off, n;                    // element range, multiples of 4
  float lr, momentum, wd;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// peer data: bypass L1 (each element is read once; never serve it from a stale line)
__device__ __forceinline__ float4 ld_peer(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_peer(float* p, float4 v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void st_peer_u2(void* p, uint2 v) {
  asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

// All threads of the CTA call this.  `phase` 0/1 selects the entry/exit flag set of the slot.
__device__ __forceinline__ void cta_barrier_all_ranks(const P2PArgs& a, int phase) {
  if (a.world == 1 || a.flags[0] == nullptr) return;
  __syncthreads();
  __shared__ uint32_t s_epoch;
  const int idx = (a.slot * 2 + phase) * P2P_MAX_BLOCKS + blockIdx.x;
  if (threadIdx.x == 0) {
    uint32_t e = a.epochs[idx] + 1;
    a.epochs[idx] = e;
    s_epoch = e;
    __threadfence_system();          // everything this CTA (and earlier kernels) wrote is visible first
  }
  __syncthreads();
  const uint32_t e = s_epoch;
  if ((int)threadIdx.x < a.world) {
    const int r = threadIdx.x;
    st_release_sys(a.flags[r] + (size_t)idx * P2P_MAX_RANKS + a.rank, e);
    const uint32_t* mine = a.flags[a.rank] + (size_t)idx * P2P_MAX_RANKS + r;
    const unsigned long long t0 = global_timer_ns();
    // signed distance handles the (theoretical) wrap of the 32-bit epoch
    while ((int32_t)(ld_acquire_sys(mine) - e) < 0) {
      if (global_timer_ns() - t0 > 20000000000ull) {      // 20 s: a peer died or never launched
        printf("ddpb: rank %d timed out waiting for rank %d (slot %d phase %d block %d epoch %u)\n", a.rank, r,
               a.slot, phase, blockIdx.x, e);
        __trap();
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ void slice_of(long long n4, int world, int r, long long* lo, long long* hi) {
  long long per = (n4 + world - 1) / world;
  long long a = per * r;
  if (a > n4) a = n4;
  long long b = a + per;
  if (b > n4) b = n4;
  *lo = a; *hi = b;
}

__device__ __forceinline__ float4 sgd_update(float4 g, float4& p, float4& m, float scale, float lr, float mom,
                                             float wd) {
  g.x = g.x * scale + wd * p.x; g.y = g.y * scale + wd * p.y;
  g.z = g.z * scale + wd * p.z; g.w = g.w * scale + wd * p.w;
  m.x = mom * m.x + g.x; m.y = mom * m.y + g.y; m.z = mom * m.z + g.z; m.w = mom * m.w + g.w;
  p.x -= lr * m.x; p.y -= lr * m.y; p.z -= lr * m.z; p.w -= lr * m.w;
  return p;
}

enum { MODE_ONESHOT = 0, MODE_TWOSHOT = 1, MODE_ROOT = 2, MODE_BCAST = 3, MODE_ALLGATHER = 4 };

template <int MODE>
__global__ void __launch_bounds__(512)
p2p_kernel(const P2PArgs a) {
  cta_barrier_all_ranks(a, 0);
  const long long n4 = a.n / 4;
  const long long base4 = a.off / 4;
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long nthr = (long long)gridDim.x * blockDim.x;
  const float inv_w = 1.f / (float)a.world;

  if (MODE == MODE_ONESHOT || MODE == MODE_TWOSHOT || MODE == MODE_ROOT) {
    long long lo = 0, hi = n4;
    if (MODE == MODE_TWOSHOT) slice_of(n4, a.world, a.rank, &lo, &hi);
    if (MODE == MODE_ROOT && a.rank != 0) hi = lo;          // only the root computes
    // A peer load takes ~2 us over NVLink: each thread keeps U independent 16-byte loads per rank in
    // flight (U elements `nthr` apart), otherwise the loop is bound by latency, not by the link.
    constexpr int U = 4;
    for (long long i0 = lo + tid; i0 < hi; i0 += nthr * U) {
      float4 s[U];
#pragma unroll
      for (int u = 0; u < U; ++u) s[u] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int r = 0; r < P2P_MAX_RANKS; ++r) {          // fixed rank order: identical sums on every rank
        if (r < a.world) {
          float4 g[U];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const long long i = i0 + u * nthr;
            const long long e = (base4 + (i < hi ? i : lo)) * 4;      // clamp: a harmless in-range re-read
            g[u] = (r == a.rank) ? *reinterpret_cast<const float4*>(a.grad[r] + e) : ld_peer(a.grad[r] + e);
          }
#pragma unroll
          for (int u = 0; u < U; ++u) { s[u].x += g[u].x; s[u].y += g[u].y; s[u].z += g[u].z; s[u].w += g[u].w; }
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = i0 + u * nthr;
        if (i >= hi) break;
        const long long e = (base4 + i) * 4;
        float4 p = *reinterpret_cast<const float4*>(a.master[a.rank] + e);
        float4 m = *reinterpret_cast<const float4*>(a.mom + e);
        sgd_update(s[u], p, m, inv_w, a.lr, a.momentum, a.wd);
        *reinterpret_cast<float4*>(a.mom + e) = m;
        uint2 sh = make_uint2(pack_bf16x2(p.x, p.y), pack_bf16x2(p.z, p.w));
        if (MODE == MODE_ONESHOT) {
          *reinterpret_cast<float4*>(a.master[a.rank] + e) = p;
          if (a.shadow[a.rank]) *reinterpret_cast<uint2*>(a.shadow[a.rank] + e) = sh;
        } else {
          for (int r = 0; r < a.world; ++r) {
            st_peer(a.master[r] + e, p);
            if (a.shadow[r]) st_peer_u2(a.shadow[r] + e, sh);
          }
        }
      }
    }
  } else if (MODE == MODE_BCAST) {
    // `grad` holds the buffer being broadcast (any fp32 range of the symmetric arena)
    if (a.rank != 0)
      for (long long i = tid; i < n4; i += nthr) {
        const long long e = (base4 + i) * 4;
        *reinterpret_cast<float4*>(a.grad[a.rank] + e) = ld_peer(a.grad[0] + e);
      }
  } else if (MODE == MODE_ALLGATHER) {
    long long lo, hi;
    slice_of(n4, a.world, a.rank, &lo, &hi);
    for (long long i = lo + tid; i < hi; i += nthr) {
      const long long e = (base4 + i) * 4;
      float4 v = *reinterpret_cast<const float4*>(a.grad[a.rank] + e);
      for (int r = 0; r < a.world; ++r)
        if (r != a.rank) st_peer(a.grad[r] + e, v);
    }
  }
  if (!a.skip_exit_barrier) cta_barrier_all_ranks(a, 1);
}


// ------------------------------------------------------------------------------------------
// K-S3  NVLS all-reduce: the reduction happens INSIDE the NVSwitch.  `mc_*` are mappings of the
// multicast object every rank's arena is bound to (runtime/vmm_arena.cpp): a multimem.ld_reduce on
// such an address returns the sum of that element over all ranks, a multimem.st writes it to all
// ranks.  Rank r owns slice r: one switch-reduced load, the SGD update on its momentum shard, one
// multicast store of the new fp32 weight and one of the bf16 copy - about S bytes per direction per
// GPU instead of 2(W-1)/W*S for the two-shot kernel, and no per-peer loop at all.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 multimem_ld_reduce_add(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void multimem_st_bf16x4(void* mc, uint2 v) {
  asm volatile("multimem.st.relaxed.sys.global.v2.bf16x2 [%0], {%1, %2};" ::"l"(mc), "r"(v.x), "r"(v.y) : "memory");
}

struct NvlsArgs {
  P2PArgs base;            // flags / epochs / rank / world / slot / range / hyper-parameters, master[rank] = local
  float* mc_grad;
  float* mc_master;
  bf16* mc_shadow;         // may be null (fp32 compute)
};

__global__ void __launch_bounds__(512)
p2p_nvls_kernel(const NvlsArgs n) {
  const P2PArgs& a = n.base;
  cta_barrier_all_ranks(a, 0);
  const long long n4 = a.n / 4, base4 = a.off / 4;
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long nthr = (long long)gridDim.x * blockDim.x;
  const float inv_w = 1.f / (float)a.world;
  long long lo, hi;
  slice_of(n4, a.world, a.rank, &lo, &hi);
  constexpr int U = 4;
  for (long long i0 = lo + tid; i0 < hi; i0 += nthr * U) {
    float4 s[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * nthr;
      s[u] = multimem_ld_reduce_add(n.mc_grad + (base4 + (i < hi ? i : lo)) * 4);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * nthr;
      if (i >= hi) break;
      const long long e = (base4 + i) * 4;
      float4 p = *reinterpret_cast<const float4*>(a.master[a.rank] + e);
      float4 m = *reinterpret_cast<const float4*>(a.mom + e);
      sgd_update(s[u], p, m, inv_w, a.lr, a.momentum, a.wd);
      *reinterpret_cast<float4*>(a.mom + e) = m;
      multimem_st(n.mc_master + e, p);
      if (n.mc_shadow) multimem_st_bf16x4(n.mc_shadow + e, make_uint2(pack_bf16x2(p.x, p.y), pack_bf16x2(p.z, p.w)));
    }
  }
  if (!a.skip_exit_barrier) cta_barrier_all_ranks(a, 1);
}

}  // namespace ddpb

using namespace ddpb;

// One fused synchronisation step over the element range [off, off+n) of the symmetric arenas.
// Pointer tables have `world` entries (index = rank); flags == null disables the cross-rank
// barriers (single-GPU emulation in the unit tests).  See include/ddp_b200.h for the modes.
extern "C" int ddpb_p2p_sync(int mode, void* const* grad, void* const* master, void* const* shadow,
                             void* const* flags, float* mom, void* epochs, int rank, int world, int slot,
                             long long off, long long n, float lr, float momentum, float wd, int blocks,
                             cudaStream_t stream) {
  DDPB_REQUIRE(world >= 1 && world <= P2P_MAX_RANKS && rank >= 0 && rank < world, "p2p: bad rank/world %d/%d",
               rank, world);
  DDPB_REQUIRE(off % 4 == 0 && n % 4 == 0, "p2p: range must be float4-aligned (off=%lld n=%lld)", off, n);
  DDPB_REQUIRE(slot >= 0 && slot < DDPB_P2P_MAX_SLOTS, "p2p: slot %d out of range", slot);
  if (blocks <= 0) {
    long long want = (n / 4 + 512 * 8 - 1) / (512 * 8);
    blocks = (int)(want < 1 ? 1 : (want > 32 ? 32 : want));
  }
  DDPB_REQUIRE(blocks <= P2P_MAX_BLOCKS, "p2p: at most %d blocks", P2P_MAX_BLOCKS);
  P2PArgs a;
  memset(&a, 0, sizeof(a));
  for (int r = 0; r < world; ++r) {
    a.grad[r] = (float*)grad[r];
    a.master[r] = master ? (float*)master[r] : nullptr;
    a.shadow[r] = shadow ? (bf16*)shadow[r] : nullptr;
    a.flags[r] = flags ? (uint32_t*)flags[r] : nullptr;
  }
  a.mom = mom; a.epochs = (uint32_t*)epochs;
  a.rank = rank; a.world = world; a.slot = slot; a.off = off; a.n = n;
  a.lr = lr; a.momentum = momentum; a.wd = wd;
  a.skip_exit_barrier = (mode & DDPB_P2P_NO_EXIT_BARRIER) ? 1 : 0;
  mode &= ~DDPB_P2P_NO_EXIT_BARRIER;
  switch (mode) {
    case DDPB_P2P_ONESHOT: p2p_kernel<MODE_ONESHOT><<<blocks, 512, 0, stream>>>(a); break;
    case DDPB_P2P_TWOSHOT: p2p_kernel<MODE_TWOSHOT><<<blocks, 512, 0, stream>>>(a); break;
    case DDPB_P2P_ROOT: p2p_kernel<MODE_ROOT><<<blocks, 512, 0, stream>>>(a); break;
    case DDPB_P2P_BCAST: p2p_kernel<MODE_BCAST><<<blocks, 512, 0, stream>>>(a); break;
    case DDPB_P2P_ALLGATHER: p2p_kernel<MODE_ALLGATHER><<<blocks, 512, 0, stream>>>(a); break;
    default: DDPB_REQUIRE(false, "p2p: unknown mode %d", mode);
  }
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}

// NVLS variant of the two-shot fused reduce+SGD (see p2p_nvls_kernel).  mc_* point into the multicast
// mapping of the symmetric arena at the same offsets as grad / master / shadow; master_local and mom
// are this rank's own buffers; `flags` is the usual per-rank table of signal pads.
extern "C" int ddpb_p2p_sync_nvls(void* mc_grad, void* mc_master, void* mc_shadow, float* master_local,
                                  void* const* flags, float* mom, void* epochs, int rank, int world, int slot,
                                  long long off, long long n, float lr, float momentum, float wd, int blocks,
                                  int skip_exit_barrier, cudaStream_t stream) {
  DDPB_REQUIRE(world >= 2 && world <= P2P_MAX_RANKS && rank >= 0 && rank < world, "nvls: bad rank/world %d/%d", rank,
               world);
  DDPB_REQUIRE(off % 4 == 0 && n % 4 == 0, "nvls: range must be float4-aligned");
  DDPB_REQUIRE(slot >= 0 && slot < DDPB_P2P_MAX_SLOTS, "nvls: slot %d out of range", slot);
  DDPB_REQUIRE(mc_grad && mc_master && master_local && mom && flags, "nvls: null pointer");
  // CTA shape of the NVLS kernel (it runs next to the backward kernels; DDPB_NVLS_THREADS / DDPB_NVLS_MAX_BLOCKS)
  static int threads = 0, cap = 0;
  if (!threads) {
    const char* t = getenv("DDPB_NVLS_THREADS");
    threads = t ? atoi(t) : 256;        // measured at 2 GPUs: 256 x 32 CTAs 0.575 ms, 512 x 32 0.589, 512 x 64 0.631, 256 x 16 0.624
    if (threads != 128 && threads != 256 && threads != 512) threads = 256;
    const char* c = getenv("DDPB_NVLS_MAX_BLOCKS");
    cap = c ? atoi(c) : 32;
    if (cap < 1 || cap > P2P_MAX_BLOCKS) cap = 32;
  }
  if (blocks <= 0) {
    long long want = (n / 4 / world + threads * 4 - 1) / (threads * 4);
    blocks = (int)(want < 1 ? 1 : (want > cap ? cap : want));
  }
  DDPB_REQUIRE(blocks <= P2P_MAX_BLOCKS, "nvls: at most %d blocks", P2P_MAX_BLOCKS);
  NvlsArgs a;
  memset(&a, 0, sizeof(a));
  for (int r = 0; r < world; ++r) a.base.flags[r] = (uint32_t*)flags[r];
  a.base.master[rank] = master_local;
  a.base.mom = mom; a.base.epochs = (uint32_t*)epochs;
  a.base.rank = rank; a.base.world = world; a.base.slot = slot; a.base.off = off; a.base.n = n;
  a.base.lr = lr; a.base.momentum = momentum; a.base.wd = wd;
  a.base.skip_exit_barrier = skip_exit_barrier;
  a.mc_grad = (float*)mc_grad; a.mc_master = (float*)mc_master; a.mc_shadow = (bf16*)mc_shadow;
  p2p_nvls_kernel<<<blocks, threads, 0, stream>>>(a);
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}
