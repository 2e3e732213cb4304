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
//                        owns (its momentum shard), then writes the new weights into every peer
//                        (all-gather).
//   MODE_ROOT      K-S4  Part 2a semantics (ref Part 2a/main.py:117-127): rank 0 gathers every
//                        rank's gradient, takes the mean, updates, scatters the result.
//   MODE_BCAST     K-S5  every rank copies rank 0's range (DDP's init / BN-buffer broadcast,
//                        ref Part 3/main.py:61).
//   MODE_ALLGATHER       rank r publishes its two-shot slice of a buffer (momentum / master shards
//                        at checkpoint time).
//   NVLS           K-S3  the two-shot (or one-shot) mode with the reduction done inside the NVSwitch.
//
// What crosses NVLink in the all-gather half.  The convolution kernels read their weights from the
// bf16 copy only; the fp32 master of a conv weight is read by nobody but the optimizer.  So for the
// element ranges the caller marks as "sharded" (the conv weights: 99.85 % of the bytes) the owner
// keeps the fp32 master to itself, exactly like the momentum, and publishes 2 bytes per element
// instead of 6.  Everything else (BN scale / shift, biases, the classifier: read as fp32 by the
// forward kernels) is published in both precisions.  Stale master slices are brought up to date
// with MODE_ALLGATHER before anyone reads a full state_dict.
//
// K-S7 barriers: each CTA b owns flags[slot][b][rank] in every rank's signal pad.  A barrier is
// "fence, store my epoch into everybody's pad, spin until everybody's epoch shows up in mine,
// fence".  Epochs are per-CTA counters kept in device memory, so a captured CUDA graph can replay
// the kernels unchanged.  Spins are bounded (DDPB_SYNC_TIMEOUT_S, default 120 s): a dead peer leaves
// a (rank, peer, slot, phase, block, epoch) record in host-mapped memory and traps the kernel
// instead of hanging the GPU.
//
// The reduction order is fixed (rank 0..W-1, or the switch's) so every rank computes identical sums.
#include <stdlib.h>

#include <mutex>

#include "../common.cuh"
#include "../../include/ddp_b200.h"

namespace ddpb {

constexpr int P2P_MAX_RANKS = DDPB_MAX_RANKS;
constexpr int P2P_MAX_BLOCKS = DDPB_P2P_MAX_BLOCKS;
constexpr int P2P_MAX_SHARD = DDPB_P2P_MAX_SHARD_RANGES;

struct P2PArgs {
  float* grad[P2P_MAX_RANKS];
  float* master[P2P_MAX_RANKS];
  bf16* shadow[P2P_MAX_RANKS];
  uint32_t* flags[P2P_MAX_RANKS];     // [slots][P2P_MAX_BLOCKS][P2P_MAX_RANKS]
  float* mom;                          // local momentum (same offsets as the arenas)
  uint32_t* epochs;                    // local [slots][P2P_MAX_BLOCKS]
  int rank, world, slot;
  int skip_exit_barrier;               // 1: a later launch of the same step ends with the barrier instead
  long long off, n;                    // element range, multiples of 8
  float lr, momentum, wd;
  unsigned long long timeout_ns;
  unsigned int* err;                   // host-mapped [8] time-out record, or null
  int kflags;                          // launch flags: trace id << 1 (see common.cuh)
  int n_shard;                         // element ranges whose fp32 master stays with its owner
  long long shard[2 * P2P_MAX_SHARD];  // [lo, hi) pairs, multiples of 8
};

__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
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
__device__ __forceinline__ void st_peer_u4(void* p, uint4 v) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// All threads of the CTA call this.  `phase` 0/1 selects the entry/exit flag set of the slot.
//   release: bar.sync (the CTA's accesses are done) -> thread 0: fence.sys, then the W flag stores
//   acquire: thread r polls rank r's flag with relaxed loads, then fence.sys -> bar.sync
// One fence per side instead of a release store / acquire load per flag and per poll.
__device__ __forceinline__ void cta_barrier_all_ranks(const P2PArgs& a, int phase) {
  if (a.world == 1 || a.flags[0] == nullptr) return;
  __syncthreads();
  __shared__ uint32_t s_epoch;
  const int idx = (a.slot * 2 + phase) * P2P_MAX_BLOCKS + blockIdx.x;
  if (threadIdx.x == 0) {
    const uint32_t e = a.epochs[idx] + 1;
    a.epochs[idx] = e;
    s_epoch = e;
    fence_sys();                     // everything this CTA (and earlier kernels) wrote is visible first
    for (int r = 0; r < a.world; ++r) st_relaxed_sys(a.flags[r] + (size_t)idx * P2P_MAX_RANKS + a.rank, e);
  }
  __syncthreads();
  const uint32_t e = s_epoch;
  if ((int)threadIdx.x < a.world) {
    const int r = threadIdx.x;
    const uint32_t* mine = a.flags[a.rank] + (size_t)idx * P2P_MAX_RANKS + r;
    // signed distance handles the (theoretical) wrap of the 32-bit epoch
    if ((int32_t)(ld_relaxed_sys(mine) - e) < 0) {
      const unsigned long long t0 = global_timer_ns();
      while ((int32_t)(ld_relaxed_sys(mine) - e) < 0) {
        if (global_timer_ns() - t0 > a.timeout_ns) {       // a peer died or never launched
          if (a.err) {
            a.err[1] = a.rank; a.err[2] = r; a.err[3] = a.slot; a.err[4] = phase; a.err[5] = blockIdx.x; a.err[6] = e;
            __threadfence_system();
            a.err[0] = 1;
            __threadfence_system();
          }
          printf("ddpb: rank %d timed out waiting for rank %d (slot %d phase %d block %d epoch %u)\n", a.rank, r,
                 a.slot, phase, blockIdx.x, e);
          __trap();
        }
      }
    }
    fence_sys();
  }
  __syncthreads();
}

__device__ __forceinline__ void slice_of(long long units, int world, int r, long long* lo, long long* hi) {
  long long per = (units + world - 1) / world;
  long long a = per * r;
  if (a > units) a = units;
  long long b = a + per;
  if (b > units) b = units;
  *lo = a; *hi = b;
}

// (slices are cut in units of 4 elements; every range handed in is a multiple of 8 elements)
// is arena element e inside one of the ranges whose fp32 master is not published?
__device__ __forceinline__ bool is_sharded(const P2PArgs& a, long long e) {
  bool in = false;
  for (int k = 0; k < a.n_shard; ++k) in |= (e >= a.shard[2 * k]) & (e < a.shard[2 * k + 1]);
  return in;
}

__device__ __forceinline__ void sgd_update(float4 g, float4& p, float4& m, float scale, float lr, float mom, float wd) {
  g.x = g.x * scale + wd * p.x; g.y = g.y * scale + wd * p.y;
  g.z = g.z * scale + wd * p.z; g.w = g.w * scale + wd * p.w;
  m.x = mom * m.x + g.x; m.y = mom * m.y + g.y; m.z = mom * m.z + g.z; m.w = mom * m.w + g.w;
  p.x -= lr * m.x; p.y -= lr * m.y; p.z -= lr * m.z; p.w -= lr * m.w;
}
__device__ __forceinline__ uint4 pack8(const float4& a, const float4& b) {
  return make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
}

enum { MODE_ONESHOT = 0, MODE_TWOSHOT = 1, MODE_ROOT = 2, MODE_BCAST = 3, MODE_ALLGATHER = 4 };

// The unit of work is one 16-byte access (4 elements) per thread with consecutive lanes on consecutive
// addresses: a warp's load instruction covers 512 contiguous bytes.  (Two adjacent 16-byte loads per thread
// - a 32-byte stride between lanes - ran at HALF the NVLink read rate: 360 instead of 690 GB/s, measured
// with tools/peer_read_probe.py.)  Shapes as for the NVLS kernel below; U units are in flight per thread.
template <int MODE, int U, int THREADS>
__global__ void __launch_bounds__(THREADS)
p2p_kernel(const P2PArgs a) {
  ktrace_begin(a.kflags);
  cta_barrier_all_ranks(a, 0);
  const long long n4 = a.n / 4;
  const long long base4 = a.off / 4;
  const long long tid = (long long)blockIdx.x * THREADS + threadIdx.x;
  const long long nthr = (long long)gridDim.x * THREADS;
  const float inv_w = 1.f / (float)a.world;

  if (MODE == MODE_ONESHOT || MODE == MODE_TWOSHOT || MODE == MODE_ROOT) {
    long long lo = 0, hi = n4;
    if (MODE == MODE_TWOSHOT) slice_of(n4, a.world, a.rank, &lo, &hi);
    if (MODE == MODE_ROOT && a.rank != 0) hi = lo;          // only the root computes
    float* const master = a.master[a.rank];
    for (long long i0 = lo + tid; i0 < hi; i0 += nthr * U) {
      float4 s[U], pp[U], mm[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        s[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        // master weights and momentum do not depend on the peers: fetch them under the same round trip
        const long long i = i0 + u * nthr;
        const long long e = (base4 + (i < hi ? i : lo)) * 4;      // clamp: a harmless in-range re-read
        pp[u] = *reinterpret_cast<const float4*>(master + e);
        mm[u] = *reinterpret_cast<const float4*>(a.mom + e);
      }
#pragma unroll
      for (int r = 0; r < P2P_MAX_RANKS; ++r) {          // fixed rank order: identical sums on every rank
        if (r < a.world) {
          float4 g[U];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const long long i = i0 + u * nthr;
            const long long e = (base4 + (i < hi ? i : lo)) * 4;
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
        float4 p = pp[u], m = mm[u];
        sgd_update(s[u], p, m, inv_w, a.lr, a.momentum, a.wd);
        *reinterpret_cast<float4*>(a.mom + e) = m;
        const uint2 sh = make_uint2(pack_bf16x2(p.x, p.y), pack_bf16x2(p.z, p.w));
        if (MODE == MODE_ONESHOT) {
          *reinterpret_cast<float4*>(master + e) = p;
          if (a.shadow[a.rank]) *reinterpret_cast<uint2*>(a.shadow[a.rank] + e) = sh;
        } else {
          // fp32 master of a sharded range stays here; peers only ever read its bf16 copy
          const bool keep = a.shadow[a.rank] != nullptr && is_sharded(a, e);
          if (keep) *reinterpret_cast<float4*>(master + e) = p;
          for (int r = 0; r < a.world; ++r) {
            if (!keep) st_peer(a.master[r] + e, p);
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
  ktrace_end(a.kflags);
}


// ------------------------------------------------------------------------------------------
// K-S3  NVLS all-reduce: the reduction happens INSIDE the NVSwitch.  `mc_*` are mappings of the
// multicast object every rank's arena is bound to (runtime/vmm_arena.cpp): a multimem.ld_reduce on
// such an address returns the sum of that element over all ranks, a multimem.st writes it to all
// ranks.  Rank r owns slice r: one switch-reduced load, the SGD update on its momentum shard, one
// multicast store of the bf16 copy (and of the fp32 weight for the few ranges that are not sharded):
// about S bytes out and S/2 bytes in per GPU for an S-byte gradient, and no per-peer loop at all.
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
__device__ __forceinline__ void multimem_st_bf16x8(void* mc, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1, %2, %3, %4};" ::"l"(mc), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}

struct NvlsArgs {
  P2PArgs base;            // flags / epochs / rank / world / slot / range / hyper-parameters, master[rank] = local
  float* mc_grad;
  float* mc_master;
  bf16* mc_shadow;         // may be null (fp32 compute)
  int one_shot;            // 1: every rank reduces the WHOLE range through the switch and updates its own replica
  int root;                // 1: rank 0 reduces and updates the whole range and multicasts the result (Part 2a's data flow)
};

// U units (one 16-byte access each, lanes contiguous) per thread and iteration.  ALL loads of an iteration -
// the switch-reduced gradients and the local master weights and momentum - are issued before anything is
// consumed, so one round trip covers them all.  Two shapes: the lean one runs next to the
// backward kernels (see lean_threads: the cap is on requests in flight); THREADS = 512 on a caller-chosen grid is
// for a caller that owns the GPU (the bandwidth sweep).
template <int U, int THREADS>
__global__ void __launch_bounds__(THREADS)
p2p_nvls_kernel(const NvlsArgs n) {
  const P2PArgs& a = n.base;
  ktrace_begin(a.kflags);
  cta_barrier_all_ranks(a, 0);
  const long long n4 = a.n / 4, base4 = a.off / 4;
  const long long tid = (long long)blockIdx.x * THREADS + threadIdx.x;
  const long long nthr = (long long)gridDim.x * THREADS;
  const float inv_w = 1.f / (float)a.world;
  long long lo = 0, hi = n4;
  if (n.root) { if (a.rank != 0) hi = lo; }
  else if (!n.one_shot) slice_of(n4, a.world, a.rank, &lo, &hi);
  float* const master = a.master[a.rank];
  for (long long i0 = lo + tid; i0 < hi; i0 += nthr * U) {
    float4 s[U], p[U], m[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * nthr;
      const long long e = (base4 + (i < hi ? i : lo)) * 4;        // clamp: a harmless in-range re-read
      s[u] = multimem_ld_reduce_add(n.mc_grad + e);
      p[u] = *reinterpret_cast<const float4*>(master + e);
      m[u] = *reinterpret_cast<const float4*>(a.mom + e);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * nthr;
      if (i >= hi) break;
      const long long e = (base4 + i) * 4;
      sgd_update(s[u], p[u], m[u], inv_w, a.lr, a.momentum, a.wd);
      *reinterpret_cast<float4*>(a.mom + e) = m[u];
      const uint2 sh = make_uint2(pack_bf16x2(p[u].x, p[u].y), pack_bf16x2(p[u].z, p[u].w));
      if (n.one_shot) {
        // every rank computed the same update from the same switch-reduced sum: purely local stores
        *reinterpret_cast<float4*>(master + e) = p[u];
        if (a.shadow[a.rank]) *reinterpret_cast<uint2*>(a.shadow[a.rank] + e) = sh;
        continue;
      }
      if (n.mc_shadow != nullptr && is_sharded(a, e)) *reinterpret_cast<float4*>(master + e) = p[u];
      else multimem_st(n.mc_master + e, p[u]);
      if (n.mc_shadow) multimem_st_bf16x4(n.mc_shadow + e, sh);
    }
  }
  if (!a.skip_exit_barrier) cta_barrier_all_ranks(a, 1);
  ktrace_end(a.kflags);
}

// Diagnostic: how fast can this GPU READ a buffer (peer memory, or the switch-reduced multicast view) with
// a given CTA shape?  kind 0 = plain 16-byte loads, 1 = ld.relaxed.sys (what the sync kernels use), 2 =
// multimem.ld_reduce.  Every thread keeps U loads in flight and folds the values into one float so that
// nothing is optimised away.  tools/peer_read_probe.py; no reference counterpart.
template <int U>
__global__ void __launch_bounds__(512)
peer_read_probe_kernel(const float* __restrict__ src, float* __restrict__ sink, long long n8, int kind, bf16* dst,
                       float* local_rw) {
  // unit = 8 floats: two 16-byte loads; kinds >= 3 also store the 8 values as bf16 (16 bytes) to dst:
  // 3 = multimem.ld_reduce + multimem.st, 4 = ld.relaxed.sys + st.relaxed.sys, 5 = kind 3 plus the optimizer's
  // local traffic (read 64 B, write 64 B per unit), 6 = kind 3 without any load (stores only)
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long nthr = (long long)gridDim.x * blockDim.x;
  float acc = 0.f;
  for (long long i0 = tid; i0 < n8; i0 += nthr * U) {
    float4 v[U][2], w[U][2];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * nthr;
      const float* p = src + (i < n8 ? i : tid) * 8;
      if (kind == 0) { v[u][0] = *reinterpret_cast<const float4*>(p); v[u][1] = *reinterpret_cast<const float4*>(p + 4); }
      else if (kind == 1 || kind == 4) { v[u][0] = ld_peer(p); v[u][1] = ld_peer(p + 4); }
      else if (kind == 6) { v[u][0] = v[u][1] = make_float4(1.f, 2.f, 3.f, 4.f); }
      else { v[u][0] = multimem_ld_reduce_add(p); v[u][1] = multimem_ld_reduce_add(p + 4); }
      if (kind == 5) {
        const float* q = local_rw + (i < n8 ? i : tid) * 16;
        w[u][0] = *reinterpret_cast<const float4*>(q); w[u][1] = *reinterpret_cast<const float4*>(q + 4);
        const float4 w2 = *reinterpret_cast<const float4*>(q + 8), w3 = *reinterpret_cast<const float4*>(q + 12);
        w[u][0].x += w2.x; w[u][1].x += w3.x;
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * nthr;
      acc += v[u][0].x + v[u][1].w;
      if (kind >= 3 && i < n8) {
        const uint4 pk = pack8(v[u][0], v[u][1]);
        if (kind == 4) st_peer_u4(dst + i * 8, pk);
        else multimem_st_bf16x8(dst + i * 8, pk);
        if (kind == 5) {
          float* q = local_rw + i * 16;
          w[u][0].x += v[u][0].x; w[u][1].y += v[u][1].y;
          *reinterpret_cast<float4*>(q) = w[u][0]; *reinterpret_cast<float4*>(q + 4) = w[u][1];
          *reinterpret_cast<float4*>(q + 8) = w[u][1]; *reinterpret_cast<float4*>(q + 12) = w[u][0];
        }
      }
    }
  }
  if (acc == 123.456f) sink[tid] = acc;
}

// ---- host side -----------------------------------------------------------------------------
static unsigned int* g_err_host = nullptr;       // pinned, mapped, portable: one record per process
static std::mutex g_err_mutex;

static unsigned int* err_record() {
  std::lock_guard<std::mutex> lock(g_err_mutex);
  if (!g_err_host) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, 64, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) {
      (void)cudaGetLastError();
      return nullptr;
    }
    memset(p, 0, 64);
    g_err_host = (unsigned int*)p;
  }
  return g_err_host;        // unified addressing: the host pointer is valid on the device
}

// The in-step ("lean") shape.  What matters next to the backward pass is how many NVLink requests are IN FLIGHT,
// not how many SMs the kernel touches: with ~150 K sixteen-byte requests queued (148 CTAs x 128 threads x 4 units x
// 2) every kernel boundary of EVERY other stream took 8-11 us instead of 1.3 us - its completion has to travel
// through the same queues - while 8 CTAs x 512 threads (32 K requests) moved the same bytes per second through the
// switch and left the boundaries at 1.5 us (tools/launch_lag_probe.py, profiles/r2/launch_lag_w2.txt).
// In the step itself the two kernels want different shapes (bench.py at 2 GPUs, 100 steps: the unicast kernel
// 0.581 ms with 32 x 256 against 0.594 with 148 x 128; the NVLS kernel 0.599 with 148 x 128 against 0.625 with
// 32 x 256 - its switch-reduced loads are the slow ones and need the requests): those are the defaults.
// DDPB_SYNC_THREADS = 128 | 256 | 512 per CTA and DDPB_SYNC_MAX_BLOCKS = CTA cap override both.
static int lean_threads(bool nvls) {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("DDPB_SYNC_THREADS");
    const int t = e ? atoi(e) : 0;
    v = (t == 128 || t == 256 || t == 512) ? t : 0;
  }
  return v ? v : (nvls ? 128 : 256);
}
static int lean_cap(bool nvls) {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("DDPB_SYNC_MAX_BLOCKS");
    v = e ? atoi(e) : 0;
    if (v < 1 || v > P2P_MAX_BLOCKS) v = 0;
  }
  if (v) return v;
  const int sms = sm_count();
  return nvls ? (sms < P2P_MAX_BLOCKS ? sms : P2P_MAX_BLOCKS) : 32;
}
// CTAs for `units` 16-byte units per rank: at least two iterations of 4 units per thread before adding a CTA
static int lean_blocks(long long units, bool nvls) {
  const long long per_cta = (long long)lean_threads(nvls) * 8;
  const long long want = (units + per_cta - 1) / per_cta;
  const int cap = lean_cap(nvls);
  return (int)(want < 1 ? 1 : (want > cap ? cap : want));
}

// blocks < 0: the fat shape (512 threads) on a grid sized by the data - a caller that owns the GPU but has little
// to move should not pay a 148-CTA cross-GPU barrier (4 MB over 8 GPUs: 65 us with 148 CTAs)
static int fat_blocks(long long units) {
  const long long per_cta = 512 * 8;
  const long long want = (units + per_cta - 1) / per_cta;
  const int cap = sm_count() < P2P_MAX_BLOCKS ? sm_count() : P2P_MAX_BLOCKS;
  return (int)(want < 1 ? 1 : (want > cap ? cap : want));
}

static unsigned long long timeout_ns() {
  static unsigned long long v = 0;
  if (!v) {
    const char* e = getenv("DDPB_SYNC_TIMEOUT_S");
    double s = e ? atof(e) : 120.0;
    if (!(s > 0.0)) s = 120.0;
    v = (unsigned long long)(s * 1e9);
  }
  return v;
}

static int fill_common(P2PArgs& a, void* const* flags, float* mom, void* epochs, int rank, int world, int slot,
                       long long off, long long n, float lr, float momentum, float wd, const long long* shard,
                       int n_shard) {
  DDPB_REQUIRE(off % 8 == 0 && n % 8 == 0, "p2p: range must be a multiple of 8 elements (off=%lld n=%lld)", off, n);
  DDPB_REQUIRE(slot >= 0 && slot < DDPB_P2P_MAX_SLOTS, "p2p: slot %d out of range", slot);
  DDPB_REQUIRE(n_shard >= 0 && n_shard <= P2P_MAX_SHARD && (n_shard == 0 || shard), "p2p: at most %d sharded ranges",
               P2P_MAX_SHARD);
  for (int r = 0; r < world; ++r) a.flags[r] = flags ? (uint32_t*)flags[r] : nullptr;
  a.mom = mom; a.epochs = (uint32_t*)epochs;
  a.rank = rank; a.world = world; a.slot = slot; a.off = off; a.n = n;
  a.lr = lr; a.momentum = momentum; a.wd = wd;
  a.timeout_ns = timeout_ns();
  a.err = err_record();
  a.n_shard = 0;
  for (int k = 0; k < n_shard; ++k) {
    const long long lo = shard[2 * k], hi = shard[2 * k + 1];
    DDPB_REQUIRE(lo % 8 == 0 && hi % 8 == 0 && lo <= hi, "p2p: sharded range [%lld, %lld) must be 8-element aligned", lo, hi);
    if (hi <= off || lo >= off + n) continue;       // does not touch this launch
    a.shard[2 * a.n_shard] = lo; a.shard[2 * a.n_shard + 1] = hi;
    ++a.n_shard;
  }
  return DDPB_OK;
}

}  // namespace ddpb
DDPB_KTRACE_SETTER(sync_p2p)

using namespace ddpb;

// One fused synchronisation step over the element range [off, off+n) of the symmetric arenas.
// Pointer tables have `world` entries (index = rank); flags == null disables the cross-rank
// barriers (single-GPU emulation in the unit tests).  See include/ddp_b200.h for the modes.
// shard / n_shard: [lo, hi) element ranges whose fp32 master the owner keeps to itself (two-shot and
// root modes, only when a bf16 copy exists); n_shard = 0 publishes everything.
extern "C" int ddpb_p2p_sync_sharded(int mode, void* const* grad, void* const* master, void* const* shadow,
                                     void* const* flags, float* mom, void* epochs, int rank, int world, int slot,
                                     long long off, long long n, float lr, float momentum, float wd, int blocks,
                                     const long long* shard, int n_shard, cudaStream_t stream) {
  DDPB_REQUIRE(world >= 1 && world <= P2P_MAX_RANKS && rank >= 0 && rank < world, "p2p: bad rank/world %d/%d",
               rank, world);
  P2PArgs a;
  memset(&a, 0, sizeof(a));
  int rc = fill_common(a, flags, mom, epochs, rank, world, slot, off, n, lr, momentum, wd, shard, n_shard);
  if (rc) return rc;
  // blocks = 0: the lean shape that runs next to the backward kernels (see lean_threads);
  // blocks > 0: the caller owns the GPU and gets 512-thread CTAs on the grid it asks for; blocks < 0: the same
  // CTAs on a grid sized by the data (fat_blocks)
  const bool lean = blocks == 0;
  const long long units = n / 4 / ((mode & ~DDPB_P2P_NO_EXIT_BARRIER) == DDPB_P2P_TWOSHOT ? world : 1);
  if (lean) blocks = lean_blocks(units, false);
  else if (blocks < 0) blocks = fat_blocks(units);
  DDPB_REQUIRE(blocks <= P2P_MAX_BLOCKS, "p2p: at most %d blocks", P2P_MAX_BLOCKS);
  for (int r = 0; r < world; ++r) {
    a.grad[r] = (float*)grad[r];
    a.master[r] = master ? (float*)master[r] : nullptr;
    a.shadow[r] = shadow ? (bf16*)shadow[r] : nullptr;
  }
  a.skip_exit_barrier = (mode & DDPB_P2P_NO_EXIT_BARRIER) ? 1 : 0;
  a.kflags = kernel_flags(stream, 0);
  mode &= ~DDPB_P2P_NO_EXIT_BARRIER;
#define DDPB_P2P_LAUNCH(M)                                                          \
  do {                                                                              \
    if (lean && lean_threads(false) == 128) p2p_kernel<M, 4, 128><<<blocks, 128, 0, stream>>>(a);   \
    else if (lean && lean_threads(false) == 256) p2p_kernel<M, 4, 256><<<blocks, 256, 0, stream>>>(a);   \
    else p2p_kernel<M, 4, 512><<<blocks, 512, 0, stream>>>(a);                      \
  } while (0)
  switch (mode) {
    case DDPB_P2P_ONESHOT: DDPB_P2P_LAUNCH(MODE_ONESHOT); break;
    case DDPB_P2P_TWOSHOT: DDPB_P2P_LAUNCH(MODE_TWOSHOT); break;
    case DDPB_P2P_ROOT: DDPB_P2P_LAUNCH(MODE_ROOT); break;
    case DDPB_P2P_BCAST: DDPB_P2P_LAUNCH(MODE_BCAST); break;
    case DDPB_P2P_ALLGATHER: DDPB_P2P_LAUNCH(MODE_ALLGATHER); break;
    default: DDPB_REQUIRE(false, "p2p: unknown mode %d", mode);
  }
#undef DDPB_P2P_LAUNCH
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}

extern "C" int ddpb_p2p_sync(int mode, void* const* grad, void* const* master, void* const* shadow,
                             void* const* flags, float* mom, void* epochs, int rank, int world, int slot,
                             long long off, long long n, float lr, float momentum, float wd, int blocks,
                             cudaStream_t stream) {
  return ddpb_p2p_sync_sharded(mode, grad, master, shadow, flags, mom, epochs, rank, world, slot, off, n, lr, momentum,
                               wd, blocks, nullptr, 0, stream);
}

// NVLS variant of the fused reduce+SGD (see p2p_nvls_kernel).  mc_* point into the multicast
// mapping of the symmetric arena at the same offsets as grad / master / shadow; master_local, shadow_local
// (nullable) and mom are this rank's own buffers; `flags` is the usual per-rank table of signal pads.
// mode: DDPB_P2P_TWOSHOT (owner of a slice reduces, updates, multicasts) or DDPB_P2P_ONESHOT (every rank
// pulls the switch-reduced sum of the whole range and updates its own replica: small ranges, no stores
// over NVLink at all).  blocks = 0 picks a grid meant to run NEXT TO the backward kernels; a caller that
// owns the GPU (the bandwidth sweep) passes the grid it wants (up to DDPB_P2P_MAX_BLOCKS).
extern "C" int ddpb_p2p_sync_nvls_sharded(int mode, void* mc_grad, void* mc_master, void* mc_shadow,
                                          float* master_local, void* shadow_local, void* const* flags, float* mom,
                                          void* epochs, int rank, int world, int slot, long long off, long long n,
                                          float lr, float momentum, float wd, int blocks, int skip_exit_barrier,
                                          const long long* shard, int n_shard, cudaStream_t stream) {
  DDPB_REQUIRE(world >= 2 && world <= P2P_MAX_RANKS && rank >= 0 && rank < world, "nvls: bad rank/world %d/%d", rank,
               world);
  DDPB_REQUIRE(mode == DDPB_P2P_TWOSHOT || mode == DDPB_P2P_ONESHOT || mode == DDPB_P2P_ROOT,
               "nvls: mode must be two-shot, one-shot or root");
  DDPB_REQUIRE(mc_grad && mc_master && master_local && mom && flags, "nvls: null pointer");
  NvlsArgs a;
  memset(&a, 0, sizeof(a));
  int rc = fill_common(a.base, flags, mom, epochs, rank, world, slot, off, n, lr, momentum, wd, shard, n_shard);
  if (rc) return rc;
  // blocks = 0: the lean shape that runs next to the backward kernels (see lean_threads);
  // blocks > 0: the caller owns the GPU and gets the fat shape on the grid it asks for; blocks < 0: fat, grid by size
  const bool lean = blocks == 0;
  const long long units = n / 4 / (mode == DDPB_P2P_TWOSHOT ? world : 1);
  if (lean) blocks = lean_blocks(units, true);
  else if (blocks < 0) blocks = fat_blocks(units);
  DDPB_REQUIRE(blocks <= P2P_MAX_BLOCKS, "nvls: at most %d blocks", P2P_MAX_BLOCKS);
  a.base.master[rank] = master_local;
  a.base.shadow[rank] = (bf16*)shadow_local;
  a.base.skip_exit_barrier = skip_exit_barrier;
  a.mc_grad = (float*)mc_grad; a.mc_master = (float*)mc_master; a.mc_shadow = (bf16*)mc_shadow;
  a.one_shot = mode == DDPB_P2P_ONESHOT ? 1 : 0;
  a.root = mode == DDPB_P2P_ROOT ? 1 : 0;
  a.base.kflags = kernel_flags(stream, 0);
  if (lean && lean_threads(true) == 128) p2p_nvls_kernel<4, 128><<<blocks, 128, 0, stream>>>(a);
  else if (lean && lean_threads(true) == 256) p2p_nvls_kernel<4, 256><<<blocks, 256, 0, stream>>>(a);
  else p2p_nvls_kernel<4, 512><<<blocks, 512, 0, stream>>>(a);
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}

extern "C" int ddpb_p2p_sync_nvls(void* mc_grad, void* mc_master, void* mc_shadow, float* master_local,
                                  void* const* flags, float* mom, void* epochs, int rank, int world, int slot,
                                  long long off, long long n, float lr, float momentum, float wd, int blocks,
                                  int skip_exit_barrier, cudaStream_t stream) {
  return ddpb_p2p_sync_nvls_sharded(DDPB_P2P_TWOSHOT, mc_grad, mc_master, mc_shadow, master_local, nullptr, flags, mom,
                                    epochs, rank, world, slot, off, n, lr, momentum, wd, blocks, skip_exit_barrier,
                                    nullptr, 0, stream);
}

extern "C" int ddpb_probe_peer_read(const void* src, void* sink, long long n_floats, int kind, int blocks, int threads,
                                    int unroll, void* dst_bf16, float* local_rw, cudaStream_t stream) {
  DDPB_REQUIRE(src && sink && n_floats % 8 == 0 && blocks > 0 && threads > 0 && threads <= 512, "peer_read_probe: bad argument");
  DDPB_REQUIRE(kind < 3 || dst_bf16, "peer_read_probe: kinds >= 3 need a destination");
  DDPB_REQUIRE(kind != 5 || local_rw, "peer_read_probe: kind 5 needs a local read/write buffer of 2 * n floats");
  const long long n8 = n_floats / 8;
#define DDPB_PROBE(UU) peer_read_probe_kernel<UU><<<blocks, threads, 0, stream>>>((const float*)src, (float*)sink, n8, kind, (bf16*)dst_bf16, local_rw)
  switch (unroll) {
    case 1: DDPB_PROBE(1); break;
    case 2: DDPB_PROBE(2); break;
    case 4: DDPB_PROBE(4); break;
    case 8: DDPB_PROBE(8); break;
    default: DDPB_REQUIRE(false, "peer_read_probe: unroll must be 1, 2, 4 or 8");
  }
#undef DDPB_PROBE
  DDPB_LAUNCH_CHECK();
  return DDPB_OK;
}

// The record a timed-out cross-rank wait leaves behind (rec[0] = 1 when present, then rank, peer, slot,
// phase, block, epoch): lets the caller turn the generic "unspecified launch failure" that follows the
// trap into a specific error.  Returns 1 when a record is present.
extern "C" int ddpb_p2p_last_timeout(unsigned int* rec8) {
  std::lock_guard<std::mutex> lock(g_err_mutex);
  if (!g_err_host || !g_err_host[0]) return 0;
  if (rec8) memcpy(rec8, g_err_host, 8 * sizeof(unsigned int));
  return 1;
}
