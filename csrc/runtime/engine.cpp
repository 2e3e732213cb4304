// The VGG training-step engine: owns every device buffer, runs forward / backward as a static
// sequence of sm_100a kernel launches, and schedules the gradient buckets.
//
// What it replaces in the reference: the autograd-driven loop body
//   zero_grad -> model(data) -> criterion -> loss.backward() -> [sync] -> optimizer.step()
// (ref Part 2a/main.py:86-98) plus, for Part 3, torch DDP's C++ Reducer (bucket assignment, ready
// counting, asynchronous all-reduce launch, finalize; ref Part 3/main.py:61).
//
// Design notes
//  * The network is static, so the gradient-ready order is known up front: the
//    arena that holds gradients / fp32 master weights / momentum / bf16 compute weights is laid
//    out in THAT order and every bucket is one contiguous range.  wgrad, BN-backward and the
//    classifier backward write their gradients straight into the arena (no pack/unpack copies).
//  * Bucket scheduler: when the last producer of a bucket has been enqueued on the compute
//    stream, an event is recorded; the communication stream waits on it and launches ONE fused
//    kernel (peer-to-peer reduce + 1/W + SGD momentum/weight decay + bf16 refresh, sync_p2p.cu)
//    that overlaps with the rest of backward.  The step ends by joining the two streams.
//  * The whole step (both streams) is captured once per batch size into a CUDA graph and
//    replayed, so the host cost of a step is one graph launch.
#include <cuda_runtime.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "../../include/ddp_b200.h"
#include "../common.cuh"

namespace ddpb {

static constexpr int kAlign = 64;    // arena slot alignment in elements (256 B of fp32)

static long long round_up(long long x, long long m) { return (x + m - 1) / m * m; }

struct Slot { long long off, numel; };

struct Block {
  int cin, cout, hw, pool;
  Slot conv_w, conv_b, bn_w, bn_b;
  long long stat_off;    // offset (floats) of this block's [2C] in the fwd-stats / bwd-sums scratch
  long long ch_off;      // offset of this block's [C] in mean / invstd / running stats
  void* y;               // conv output  [B,hw,hw,cout]
  void* act;             // block output [B,ohw,ohw,cout]
  void* win = nullptr;   // pooling blocks, bf16 tensor-core path: conv output that won each pool window (NaN = none)
};

struct Bucket {
  long long off, numel; int last_block; int mode;   // last_block: -1 = classifier
  int part, parts;       // sub-bucket hand-off: this is part `part` of `parts` of its layer (output-channel ranges of the conv weight)
};

typedef std::map<std::string, long long> NameCounts;
struct Graph { cudaGraphExec_t exec = nullptr; long long launches = 0; NameCounts by_name; };

struct Engine {
  ddpb_engine_config cfg;
  int esize;                       // activation element size
  std::vector<Block> blocks;
  Slot fc_w, fc_b;
  long long arena_elems = 0, total_c = 0;
  // symmetric arena
  void* arena = nullptr; size_t arena_bytes = 0; unsigned char ipc_handle[64];
  float *grad = nullptr, *master = nullptr, *mom = nullptr, *rmean = nullptr, *rvar = nullptr;
  bf16* shadow = nullptr; uint32_t* flags = nullptr;
  size_t off_grad = 0, off_master = 0, off_mom = 0, off_shadow = 0, off_rstats = 0, off_flags = 0;
  void* peer_arena[DDPB_MAX_RANKS] = {nullptr};
  bool peers_set = false;
  bool arena_external = false;     // caller-owned arena: never freed / IPC-closed here
  void* mc_arena = nullptr;        // multicast mapping of the arena (NVLS), or null
  // local buffers
  uint32_t* epochs = nullptr;
  float *scratch = nullptr, *mean = nullptr, *invstd = nullptr, *logits = nullptr, *dlogits = nullptr;
  float *loss = nullptr, *eval_loss = nullptr; int *correct = nullptr; long long* nbt = nullptr;
  size_t scratch_floats = 0;
  void *input = nullptr, *dy_buf2[2] = {nullptr, nullptr}, *dact = nullptr; long long* target = nullptr;
  bool fuse_bn_fwd = false;                   // DDPB_FUSE_BN_FWD=1: conv + grid barrier + BatchNorm forward in one kernel
                                              // (measured ~10 us per step SLOWER than the separate, PDL-launched kernel: off)
  bool fuse_bn_reduce = true;                 // DDPB_FUSE_BN_REDUCE=0: keep the separate reduction kernel
  bool tf32 = true;                           // DDPB_TF32=0: fp32 engine on the SIMT kernels (exact fp32 products)
  bool fuse_first_bwd = true;                 // DDPB_FUSE_FIRST_BWD=0: block 0 backward as BN-apply + weight-gradient kernels
  void *x0cols = nullptr, *w0pad = nullptr;   // first layer as a GEMM: patch matrix [B*1024][64], weights [64][64]
  std::vector<Bucket> buckets;
  std::vector<long long> shard;          // [lo, hi) arena ranges of the conv weights: fp32 master stays with its owner
  bool fuse_fc_da = false;               // DDPB_FUSE_FC_DA=1: the loss kernel also produces the classifier's input gradient
  bool shard_master = true;              // DDPB_SHARD_MASTER=0: publish the fp32 master of every element
  int sub_buckets = 1;                   // DDPB_SUB_BUCKETS=2|4: split the large layers' buckets by output-channel range
  long long sub_bucket_min = 1 << 20;    // ... layers with at least this many weight elements (DDPB_SUB_BUCKET_MIN)
  bool sub_bucket_force = false;         // DDPB_SUB_BUCKETS_FORCE=1: also on one GPU (tests)
  std::vector<cudaEvent_t> ev_part;      // [block][part] weight-gradient part complete
  bool nvls_after_bwd = true;            // Parts 2a / 2b on a multicast arena: in-switch reduction (DDPB_NVLS_AFTER_BWD=0: unicast)
  bool fat_after_bwd = true;             // Parts 2a / 2b: 512-thread CTAs, grid by size (DDPB_FAT_AFTER_BWD=0: the in-step shape)
  bool zero_async = true;                // DDPB_ZERO_ASYNC=0: clear the gradients at the start of backward, on its stream
  bool nvls_oneshot = true;              // DDPB_NVLS_ONESHOT=0: small buckets use the unicast one-shot kernel
  cudaStream_t comm = nullptr, comm2 = nullptr, wg = nullptr;
  bool two_comm = true, comm2_used = false;   // multi-GPU: buckets alternate between two streams (DDPB_TWO_COMM=0: one)
  std::vector<cudaEvent_t> bucket_ev, ev_dy, ev_dg, ev_wg;
  cudaEvent_t join_ev = nullptr, fork_ev = nullptr, zero_ev = nullptr, start_ev = nullptr;
  std::map<long long, Graph> graphs;     // key = kind * 1e6 + batch
  long long launches = 0, counting = 0;
  NameCounts launches_by_name, counting_by_name;    // per entry point: what the step really launched
  bool count_only = false;
  // step-timeline profiling (ddpb_engine_profile_step): timing events, only recorded when prof_on
  bool prof_on = false;
  cudaEvent_t prof_start = nullptr, prof_fwd = nullptr, prof_bwd = nullptr, prof_end = nullptr;
  std::vector<cudaEvent_t> prof_ready, prof_done;
  // per-kernel trace (ddpb_engine_trace_step): one event after every launch, on its stream
  struct TraceRec { char name[48]; int block; int stream; cudaEvent_t ev; int id_lo, id_hi; };
  int trace_mode = 0;                  // 0 eager events, 1 graph + event nodes, 2 graph + in-kernel stamps
  int trace_id_mark = 0;               // trace ids handed out before the current LAUNCH
  unsigned long long* ktrace_buf = nullptr;    // device [2 * kTraceSlots]
  bool trace_on = false;
  bool trace_external = false;         // tracing inside a stream capture: event-record NODES (cudaEventRecordExternal)
  int trace_block = -1;                // conv block the launches belong to (-1: classifier / optimizer)
  size_t trace_n = 0;
  std::vector<TraceRec> trace;         // events are created once and reused
  cudaEvent_t trace_start = nullptr;
  cudaStream_t trace_main = nullptr;
};

static void count_name(Engine* e, const char* call_text) {
  const char* p = call_text;
  if (!strncmp(p, "ddpb_", 5)) p += 5;
  size_t n = 0;
  while (p[n] && p[n] != '(') ++n;
  e->counting_by_name[std::string(p, n)] += 1;
}
static void add_counts(NameCounts& into, const NameCounts& from, long long times = 1) {
  for (const auto& kv : from) into[kv.first] += kv.second * times;
}

static constexpr int kTraceSlots = 512 * 32;      // 512 trace ids x KTRACE_LANES (common.cuh)
extern "C" {
int ddpb_ktrace_set_conv_tc(void*); int ddpb_ktrace_set_wgrad_tc(void*); int ddpb_ktrace_set_gemm_rows_tc(void*);
int ddpb_ktrace_set_bn_relu_pool(void*); int ddpb_ktrace_set_fc_ce(void*); int ddpb_ktrace_set_elementwise(void*);
int ddpb_ktrace_set_sync_p2p(void*); int ddpb_ktrace_set_first_layer_bwd(void*); int ddpb_ktrace_set_conv_tf32(void*);
}

static int trace_mark(Engine* e, cudaStream_t stream, const char* call_text) {
  if (e->trace_n == e->trace.size()) {
    Engine::TraceRec r;
    memset(&r, 0, sizeof(r));
    DDPB_CUDA_CHECK(cudaEventCreate(&r.ev));
    e->trace.push_back(r);
  }
  Engine::TraceRec& r = e->trace[e->trace_n++];
  const char* p = call_text;
  if (!strncmp(p, "ddpb_", 5)) p += 5;
  size_t n = 0;
  while (p[n] && p[n] != '(' && n + 1 < sizeof(r.name)) { r.name[n] = p[n]; ++n; }
  r.name[n] = 0;
  r.block = e->trace_block;
  r.stream = stream == e->trace_main ? 0 : (stream == e->wg ? 1 : (stream == e->comm ? 2 : (stream == e->comm2 ? 3 : 4)));
  if (e->trace_mode == 2) {            // the kernels stamp themselves: remember which trace ids this call used
    r.id_lo = e->trace_id_mark + 1; r.id_hi = ktrace_last_id();
    e->trace_id_mark = r.id_hi;
    return DDPB_OK;
  }
  if (e->trace_external) DDPB_CUDA_CHECK(cudaEventRecordWithFlags(r.ev, stream, cudaEventRecordExternal));
  else DDPB_CUDA_CHECK(cudaEventRecord(r.ev, stream));
  return DDPB_OK;
}

#define ECHECK(call)        \
  do {                      \
    int _rc = (call);       \
    if (_rc) return _rc;    \
  } while (0)

static void peer_table(Engine* e, size_t off, void** out) {
  for (int r = 0; r < e->cfg.world; ++r) {
    void* base = (r == e->cfg.rank) ? e->arena : e->peer_arena[r];
    out[r] = base ? (void*)((char*)base + off) : nullptr;
  }
}

static Slot take(long long& cursor, long long numel) {
  Slot s{cursor, numel};
  cursor = round_up(cursor + numel, kAlign);
  return s;
}

static int plan_layout(Engine* e) {
  const auto& c = e->cfg;
  e->blocks.resize(c.num_blocks);
  long long cur = 0, stat = 0, ch = 0;
  // gradient-ready order: classifier first (bias, weight), then blocks last -> first with the BN
  // pair before the conv pair and weight before bias
  e->fc_b = take(cur, 10);
  e->fc_w = take(cur, 10 * 512);
  for (int i = c.num_blocks - 1; i >= 0; --i) {
    Block& b = e->blocks[i];
    b.cin = c.cin[i]; b.cout = c.cout[i]; b.hw = c.hw[i]; b.pool = c.pool[i];
    b.bn_w = take(cur, b.cout);
    b.bn_b = take(cur, b.cout);
    b.conv_w = take(cur, (long long)b.cout * 9 * b.cin);
    b.conv_b = take(cur, b.cout);
  }
  for (int i = 0; i < c.num_blocks; ++i) {
    e->blocks[i].stat_off = stat; stat += 2 * e->blocks[i].cout;
    e->blocks[i].ch_off = ch; ch += e->blocks[i].cout;
  }
  e->arena_elems = cur;
  e->total_c = ch;
  return DDPB_OK;
}

// Buckets over the ready-ordered arena.  plan 0 = size-capped greedy (torch DDP's rule: a bucket
// closes once it reaches its cap; first cap, then the general cap), 1 = one bucket per layer
// (classifier separate), 2 = a single bucket.
static void plan_buckets(Engine* e) {
  const auto& c = e->cfg;
  struct Item { long long off, end; int block; };
  std::vector<Item> items;
  auto add = [&](Slot s, int blk) { items.push_back({s.off, round_up(s.off + s.numel, kAlign), blk}); };
  add(e->fc_b, -1); add(e->fc_w, -1);
  for (int i = c.num_blocks - 1; i >= 0; --i) {
    const Block& b = e->blocks[i];
    add(b.bn_w, i); add(b.bn_b, i); add(b.conv_w, i); add(b.conv_b, i);
  }
  e->buckets.clear();
  auto close = [&](long long off, long long end, int last) { e->buckets.push_back({off, end - off, last, 0, 0, 1}); };
  if (c.bucket_plan == 2) {
    close(0, e->arena_elems, 0);
  } else if (c.bucket_plan == 1) {
    size_t i = 0;
    while (i < items.size()) {
      size_t j = i;
      while (j + 1 < items.size() && items[j + 1].block == items[i].block) ++j;
      close(items[i].off, items[j].end, items[i].block);
      i = j + 1;
    }
  } else {
    const double mib = 1024.0 * 1024.0;
    double limit = c.first_bucket_mb * mib, acc = 0;
    long long start = 0;
    for (size_t i = 0; i < items.size(); ++i) {
      acc += (double)(items[i].end - items[i].off) * 4.0;
      bool last_item = i + 1 == items.size();
      // a bucket may only close at a layer boundary here: all four tensors of a layer are produced
      // by the same group of kernels, so splitting them buys nothing
      bool boundary = last_item || items[i + 1].block != items[i].block;
      if ((acc >= limit && boundary) || last_item) {
        close(start, items[i].end, items[i].block);
        start = items[i].end; acc = 0; limit = c.bucket_cap_mb * mib;
      }
    }
  }
  // Sub-bucket hand-off (per-layer plan): a large layer's bucket is cut at output-channel boundaries of its conv
  // weight ([co][3][3][ci] in the arena, so a channel range is contiguous); backward() launches that layer's
  // weight gradient once per range and hands each part to the communication streams as soon as it exists, so the
  // reduce of the first half runs while the second half is still being computed (DDP cannot go below a bucket).
  if (c.bucket_plan == 1 && e->sub_buckets > 1 && (c.world > 1 || e->sub_bucket_force)) {
    std::vector<Bucket> out;
    for (const Bucket& bk : e->buckets) {
      const int S = e->sub_buckets;
      const Block* blk = bk.last_block >= 0 ? &e->blocks[bk.last_block] : nullptr;
      const bool ok = blk && c.dtype == DDPB_BF16 && c.use_tc && blk->cin % 64 == 0 && blk->cout % (128 * S) == 0 &&
                      blk->conv_w.numel >= e->sub_bucket_min && blk->conv_w.off >= bk.off &&
                      blk->conv_w.off + blk->conv_w.numel <= bk.off + bk.numel;
      if (!ok) { out.push_back(bk); continue; }
      const long long per = blk->conv_w.numel / S;
      long long lo = bk.off;
      for (int p = 0; p < S; ++p) {
        const long long hi = p + 1 == S ? bk.off + bk.numel : blk->conv_w.off + per * (p + 1);
        out.push_back({lo, hi - lo, bk.last_block, 0, p, S});
        lo = hi;
      }
    }
    e->buckets.swap(out);
  }
  for (auto& b : e->buckets)
    b.mode = (c.world > 1 && b.numel * 4 >= 256 * 1024) ? DDPB_P2P_TWOSHOT : DDPB_P2P_ONESHOT;
  // conv weights are read in bf16 only (tensor-core path): their fp32 master is sharded like the momentum
  e->shard.clear();
  if (c.world > 1 && c.dtype == DDPB_BF16 && e->shard_master)
    for (int i = 0; i < c.num_blocks; ++i) {
      const Block& b = e->blocks[i];
      if (b.cin % 64 == 0 && b.cout % 64 == 0 && (int)e->shard.size() < 2 * DDPB_P2P_MAX_SHARD_RANGES) {
        e->shard.push_back(b.conv_w.off);
        e->shard.push_back(round_up(b.conv_w.off + b.conv_w.numel, 8));
      }
    }
}

// Offsets of the sub-buffers inside the symmetric arena (same on every rank).
static void carve_arena(Engine* e) {
  size_t cur = 0;
  auto carve = [&](size_t bytes) { size_t o = cur; cur = (size_t)round_up((long long)(cur + bytes), 1024); return o; };
  e->off_grad = carve(e->arena_elems * 4);
  e->off_master = carve(e->arena_elems * 4);
  e->off_mom = carve(e->arena_elems * 4);
  e->off_shadow = carve(e->arena_elems * 2);
  e->off_rstats = carve(e->total_c * 2 * 4);
  e->off_flags = carve((size_t)DDPB_P2P_MAX_SLOTS * 2 * DDPB_P2P_MAX_BLOCKS * DDPB_MAX_RANKS * 4);
  e->arena_bytes = cur;
}

static int alloc_all(Engine* e) {
  const auto& c = e->cfg;
  const long long B = c.max_batch;
  e->esize = c.dtype == DDPB_BF16 ? 2 : 4;
  // ---- symmetric arena ----
  carve_arena(e);
  if (c.external_arena) {
    // the caller owns the arena (e.g. cuMemCreate memory bound to an NVSwitch multicast object)
    DDPB_REQUIRE(c.external_arena_bytes >= e->arena_bytes, "engine: external arena too small (%zu < %zu)",
                 (size_t)c.external_arena_bytes, e->arena_bytes);
    e->arena = c.external_arena;
    e->arena_external = true;
    memset(e->ipc_handle, 0, sizeof(e->ipc_handle));
    DDPB_CUDA_CHECK(cudaMemset(e->arena, 0, e->arena_bytes));
  } else {
    ECHECK(ddpb_symm_alloc(e->arena_bytes, &e->arena, e->ipc_handle));
  }
  char* a = (char*)e->arena;
  e->grad = (float*)(a + e->off_grad); e->master = (float*)(a + e->off_master);
  e->mom = (float*)(a + e->off_mom); e->shadow = (bf16*)(a + e->off_shadow);
  e->rmean = (float*)(a + e->off_rstats); e->rvar = e->rmean + e->total_c;
  e->flags = (uint32_t*)(a + e->off_flags);
  // ---- local ----
  size_t nep = (size_t)DDPB_P2P_MAX_SLOTS * 2 * DDPB_P2P_MAX_BLOCKS * 4;
  DDPB_CUDA_CHECK(cudaMalloc(&e->epochs, nep));
  DDPB_CUDA_CHECK(cudaMemset(e->epochs, 0, nep));
  // scratch: [fwd stats 2*sumC][bwd sums 2*sumC][loss][pad]
  e->scratch_floats = (size_t)e->total_c * 4 + 64 + 64;     // ... [loss, pad 32][grid-barrier counters, one per block]
  DDPB_CUDA_CHECK(cudaMalloc(&e->scratch, e->scratch_floats * 4));
  e->loss = e->scratch + e->total_c * 4;
  DDPB_CUDA_CHECK(cudaMalloc(&e->mean, e->total_c * 4));
  DDPB_CUDA_CHECK(cudaMalloc(&e->invstd, e->total_c * 4));
  DDPB_CUDA_CHECK(cudaMalloc(&e->nbt, sizeof(long long) * (c.num_blocks + 1)));
  DDPB_CUDA_CHECK(cudaMemset(e->nbt, 0, sizeof(long long) * (c.num_blocks + 1)));
  DDPB_CUDA_CHECK(cudaMalloc(&e->logits, B * 10 * 4));
  DDPB_CUDA_CHECK(cudaMalloc(&e->dlogits, B * 10 * 4));
  DDPB_CUDA_CHECK(cudaMalloc(&e->eval_loss, 64));
  e->correct = (int*)(e->eval_loss + 1);
  DDPB_CUDA_CHECK(cudaMemset(e->eval_loss, 0, 64));
  DDPB_CUDA_CHECK(cudaMalloc(&e->input, B * 32 * 32 * 3 * e->esize));
  DDPB_CUDA_CHECK(cudaMalloc(&e->target, B * 8));
  if (c.dtype == DDPB_BF16 && c.use_tc && e->blocks[0].cin == 3 && e->blocks[0].cout % 64 == 0) {
    DDPB_CUDA_CHECK(cudaMalloc(&e->x0cols, B * 32 * 32 * 64 * 2));
    DDPB_CUDA_CHECK(cudaMemset(e->x0cols, 0, B * 32 * 32 * 64 * 2));     // columns 32..63 stay zero for good
    DDPB_CUDA_CHECK(cudaMalloc(&e->w0pad, (size_t)e->blocks[0].cout * 64 * 2));
  }
  long long max_y = 0, max_act = 512;
  for (auto& b : e->blocks) {
    long long y = (long long)b.hw * b.hw * b.cout;
    long long o = b.pool ? y / 4 : y;
    max_y = std::max(max_y, y); max_act = std::max(max_act, o);
    DDPB_CUDA_CHECK(cudaMalloc(&b.y, B * y * e->esize));
    DDPB_CUDA_CHECK(cudaMalloc(&b.act, B * o * e->esize));
    if (b.pool && c.dtype == DDPB_BF16 && c.use_tc && e->fuse_bn_reduce) DDPB_CUDA_CHECK(cudaMalloc(&b.win, B * o * e->esize));
  }
  DDPB_CUDA_CHECK(cudaMalloc(&e->dy_buf2[0], B * max_y * e->esize));
  DDPB_CUDA_CHECK(cudaMalloc(&e->dy_buf2[1], B * max_y * e->esize));
  DDPB_CUDA_CHECK(cudaMalloc(&e->dact, B * max_act * e->esize));
  if (getenv("DDPB_POISON")) {
    // debugging aid: every buffer that is meant to be fully written before it is read starts as NaN
    // (0xFF bytes), so a kernel that reads what nobody wrote shows up as NaN instead of as leftovers
    for (auto& b : e->blocks) {
      long long y = (long long)b.hw * b.hw * b.cout, o = b.pool ? y / 4 : y;
      DDPB_CUDA_CHECK(cudaMemset(b.y, 0xFF, B * y * e->esize));
      DDPB_CUDA_CHECK(cudaMemset(b.act, 0xFF, B * o * e->esize));
      if (b.win) DDPB_CUDA_CHECK(cudaMemset(b.win, 0xFF, B * o * e->esize));
    }
    DDPB_CUDA_CHECK(cudaMemset(e->dy_buf2[0], 0xFF, B * max_y * e->esize));
    DDPB_CUDA_CHECK(cudaMemset(e->dy_buf2[1], 0xFF, B * max_y * e->esize));
    DDPB_CUDA_CHECK(cudaMemset(e->dact, 0xFF, B * max_act * e->esize));
    DDPB_CUDA_CHECK(cudaMemset(e->logits, 0xFF, B * 10 * 4));
    DDPB_CUDA_CHECK(cudaMemset(e->dlogits, 0xFF, B * 10 * 4));
    DDPB_CUDA_CHECK(cudaMemset(e->input, 0xFF, B * 32 * 32 * 3 * e->esize));
    if (e->w0pad) DDPB_CUDA_CHECK(cudaMemset(e->w0pad, 0xFF, (size_t)e->blocks[0].cout * 64 * 2));
  }
  // BN defaults: gamma 1, running_var 1 (weights are loaded by the caller afterwards)
  DDPB_CUDA_CHECK(cudaMemset(e->mean, 0, e->total_c * 4));
  DDPB_CUDA_CHECK(cudaMemset(e->invstd, 0, e->total_c * 4));
  // the bucket kernels are few, small CTAs on the step's critical tail: give them scheduling priority
  int prio_lo = 0, prio_hi = 0;
  DDPB_CUDA_CHECK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
  // (DDPB_PRIO_COMM / DDPB_PRIO_WG override, clamped to the device's range; the caller's stream is the engine's
  // main stream: cs744_ddp_b200/engine.py creates it, DDPB_PRIO_MAIN)
  auto prio_env = [&](const char* name, int dflt) {
    const char* v = getenv(name);
    int p = v ? atoi(v) : dflt;
    if (p < prio_hi) p = prio_hi;
    if (p > prio_lo) p = prio_lo;
    return p;
  };
  const int prio_comm = prio_env("DDPB_PRIO_COMM", prio_hi), prio_wg = prio_env("DDPB_PRIO_WG", prio_lo);
  DDPB_CUDA_CHECK(cudaStreamCreateWithPriority(&e->comm, cudaStreamNonBlocking, prio_comm));
  DDPB_CUDA_CHECK(cudaStreamCreateWithPriority(&e->comm2, cudaStreamNonBlocking, prio_comm));
  DDPB_CUDA_CHECK(cudaStreamCreateWithPriority(&e->wg, cudaStreamNonBlocking, prio_wg));
  e->ev_dy.resize(c.num_blocks + 1); e->ev_dg.resize(c.num_blocks + 1); e->ev_wg.resize(c.num_blocks + 1);
  for (int i = 0; i <= c.num_blocks; ++i) {
    DDPB_CUDA_CHECK(cudaEventCreateWithFlags(&e->ev_dy[i], cudaEventDisableTiming));
    DDPB_CUDA_CHECK(cudaEventCreateWithFlags(&e->ev_dg[i], cudaEventDisableTiming));
    DDPB_CUDA_CHECK(cudaEventCreateWithFlags(&e->ev_wg[i], cudaEventDisableTiming));
  }
  e->ev_part.resize((size_t)c.num_blocks * 4);
  for (auto& ev : e->ev_part) DDPB_CUDA_CHECK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  e->bucket_ev.resize(e->buckets.size());
  for (auto& ev : e->bucket_ev) DDPB_CUDA_CHECK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  DDPB_CUDA_CHECK(cudaEventCreateWithFlags(&e->join_ev, cudaEventDisableTiming));
  DDPB_CUDA_CHECK(cudaEventCreateWithFlags(&e->fork_ev, cudaEventDisableTiming));
  DDPB_CUDA_CHECK(cudaEventCreateWithFlags(&e->zero_ev, cudaEventDisableTiming));
  DDPB_CUDA_CHECK(cudaEventCreateWithFlags(&e->start_ev, cudaEventDisableTiming));
  return DDPB_OK;
}

// ---------------------------------------------------------------------------------------------
// forward / backward sequences
// ---------------------------------------------------------------------------------------------
// forward / dgrad tiles need Cin, Cout multiples of 64; the wgrad tile is 128 output channels wide
static bool use_tc(const Engine* e, const Block& b) {
  return e->cfg.dtype == DDPB_BF16 && e->cfg.use_tc && b.cin % 64 == 0 && b.cout % 64 == 0;
}
// fp32 engine: the same layers on tcgen05 kind::tf32 (conv_tf32.cu); DDPB_TF32=0 keeps the SIMT oracle kernels
static bool use_tf32(const Engine* e, const Block& b) {
  return e->cfg.dtype == DDPB_F32 && e->cfg.use_tc && e->tf32 && b.cin % 64 == 0 && b.cout % 64 == 0;
}
static bool use_tc_wgrad(const Engine* e, const Block& b) {
  return use_tc(e, b) && b.cout % 128 == 0;
}
// the 3-channel first layer: im2col to a [pixels][64] matrix, then 1x1-conv forward and GEMM wgrad
static bool use_tc_first(const Engine* e, const Block& b) {
  return e->cfg.dtype == DDPB_BF16 && e->cfg.use_tc && b.cin == 3 && b.cout % 64 == 0 && b.hw == 32 && e->x0cols;
}

// Every kernel launch of the step goes through here: counted (the launch count bench.py reports) and,
// while a trace is being taken (ddpb_engine_trace_step), followed by an event on the launching stream.
#define LAUNCH(e, stream, call)                                   \
  do {                                                            \
    ECHECK(call);                                                 \
    (e)->counting += 1;                                           \
    count_name((e), #call);                                       \
    if ((e)->trace_on) ECHECK(trace_mark((e), (stream), #call));  \
  } while (0)


// zero_grads: also clear the gradient arena, on the weight-gradient stream, so that the 37 MB memset
// overlaps the forward pass instead of opening backward (only when forward and backward are issued as ONE
// unit - a whole step - because the fork is joined by backward()).
static int forward(Engine* e, int B, int training, cudaStream_t s, bool zero_grads = false) {
  const int dt = e->cfg.dtype;
  float* fstats = e->scratch;
  if (training) {
    DDPB_CUDA_CHECK(cudaMemsetAsync(e->scratch, 0, e->scratch_floats * 4, s));
  }
  if (zero_grads) {
    // everything of the previous step (all streams were joined into s) precedes this point
    DDPB_CUDA_CHECK(cudaEventRecord(e->start_ev, s));
    DDPB_CUDA_CHECK(cudaStreamWaitEvent(e->wg, e->start_ev, 0));
    DDPB_CUDA_CHECK(cudaMemsetAsync(e->grad, 0, e->arena_elems * 4, e->wg));
    DDPB_CUDA_CHECK(cudaEventRecord(e->zero_ev, e->wg));
  }
  const void* x = e->input;
  for (size_t i = 0; i < e->blocks.size(); ++i) {
    Block& b = e->blocks[i];
    e->trace_block = (int)i;
    float* st = training ? fstats + b.stat_off : nullptr;
    const float* bias = e->master + b.conv_b.off;
    int bn_fused = 0;
    if (use_tc(e, b) && training && e->fuse_bn_fwd) {
      // conv + BatchNorm + ReLU + pool in one kernel when the grid fits on the GPU at once
      unsigned int* barrier = reinterpret_cast<unsigned int*>(e->scratch + e->total_c * 4 + 32) + i;
      LAUNCH(e, s, ddpb_conv3x3_fwd_bn_tc(x, e->shadow + b.conv_w.off, bias, b.y, st, B, b.hw, b.hw, b.cin, b.cout, 0,
                                       e->master + b.bn_w.off, e->master + b.bn_b.off, b.act, b.win,
                                       e->mean + b.ch_off, e->invstd + b.ch_off, e->rmean + b.ch_off,
                                       e->rvar + b.ch_off, e->nbt + i, b.pool, 1e-5f, 0.1f, barrier, &bn_fused, s));
    } else if (use_tc(e, b))
      LAUNCH(e, s, ddpb_conv3x3_fwd_tc(x, e->shadow + b.conv_w.off, bias, b.y, st, B, b.hw, b.hw, b.cin, b.cout, 0, s));
    else if (use_tf32(e, b))
      LAUNCH(e, s, ddpb_conv3x3_fwd_tf32((const float*)x, e->master + b.conv_w.off, bias, (float*)b.y, st, B, b.hw, b.hw, b.cin,
                                         b.cout, s));
    else if (use_tc_first(e, b)) {
      LAUNCH(e, s, ddpb_im2col3x3_c3(x, e->shadow + b.conv_w.off, e->x0cols, e->w0pad, B, b.hw, b.hw, b.cout, 0, s));
      LAUNCH(e, s, ddpb_gemm_rows_fwd_tc(e->x0cols, e->w0pad, bias, b.y, st, (long long)B * b.hw * b.hw, b.cout, s));
    } else {
      LAUNCH(e, s, ddpb_conv3x3_fwd_simt(x, e->master + b.conv_w.off, bias, b.y, st, B, b.hw, b.hw, b.cin, b.cout, dt, s));
      if (st && !(b.cin == 3 && b.cout == 64)) e->counting += 1;       // separate channel_stats launch
    }
    if (!bn_fused)
    LAUNCH(e, s, ddpb_bn_relu_pool_fwd_winner(b.y, st, e->master + b.bn_w.off, e->master + b.bn_b.off, b.act,
                                           e->mean + b.ch_off, e->invstd + b.ch_off, e->rmean + b.ch_off,
                                           e->rvar + b.ch_off, e->nbt + i, training ? b.win : nullptr, B, b.hw, b.hw,
                                           b.cout, b.pool, training, 1e-5f, 0.1f, dt, s));
    x = b.act;
  }
  e->trace_block = -1;
  return DDPB_OK;
}

static int loss_fwd(Engine* e, int B, int training, cudaStream_t s) {
  const Block& last = e->blocks.back();
  // training: the loss kernel also produces the gradient of the classifier's input (DDPB_FUSE_FC_DA=0: fc_bwd does)
  LAUNCH(e, s, ddpb_fc_ce_fwd_dact(last.act, e->master + e->fc_w.off, e->master + e->fc_b.off, e->target, e->logits,
                                training ? e->dlogits : nullptr, training ? e->loss : e->eval_loss,
                                training ? nullptr : e->correct, (training && e->fuse_fc_da) ? e->dact : nullptr, B, B,
                                e->cfg.dtype, s));
  return DDPB_OK;
}

static bool overlapped(const Engine* e);
// how many communication streams the step's bucket kernels alternate between
static int bucket_streams(const Engine* e) {
  return (e->cfg.world > 1 && e->peers_set && e->two_comm && overlapped(e)) ? 2 : 1;
}

static int launch_bucket(Engine* e, int bi, cudaStream_t s) {
  const auto& c = e->cfg;
  const Bucket& b = e->buckets[bi];
  if (c.world == 1 || !e->peers_set) {
    // single process (or NCCL fallback driven from Python): plain fused SGD on the bucket
    if (c.external_sync) return DDPB_OK;
    LAUNCH(e, s, ddpb_sgd_step(e->master + b.off, e->mom + b.off, e->grad + b.off,
                            c.dtype == DDPB_BF16 ? (void*)(e->shadow + b.off) : nullptr, b.numel, c.lr, c.momentum,
                            c.wd, 1.f, s));
    return DDPB_OK;
  }
  void *g[DDPB_MAX_RANKS], *m[DDPB_MAX_RANKS], *sh[DDPB_MAX_RANKS], *f[DDPB_MAX_RANKS];
  peer_table(e, e->off_grad, g); peer_table(e, e->off_master, m);
  peer_table(e, e->off_shadow, sh); peer_table(e, e->off_flags, f);
  // Only the last bucket kernel of each communication stream needs the closing barrier: it is reached after
  // every earlier bucket kernel of that stream on this rank, so passing it proves that all peers are done
  // reading our gradients of those buckets and that all their weight pushes into our arena have landed.
  // (Buckets alternate between two streams, see buckets_ready: the last TWO kernels close, and they overlap.)
  const int nb = (int)e->buckets.size();
  const bool last = bi + 1 == nb || (bucket_streams(e) == 2 && bi + 2 == nb);
  const int mode = b.mode | (last ? 0 : DDPB_P2P_NO_EXIT_BARRIER);
  const int blocks = 0;     // the launcher's lean shape: it runs next to the backward kernels
  const long long* sh_tab = e->shard.empty() ? nullptr : e->shard.data();
  const int n_sh = (int)e->shard.size() / 2;
  if (e->mc_arena && (b.mode == DDPB_P2P_TWOSHOT || e->nvls_oneshot)) {
    // NVLS: the switch reduces; big buckets: the owner of a slice updates it and one multicast store
    // publishes the result; small buckets: every rank pulls the reduced sum and updates its own replica
    char* mc = (char*)e->mc_arena;
    LAUNCH(e, s, ddpb_p2p_sync_nvls_sharded(b.mode, mc + e->off_grad, mc + e->off_master,
                                            c.dtype == DDPB_BF16 ? (void*)(mc + e->off_shadow) : nullptr, e->master,
                                            c.dtype == DDPB_BF16 ? (void*)e->shadow : nullptr, f, e->mom, e->epochs,
                                            c.rank, c.world, bi, b.off, b.numel, c.lr, c.momentum, c.wd, 0,
                                            last ? 0 : 1, sh_tab, n_sh, s));
    return DDPB_OK;
  }
  LAUNCH(e, s, ddpb_p2p_sync_sharded(mode, g, m, sh, f, e->mom, e->epochs, c.rank, c.world, bi, b.off, b.numel, c.lr,
                                     c.momentum, c.wd, blocks, sh_tab, n_sh, s));
  return DDPB_OK;
}

// The captured step graph, before it is instantiated.  The bucket kernels (reduce + SGD) must win the CTA
// scheduler whenever an SM has room: they are short, and a peer may be waiting inside the same kernel.  The
// priority of the communication streams is written onto their kernel nodes explicitly here (DDPB_COMM_PRIO=0
// leaves the nodes as captured); DDPB_GRAPH_DUMP=1 lists every kernel node with its priority, grid and the
// kernels it depends on.
static void tune_graph(Engine* e, cudaGraph_t graph) {
  static const bool dump = getenv("DDPB_GRAPH_DUMP") && atoi(getenv("DDPB_GRAPH_DUMP")) != 0;
  static const bool set_prio = !(getenv("DDPB_COMM_PRIO") && atoi(getenv("DDPB_COMM_PRIO")) == 0);
  size_t n = 0;
  if (cudaGraphGetNodes(graph, nullptr, &n) != cudaSuccess || n == 0) { (void)cudaGetLastError(); return; }
  std::vector<cudaGraphNode_t> nodes(n);
  if (cudaGraphGetNodes(graph, nodes.data(), &n) != cudaSuccess) { (void)cudaGetLastError(); return; }
  auto name_of = [](cudaGraphNode_t nd, cudaKernelNodeParams* kp) -> std::string {
    cudaGraphNodeType t;
    if (cudaGraphNodeGetType(nd, &t) != cudaSuccess) { (void)cudaGetLastError(); return "?"; }
    if (t != cudaGraphNodeTypeKernel) return t == cudaGraphNodeTypeMemset ? "(memset)" : "(other)";
    cudaKernelNodeParams local;
    if (!kp) kp = &local;
    if (cudaGraphKernelNodeGetParams(nd, kp) != cudaSuccess) { (void)cudaGetLastError(); return "?"; }
    const char* nm = nullptr;
    if (cudaFuncGetName(&nm, kp->func) != cudaSuccess || !nm) { (void)cudaGetLastError(); return "?"; }
    return nm;
  };
  int lo = 0, hi = 0;
  cudaDeviceGetStreamPriorityRange(&lo, &hi);
  for (size_t i = 0; i < n; ++i) {
    cudaKernelNodeParams kp;
    memset(&kp, 0, sizeof(kp));
    const std::string nm = name_of(nodes[i], &kp);
    if (nm[0] == '(' || nm[0] == '?') continue;
    const bool comm = nm.find("p2p_") != std::string::npos || nm.find("sgd_") != std::string::npos;
    cudaKernelNodeAttrValue v;
    memset(&v, 0, sizeof(v));
    int before = 0;
    if (cudaGraphKernelNodeGetAttribute(nodes[i], cudaKernelNodeAttributePriority, &v) == cudaSuccess) before = v.priority;
    else (void)cudaGetLastError();
    if (comm && set_prio && e->cfg.world > 1) {
      v.priority = hi;
      if (cudaGraphKernelNodeSetAttribute(nodes[i], cudaKernelNodeAttributePriority, &v) != cudaSuccess) (void)cudaGetLastError();
    }
    if (dump) {
      int after = before;
      if (cudaGraphKernelNodeGetAttribute(nodes[i], cudaKernelNodeAttributePriority, &v) == cudaSuccess) after = v.priority;
      else (void)cudaGetLastError();
      fprintf(stderr, "[ddpb graph] %-3zu %-40.40s grid %5u x %-3u prio %d -> %d   after:", i, nm.c_str() + (nm.compare(0, 8, "_ZN4ddpb") == 0 ? 10 : 0),
              kp.gridDim.x, kp.blockDim.x, before, after);
      size_t nd = 0;
      if (cudaGraphNodeGetDependencies_v2(nodes[i], nullptr, nullptr, &nd) == cudaSuccess && nd) {
        std::vector<cudaGraphNode_t> deps(nd);
        std::vector<cudaGraphEdgeData> ed(nd);
        if (cudaGraphNodeGetDependencies_v2(nodes[i], deps.data(), ed.data(), &nd) == cudaSuccess)
          for (size_t k = 0; k < nd; ++k) {
            size_t idx = 0;
            while (idx < n && nodes[idx] != deps[k]) ++idx;
            fprintf(stderr, " %zu%s", idx, ed[k].type == cudaGraphDependencyTypeProgrammatic ? "p" : "");
          }
        else (void)cudaGetLastError();
      } else (void)cudaGetLastError();
      fprintf(stderr, "\n");
    }
  }
}

// Is the optimizer / gradient-sync work overlapped with backward?  The bucketed strategy by
// definition; a single process as well (its optimizer step has no cross-rank dependency, so stepping
// a layer as soon as its gradients are final is observably identical to stepping after backward).
static bool overlapped(const Engine* e) {
  const auto& c = e->cfg;
  if (c.external_sync) return false;
  return c.strategy == 3 || c.world == 1;
}

// into how many output-channel parts backward() cuts the weight gradient of block `blk` (see plan_buckets)
static int layer_parts(const Engine* e, int blk) {
  for (const Bucket& b : e->buckets)
    if (b.last_block == blk) return b.parts;
  return 1;
}

// Block `blk` (-1 = classifier) is complete: its gradients have been produced (event ev_a) and every
// reader of its current weights has run (event ev_b).  Hand each bucket that ends with this block to
// the communication stream, where its fused reduce + optimizer kernel overlaps the rest of backward.
static int buckets_ready(Engine* e, int blk, cudaEvent_t ev_a, cudaEvent_t ev_b, int part = -1) {
  if (!overlapped(e)) return DDPB_OK;
  for (size_t i = 0; i < e->buckets.size(); ++i) {
    if (e->buckets[i].last_block != blk) continue;
    if (part >= 0 && e->buckets[i].part != part) continue;
    // Multi-GPU: a bucket kernel is mostly NVLink latency (cross-rank barrier, first loads), so two of
    // them in flight on alternating streams overlap their latencies.  Each stream's last kernel carries the
    // closing barrier for the buckets of its stream (launch_bucket); the step joins both streams at its end.
    cudaStream_t cs = (bucket_streams(e) == 2 && (i & 1)) ? e->comm2 : e->comm;
    if (ev_a) DDPB_CUDA_CHECK(cudaStreamWaitEvent(cs, ev_a, 0));
    if (ev_b) DDPB_CUDA_CHECK(cudaStreamWaitEvent(cs, ev_b, 0));
    if (cs == e->comm2) e->comm2_used = true;
    if (e->prof_on) DDPB_CUDA_CHECK(cudaEventRecord(e->prof_ready[i], cs));
    ECHECK(launch_bucket(e, (int)i, cs));
    if (e->prof_on) DDPB_CUDA_CHECK(cudaEventRecord(e->prof_done[i], cs));
  }
  return DDPB_OK;
}

// Backward runs on three streams:
//   s    (critical path)  fc-bwd -> [BN-bwd(i) -> dgrad(i)] for i = last..0
//   wg   (off the path)   wgrad(i), as soon as BN-bwd(i) has produced dY(i); overlaps the next
//                         layers' BN-bwd / dgrad, which leave most SMs idle on the small layers
//   comm                  the buckets' fused reduce + SGD kernels
// dY is double-buffered so that BN-bwd(i-1) does not overwrite what wgrad(i) is still reading.
static int backward(Engine* e, int B, cudaStream_t s, bool grads_zeroed = false) {
  const auto& c = e->cfg;
  const int dt = c.dtype;
  const int L = (int)e->blocks.size();
  float* bsums = e->scratch + e->total_c * 2;
  if (grads_zeroed) DDPB_CUDA_CHECK(cudaStreamWaitEvent(s, e->zero_ev, 0));    // cleared during forward(), on wg
  else DDPB_CUDA_CHECK(cudaMemsetAsync(e->grad, 0, e->arena_elems * 4, s));
  const Block& last = e->blocks.back();
  if (e->fuse_fc_da && grads_zeroed) {
    // dA came with the loss kernel: the classifier's weight / bias gradient is all that is left of its backward,
    // and it runs on the weight-gradient stream, off the critical path.  (Only in the whole-step graph, where the
    // gradient arena was cleared on that stream: when the memset sits on `s` right above, a normally launched kernel
    // has to follow it - a programmatic launch cannot depend on a memset node - and fc_bwd is that kernel.)
    DDPB_CUDA_CHECK(cudaEventRecord(e->ev_dy[L], s));                 // the loss kernel has run, the gradient arena is zero
    DDPB_CUDA_CHECK(cudaStreamWaitEvent(e->wg, e->ev_dy[L], 0));
    LAUNCH(e, e->wg, ddpb_fc_bwd(last.act, e->master + e->fc_w.off, e->dlogits, nullptr, e->grad + e->fc_w.off,
                              e->grad + e->fc_b.off, B, dt, e->wg));
    DDPB_CUDA_CHECK(cudaEventRecord(e->ev_dg[L], e->wg));
  } else {
    LAUNCH(e, s, ddpb_fc_bwd(last.act, e->master + e->fc_w.off, e->dlogits, e->fuse_fc_da ? nullptr : e->dact,
                          e->grad + e->fc_w.off, e->grad + e->fc_b.off, B, dt, s));
    DDPB_CUDA_CHECK(cudaEventRecord(e->ev_dg[L], s));
  }
  ECHECK(buckets_ready(e, -1, e->ev_dg[L], nullptr));
  bool reduced_by_dgrad = false;
  for (int i = L - 1; i >= 0; --i) {
    Block& b = e->blocks[i];
    e->trace_block = i;
    const void* x_in = i == 0 ? e->input : e->blocks[i - 1].act;
    void* dy = e->dy_buf2[i & 1];
    if (i + 2 < L) DDPB_CUDA_CHECK(cudaStreamWaitEvent(s, e->ev_wg[i + 2], 0));   // previous reader of this dY buffer
    if (i == 0 && reduced_by_dgrad && e->fuse_first_bwd && use_tc_first(e, b) && b.pool && b.cout == 64) {
      // Block 0 has no data gradient: its BatchNorm-backward apply pass and its weight gradient run as ONE kernel
      // and the gradient of its conv output (the largest tensor of the backward pass) never touches HBM
      LAUNCH(e, s, ddpb_first_layer_bwd_fused(e->dact, b.y, e->x0cols, e->mean + b.ch_off, e->invstd + b.ch_off,
                                              e->master + b.bn_w.off, e->master + b.bn_b.off, bsums + b.stat_off,
                                              e->grad + b.conv_w.off, e->grad + b.bn_w.off, e->grad + b.bn_b.off,
                                              e->grad + b.conv_b.off, B, b.hw, b.hw, b.cout, 27, 27, 0, s));
      DDPB_CUDA_CHECK(cudaEventRecord(e->ev_dy[0], s));
      DDPB_CUDA_CHECK(cudaEventRecord(e->ev_dg[0], s));
      DDPB_CUDA_CHECK(cudaEventRecord(e->ev_wg[0], s));
      ECHECK(buckets_ready(e, 0, e->ev_wg[0], nullptr));
      reduced_by_dgrad = false;
      continue;
    }
    if (reduced_by_dgrad) {
      // the reduction pass already ran in the epilogue of dgrad(i + 1)
      LAUNCH(e, s, ddpb_bn_relu_pool_bwd_apply(e->dact, b.y, e->mean + b.ch_off, e->invstd + b.ch_off,
                                            e->master + b.bn_w.off, e->master + b.bn_b.off, bsums + b.stat_off, dy,
                                            e->grad + b.bn_w.off, e->grad + b.bn_b.off, e->grad + b.conv_b.off, B,
                                            b.hw, b.hw, b.cout, b.pool, dt, s));
    } else {
      LAUNCH(e, s, ddpb_bn_relu_pool_bwd(e->dact, b.y, e->mean + b.ch_off, e->invstd + b.ch_off, e->master + b.bn_w.off,
                                      e->master + b.bn_b.off, bsums + b.stat_off, dy, e->grad + b.bn_w.off,
                                      e->grad + b.bn_b.off, e->grad + b.conv_b.off, B, b.hw, b.hw, b.cout, b.pool,
                                      dt, s));
      e->counting += 1;     // the BN backward is two launches (reduce + apply)
    }
    reduced_by_dgrad = false;
    DDPB_CUDA_CHECK(cudaEventRecord(e->ev_dy[i], s));
    // dgrad reads this layer's CURRENT weights; the bucket's fused reduce+SGD rewrites them, so the
    // bucket may only start after dgrad (ev_dg) as well as after wgrad (ev_wg).
    if (i > 0) {
      if (use_tc(e, b) && e->fuse_bn_reduce) {
        // dgrad(i) also reduces the BN-backward sums of block i - 1 (whose output gradient it produces)
        const Block& pb = e->blocks[i - 1];
        LAUNCH(e, s, ddpb_conv3x3_dgrad_bnred_tc(dy, e->shadow + b.conv_w.off, e->dact, B, b.hw, b.hw, b.cin, b.cout, 0,
                                              pb.pool ? pb.win : pb.y, e->mean + pb.ch_off, e->invstd + pb.ch_off,
                                              e->master + pb.bn_w.off, e->master + pb.bn_b.off, bsums + pb.stat_off,
                                              s));
        reduced_by_dgrad = true;
      } else if (use_tc(e, b))
        LAUNCH(e, s, ddpb_conv3x3_dgrad_tc(dy, e->shadow + b.conv_w.off, e->dact, B, b.hw, b.hw, b.cin, b.cout, 0, s));
      else if (use_tf32(e, b))
        LAUNCH(e, s, ddpb_conv3x3_dgrad_tf32((const float*)dy, e->master + b.conv_w.off, (float*)e->dact, B, b.hw, b.hw, b.cin,
                                             b.cout, s));
      else
        LAUNCH(e, s, ddpb_conv3x3_dgrad_simt(dy, e->master + b.conv_w.off, e->dact, B, b.hw, b.hw, b.cin, b.cout, dt, s));
    }
    DDPB_CUDA_CHECK(cudaEventRecord(e->ev_dg[i], s));
    DDPB_CUDA_CHECK(cudaStreamWaitEvent(e->wg, e->ev_dy[i], 0));
    const int parts = overlapped(e) ? layer_parts(e, i) : 1;
    if (parts > 1 && use_tc_wgrad(e, b)) {
      // sub-bucket hand-off: one launch per output-channel range, each part's bucket leaves as soon as it exists
      const int cn = b.cout / parts;
      for (int p = 0; p < parts; ++p) {
        LAUNCH(e, e->wg, ddpb_conv3x3_wgrad_tc_range(dy, x_in, e->grad + b.conv_w.off, B, b.hw, b.hw, b.cin, b.cout, p * cn,
                                                     cn, 0, e->wg));
        cudaEvent_t ev = p + 1 == parts ? e->ev_wg[i] : e->ev_part[(size_t)i * 4 + p];
        DDPB_CUDA_CHECK(cudaEventRecord(ev, e->wg));
        ECHECK(buckets_ready(e, i, ev, e->ev_dg[i], p));
      }
      continue;
    }
    if (use_tc_wgrad(e, b))
      LAUNCH(e, e->wg, ddpb_conv3x3_wgrad_tc(dy, x_in, e->grad + b.conv_w.off, B, b.hw, b.hw, b.cin, b.cout, 0, e->wg));
    else if (use_tf32(e, b) && b.cout % 128 == 0)
      LAUNCH(e, e->wg, ddpb_conv3x3_wgrad_tf32((const float*)dy, (const float*)x_in, e->grad + b.conv_w.off, B, b.hw, b.hw, b.cin,
                                               b.cout, 0, e->wg));
    else if (use_tc_first(e, b))
      LAUNCH(e, e->wg, ddpb_patch_wgrad_tc(dy, e->x0cols, e->grad + b.conv_w.off, B, b.hw, b.hw, b.cout, 27, 27, 0, e->wg));
    else
      LAUNCH(e, e->wg, ddpb_conv3x3_wgrad_simt(dy, x_in, e->grad + b.conv_w.off, B, b.hw, b.hw, b.cin, b.cout, dt, e->wg));
    DDPB_CUDA_CHECK(cudaEventRecord(e->ev_wg[i], e->wg));
    ECHECK(buckets_ready(e, i, e->ev_wg[i], e->ev_dg[i]));
  }
  e->trace_block = -1;
  // join the weight-gradient stream (wgrads of blocks >= 2 were already waited for above)
  DDPB_CUDA_CHECK(cudaStreamWaitEvent(s, e->ev_wg[0], 0));
  if (L > 1) DDPB_CUDA_CHECK(cudaStreamWaitEvent(s, e->ev_wg[1], 0));
  if (e->prof_on) DDPB_CUDA_CHECK(cudaEventRecord(e->prof_bwd, s));   // all gradients of the step exist
  const bool p2p = c.world > 1 && e->peers_set;
  if (overlapped(e)) {
    // join: the step is complete when every bucket kernel has finished
    if (e->comm2_used) {
      DDPB_CUDA_CHECK(cudaEventRecord(e->fork_ev, e->comm2));
      DDPB_CUDA_CHECK(cudaStreamWaitEvent(e->comm, e->fork_ev, 0));
      e->comm2_used = false;
    }
    DDPB_CUDA_CHECK(cudaEventRecord(e->join_ev, e->comm));
    DDPB_CUDA_CHECK(cudaStreamWaitEvent(s, e->join_ev, 0));
  } else if (!c.external_sync) {
    // Parts 2a / 2b: synchronise strictly after backward, like the reference (Part 2a/main.py:94-96)
    if (p2p) {
      void *g[DDPB_MAX_RANKS], *m[DDPB_MAX_RANKS], *sh[DDPB_MAX_RANKS], *f[DDPB_MAX_RANKS];
      peer_table(e, e->off_grad, g); peer_table(e, e->off_master, m);
      peer_table(e, e->off_shadow, sh); peer_table(e, e->off_flags, f);
      // nothing else runs any more: the fat CTA shape on a grid sized by the data (blocks = -1), and the
      // in-switch reduction where the arena is bound to a multicast object
      int mode = c.strategy == 1 ? DDPB_P2P_ROOT : DDPB_P2P_TWOSHOT;
      const long long* sh_tab = e->shard.empty() ? nullptr : e->shard.data();
      const int n_sh = (int)e->shard.size() / 2;
      if (e->mc_arena && e->nvls_after_bwd) {
        char* mc = (char*)e->mc_arena;
        LAUNCH(e, s, ddpb_p2p_sync_nvls_sharded(mode, mc + e->off_grad, mc + e->off_master,
                                                dt == DDPB_BF16 ? (void*)(mc + e->off_shadow) : nullptr, e->master,
                                                dt == DDPB_BF16 ? (void*)e->shadow : nullptr, f, e->mom, e->epochs,
                                                c.rank, c.world, DDPB_P2P_MAX_SLOTS - 1, 0, e->arena_elems, c.lr,
                                                c.momentum, c.wd, -1, 0, sh_tab, n_sh, s));
      } else {
        LAUNCH(e, s, ddpb_p2p_sync_sharded(mode, g, m, sh, f, e->mom, e->epochs, c.rank, c.world, DDPB_P2P_MAX_SLOTS - 1,
                                           0, e->arena_elems, c.lr, c.momentum, c.wd, e->fat_after_bwd ? -1 : 0, sh_tab,
                                           n_sh, s));
      }
    } else {
      LAUNCH(e, s, ddpb_sgd_step(e->master, e->mom, e->grad, dt == DDPB_BF16 ? (void*)e->shadow : nullptr,
                              e->arena_elems, c.lr, c.momentum, c.wd, 1.f, s));
    }
  }
  return DDPB_OK;
}

enum { KIND_FWD = 1, KIND_BWD = 2, KIND_STEP = 3 };

// NVTX ranges around the host-side launch sequences (DDPB_NVTX=1): they show up in nsys / ncu timelines as
// ddpb.forward / ddpb.backward / ddpb.graph_launch (the reference has only wall-clock prints).
static bool nvtx_on() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("DDPB_NVTX"); v = (e && e[0] == '1') ? 1 : 0; }
  return v == 1;
}
struct NvtxRange {
  bool on;
  explicit NvtxRange(const char* name) : on(nvtx_on()) { if (on) nvtxRangePushA(name); }
  ~NvtxRange() { if (on) nvtxRangePop(); }
};

static int run_kind(Engine* e, int kind, int B, cudaStream_t s) {
  const bool whole = kind == KIND_STEP && e->zero_async;
  if (kind & KIND_FWD) {
    NvtxRange r("ddpb.forward");
    ECHECK(forward(e, B, 1, s, whole));
    ECHECK(loss_fwd(e, B, 1, s));
  }
  if (kind & KIND_BWD) {
    NvtxRange r("ddpb.backward+sync+sgd");
    ECHECK(backward(e, B, s, whole));
  }
  return DDPB_OK;
}

static int run_graphed(Engine* e, int kind, int B, cudaStream_t s) {
  if (!e->cfg.use_graph) {
    e->counting = 0; e->counting_by_name.clear();
    ECHECK(run_kind(e, kind, B, s));
    e->launches += e->counting; add_counts(e->launches_by_name, e->counting_by_name);
    return DDPB_OK;
  }
  long long key = (long long)kind * 1000000 + B;
  auto it = e->graphs.find(key);
  if (it == e->graphs.end()) {
    // one eager run first: it performs the one-time cudaFuncSetAttribute calls (not capturable)
    e->counting = 0; e->counting_by_name.clear();
    ECHECK(run_kind(e, kind, B, s));
    e->launches += e->counting; add_counts(e->launches_by_name, e->counting_by_name);
    DDPB_CUDA_CHECK(cudaStreamSynchronize(s));
    DDPB_CUDA_CHECK(cudaStreamSynchronize(e->comm));
    DDPB_CUDA_CHECK(cudaStreamSynchronize(e->wg));
    // NOTE: that eager run was a real step; the caller asked for exactly one, so capture only and
    // launch the graph from the NEXT call on.
    Graph g;
    cudaGraph_t graph = nullptr;
    e->counting = 0; e->counting_by_name.clear();
    cudaError_t ce = cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
    if (ce != cudaSuccess) {
      // e.g. the legacy default stream cannot be captured: keep running eagerly, remember why
      (void)cudaGetLastError();
      set_last_error("CUDA graph capture unavailable (%s); running eagerly", cudaGetErrorString(ce));
      e->cfg.use_graph = 0;
      return DDPB_OK;
    }
    int rc = run_kind(e, kind, B, s);
    ce = cudaStreamEndCapture(s, &graph);
    if (rc || ce != cudaSuccess || !graph) {
      (void)cudaGetLastError();
      if (graph) cudaGraphDestroy(graph);
      e->cfg.use_graph = 0;
      if (rc) return rc;
      set_last_error("CUDA graph capture failed (%s); running eagerly", cudaGetErrorString(ce));
      return DDPB_OK;
    }
    g.launches = e->counting; g.by_name = e->counting_by_name;
    tune_graph(e, graph);
    ce = cudaGraphInstantiate(&g.exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ce != cudaSuccess) {
      (void)cudaGetLastError();
      e->cfg.use_graph = 0;
      set_last_error("cudaGraphInstantiate failed (%s); running eagerly", cudaGetErrorString(ce));
      return DDPB_OK;
    }
    e->graphs[key] = g;
    return DDPB_OK;
  }
  {
    NvtxRange r("ddpb.graph_launch");
    DDPB_CUDA_CHECK(cudaGraphLaunch(it->second.exec, s));
  }
  e->launches += it->second.launches; add_counts(e->launches_by_name, it->second.by_name);
  return DDPB_OK;
}

}  // namespace ddpb

using namespace ddpb;

// Size of the symmetric arena an engine with this configuration needs (for callers that bring their
// own arena through cfg.external_arena).  Allocates nothing.
extern "C" int ddpb_engine_arena_bytes(const ddpb_engine_config* cfg, size_t* bytes) {
  DDPB_REQUIRE(cfg && bytes, "engine_arena_bytes: null argument");
  DDPB_REQUIRE(cfg->num_blocks > 0 && cfg->num_blocks <= DDPB_MAX_BLOCKS, "engine_arena_bytes: bad block count");
  Engine tmp;
  tmp.cfg = *cfg;
  plan_layout(&tmp);
  carve_arena(&tmp);
  *bytes = tmp.arena_bytes;
  return DDPB_OK;
}

extern "C" int ddpb_engine_create(const ddpb_engine_config* cfg, void** handle) {
  DDPB_REQUIRE(cfg && handle, "engine_create: null argument");
  DDPB_REQUIRE(cfg->num_blocks > 0 && cfg->num_blocks <= DDPB_MAX_BLOCKS, "engine_create: bad block count");
  DDPB_REQUIRE(cfg->world >= 1 && cfg->world <= DDPB_MAX_RANKS, "engine_create: world size %d unsupported", cfg->world);
  Engine* e = new Engine();
  e->cfg = *cfg;
  if (const char* fb = getenv("DDPB_FUSE_BN_REDUCE")) e->fuse_bn_reduce = fb[0] != '0';
  if (const char* tc = getenv("DDPB_TWO_COMM")) e->two_comm = tc[0] != '0';
  if (const char* ff = getenv("DDPB_FUSE_BN_FWD")) e->fuse_bn_fwd = ff[0] != '0';
  if (const char* f0 = getenv("DDPB_FUSE_FIRST_BWD")) e->fuse_first_bwd = f0[0] != '0';
  if (const char* tf = getenv("DDPB_TF32")) e->tf32 = tf[0] != '0';
  if (const char* sm = getenv("DDPB_SHARD_MASTER")) e->shard_master = sm[0] != '0';
  if (const char* v = getenv("DDPB_FUSE_FC_DA")) e->fuse_fc_da = v[0] != '0';
  if (const char* v = getenv("DDPB_SUB_BUCKETS")) { const int n = atoi(v); e->sub_buckets = (n == 2 || n == 4) ? n : 1; }
  if (const char* v = getenv("DDPB_SUB_BUCKET_MIN")) { const long long n = atoll(v); if (n > 0) e->sub_bucket_min = n; }
  if (const char* v = getenv("DDPB_SUB_BUCKETS_FORCE")) e->sub_bucket_force = v[0] == '1';
  if (const char* v = getenv("DDPB_NVLS_AFTER_BWD")) e->nvls_after_bwd = v[0] != '0';
  if (const char* v = getenv("DDPB_FAT_AFTER_BWD")) e->fat_after_bwd = v[0] != '0';
  if (const char* za = getenv("DDPB_ZERO_ASYNC")) e->zero_async = za[0] != '0';
  if (const char* no = getenv("DDPB_NVLS_ONESHOT")) e->nvls_oneshot = no[0] != '0';
  int rc = plan_layout(e);
  if (!rc) { plan_buckets(e); rc = alloc_all(e); }
  if (rc) { delete e; return rc; }
  DDPB_REQUIRE((int)e->buckets.size() < DDPB_P2P_MAX_SLOTS - 1, "engine_create: too many buckets");
  *handle = e;
  return DDPB_OK;
}

extern "C" int ddpb_engine_destroy(void* handle) {
  Engine* e = (Engine*)handle;
  if (!e) return DDPB_OK;
  cudaDeviceSynchronize();
  for (auto& kv : e->graphs) if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
  for (int r = 0; r < e->cfg.world; ++r)
    if (r != e->cfg.rank && e->peer_arena[r] && !e->arena_external) cudaIpcCloseMemHandle(e->peer_arena[r]);
  for (auto& b : e->blocks) { cudaFree(b.y); cudaFree(b.act); if (b.win) cudaFree(b.win); }
  if (e->ktrace_buf) cudaFree(e->ktrace_buf);
  void* frees[] = {e->epochs, e->scratch, e->mean, e->invstd, e->nbt, e->logits, e->dlogits, e->eval_loss,
                   e->input, e->target, e->dy_buf2[0], e->dy_buf2[1], e->dact, e->x0cols, e->w0pad,
                   e->arena_external ? nullptr : e->arena};
  for (void* p : frees) if (p) cudaFree(p);
  for (auto& ev : e->bucket_ev) cudaEventDestroy(ev);
  if (e->join_ev) cudaEventDestroy(e->join_ev);
  if (e->fork_ev) cudaEventDestroy(e->fork_ev);
  if (e->zero_ev) cudaEventDestroy(e->zero_ev);
  if (e->start_ev) cudaEventDestroy(e->start_ev);
  for (auto& ev : e->ev_dy) cudaEventDestroy(ev);
  for (auto& ev : e->ev_dg) cudaEventDestroy(ev);
  for (auto& ev : e->ev_wg) cudaEventDestroy(ev);
  for (auto& ev : e->ev_part) cudaEventDestroy(ev);
  if (e->comm) cudaStreamDestroy(e->comm);
  if (e->comm2) cudaStreamDestroy(e->comm2);
  if (e->wg) cudaStreamDestroy(e->wg);
  delete e;
  return DDPB_OK;
}

extern "C" int ddpb_engine_info(void* handle, ddpb_engine_info_t* out) {
  Engine* e = (Engine*)handle;
  DDPB_REQUIRE(e && out, "engine_info: null argument");
  memset(out, 0, sizeof(*out));
  out->arena = e->arena; out->arena_bytes = e->arena_bytes;
  memcpy(out->ipc_handle, e->ipc_handle, 64);
  out->arena_elems = e->arena_elems; out->total_channels = e->total_c;
  out->grad = e->grad; out->master = e->master; out->mom = e->mom; out->shadow = e->shadow;
  out->running_mean = e->rmean; out->running_var = e->rvar; out->num_batches_tracked = e->nbt;
  out->input = e->input; out->target = e->target; out->loss = e->loss; out->eval_loss = e->eval_loss;
  out->correct = e->correct; out->logits = e->logits;
  out->num_buckets = (int)e->buckets.size();
  for (size_t i = 0; i < e->buckets.size() && i < DDPB_P2P_MAX_SLOTS; ++i) {
    out->bucket_off[i] = e->buckets[i].off; out->bucket_numel[i] = e->buckets[i].numel;
    out->bucket_mode[i] = e->buckets[i].mode; out->bucket_last_block[i] = e->buckets[i].last_block;
  }
  out->fc_w_off = e->fc_w.off; out->fc_b_off = e->fc_b.off;
  for (size_t i = 0; i < e->blocks.size(); ++i) {
    const Block& b = e->blocks[i];
    out->conv_w_off[i] = b.conv_w.off; out->conv_b_off[i] = b.conv_b.off;
    out->bn_w_off[i] = b.bn_w.off; out->bn_b_off[i] = b.bn_b.off; out->ch_off[i] = b.ch_off;
    out->y[i] = b.y; out->act[i] = b.act;
  }
  out->multicast = e->mc_arena != nullptr;
  out->launches = e->launches;
  out->graph_active = e->cfg.use_graph && !e->graphs.empty();
  return DDPB_OK;
}

// "name=count;name=count;..." of every entry point launched since the engine was created (graph replays
// included): the driver-visible proof of which sync kernels a multi-GPU run really used.  Returns the
// length needed (including the terminating NUL); writes at most cap bytes.
extern "C" int ddpb_engine_launch_counts(void* handle, char* buf, int cap) {
  Engine* e = (Engine*)handle;
  if (!e) return 0;
  std::string out;
  for (const auto& kv : e->launches_by_name) {
    if (!out.empty()) out += ';';
    out += kv.first + "=" + std::to_string(kv.second);
  }
  if (buf && cap > 0) {
    const size_t n = out.size() < (size_t)cap - 1 ? out.size() : (size_t)cap - 1;
    memcpy(buf, out.data(), n);
    buf[n] = 0;
  }
  return (int)out.size() + 1;
}

// peer_arena[r] = this rank's mapping of rank r's arena (ddpb_symm_open), entry `rank` ignored.
extern "C" int ddpb_engine_set_peers(void* handle, void* const* peer_arena) {
  Engine* e = (Engine*)handle;
  DDPB_REQUIRE(e && peer_arena, "engine_set_peers: null argument");
  for (int r = 0; r < e->cfg.world; ++r) e->peer_arena[r] = r == e->cfg.rank ? e->arena : peer_arena[r];
  e->peers_set = true;
  return DDPB_OK;
}

// Multicast (NVLS) mapping of this rank's arena: same layout, every store reaches all ranks, every
// multimem.ld_reduce returns the sum over all ranks.  Pass null to go back to the unicast kernels.
extern "C" int ddpb_engine_set_multicast(void* handle, void* mc_arena) {
  Engine* e = (Engine*)handle;
  DDPB_REQUIRE(e, "engine_set_multicast: null handle");
  e->mc_arena = mc_arena;
  for (auto& kv : e->graphs) if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
  e->graphs.clear();
  return DDPB_OK;
}

extern "C" int ddpb_engine_set_lr(void* handle, float lr) {
  Engine* e = (Engine*)handle;
  DDPB_REQUIRE(e, "engine_set_lr: null");
  if (lr != e->cfg.lr) {       // baked into captured graphs: drop them
    e->cfg.lr = lr;
    for (auto& kv : e->graphs) if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    e->graphs.clear();
  }
  return DDPB_OK;
}

// Refresh the bf16 compute copy from the fp32 master weights (after loading a checkpoint).
extern "C" int ddpb_engine_refresh_shadow(void* handle, cudaStream_t s) {
  Engine* e = (Engine*)handle;
  return ddpb_cast_f32_bf16(e->master, e->shadow, e->arena_elems, s);
}

extern "C" int ddpb_engine_forward(void* handle, int batch, cudaStream_t s) {
  Engine* e = (Engine*)handle;
  DDPB_REQUIRE(batch > 0 && batch <= e->cfg.max_batch, "engine_forward: batch %d out of range", batch);
  return run_graphed(e, KIND_FWD, batch, s);
}
extern "C" int ddpb_engine_backward(void* handle, int batch, cudaStream_t s) {
  Engine* e = (Engine*)handle;
  DDPB_REQUIRE(batch > 0 && batch <= e->cfg.max_batch, "engine_backward: batch %d out of range", batch);
  return run_graphed(e, KIND_BWD, batch, s);
}
extern "C" int ddpb_engine_train_step(void* handle, int batch, cudaStream_t s) {
  Engine* e = (Engine*)handle;
  DDPB_REQUIRE(batch > 0 && batch <= e->cfg.max_batch, "engine_train_step: batch %d out of range", batch);
  return run_graphed(e, KIND_STEP, batch, s);
}

// One eager (non-graph) training step instrumented with CUDA events.  out[0] = forward end,
// out[1] = last compute kernel of backward, out[2] = step end (after the gradient sync joined),
// then per bucket i: out[3+2i] = its gradients ready, out[4+2i] = its fused reduce+update done;
// all in milliseconds since the start of the step.  Exposed communication = out[2] - out[1].
// Synchronises the stream.  Returns the number of floats written through *n_out.
extern "C" int ddpb_engine_profile_step(void* handle, int batch, cudaStream_t s, float* out, int max_out,
                                        int* n_out) {
  Engine* e = (Engine*)handle;
  DDPB_REQUIRE(e && out && n_out, "engine_profile_step: null argument");
  DDPB_REQUIRE(batch > 0 && batch <= e->cfg.max_batch, "engine_profile_step: batch %d out of range", batch);
  const int nb = (int)e->buckets.size();
  DDPB_REQUIRE(max_out >= 3 + 2 * nb, "engine_profile_step: output buffer too small (%d < %d)", max_out, 3 + 2 * nb);
  if (!e->prof_start) {
    DDPB_CUDA_CHECK(cudaEventCreate(&e->prof_start)); DDPB_CUDA_CHECK(cudaEventCreate(&e->prof_fwd));
    DDPB_CUDA_CHECK(cudaEventCreate(&e->prof_bwd)); DDPB_CUDA_CHECK(cudaEventCreate(&e->prof_end));
    e->prof_ready.resize(nb); e->prof_done.resize(nb);
    for (int i = 0; i < nb; ++i) {
      DDPB_CUDA_CHECK(cudaEventCreate(&e->prof_ready[i]));
      DDPB_CUDA_CHECK(cudaEventCreate(&e->prof_done[i]));
    }
  }
  e->prof_on = true;
  e->counting = 0; e->counting_by_name.clear();
  int rc = DDPB_OK;
  cudaEventRecord(e->prof_start, s);
  rc = forward(e, batch, 1, s, e->zero_async);
  if (!rc) rc = loss_fwd(e, batch, 1, s);
  cudaEventRecord(e->prof_fwd, s);
  if (!rc) rc = backward(e, batch, s, e->zero_async);
  cudaEventRecord(e->prof_end, s);
  e->prof_on = false;
  e->launches += e->counting; add_counts(e->launches_by_name, e->counting_by_name);
  if (rc) return rc;
  DDPB_CUDA_CHECK(cudaStreamSynchronize(s));
  DDPB_CUDA_CHECK(cudaStreamSynchronize(e->comm));
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e->prof_start, e->prof_fwd); out[0] = ms;
  cudaEventElapsedTime(&ms, e->prof_start, e->prof_bwd); out[1] = ms;
  cudaEventElapsedTime(&ms, e->prof_start, e->prof_end); out[2] = ms;
  const bool bucketed = overlapped(e);
  for (int i = 0; i < nb; ++i) {
    out[3 + 2 * i] = out[4 + 2 * i] = -1.f;
    if (!bucketed) continue;
    if (cudaEventElapsedTime(&ms, e->prof_start, e->prof_ready[i]) == cudaSuccess) out[3 + 2 * i] = ms;
    if (cudaEventElapsedTime(&ms, e->prof_start, e->prof_done[i]) == cudaSuccess) out[4 + 2 * i] = ms;
  }
  (void)cudaGetLastError();
  *n_out = 3 + 2 * nb;
  return DDPB_OK;
}

// One eager training step with an event after EVERY kernel launch: the per-kernel timeline of the step
// as the GPU really runs it (warm caches, all streams concurrent), which an ncu launch list - serialised,
// cold caches - cannot show.  Row i: names + 48 * i (NUL-terminated entry point name), block[i] (conv block,
// -1 = classifier / optimizer), stream_id[i] (0 main, 1 weight-gradient, 2 / 3 communication), end_ms[i]
// (completion time since the start of the step).  graphed = 1 runs the step as a CUDA graph with event-record
// nodes (no host in the loop, three launches, the last one is read); 0 launches eagerly.  Real steps: weights
// are updated.  Synchronises.
extern "C" int ddpb_engine_trace_step(void* handle, int batch, cudaStream_t s, int graphed, char* names, int* block,
                                      int* stream_id, float* begin_ms, float* end_ms, int max_rows, int* n_rows) {
  Engine* e = (Engine*)handle;
  DDPB_REQUIRE(e && names && block && stream_id && end_ms && n_rows, "engine_trace_step: null argument");
  DDPB_REQUIRE(batch > 0 && batch <= e->cfg.max_batch, "engine_trace_step: batch %d out of range", batch);
  DDPB_REQUIRE(graphed >= 0 && graphed <= 2, "engine_trace_step: mode must be 0, 1 or 2");
  if (!e->trace_start) DDPB_CUDA_CHECK(cudaEventCreate(&e->trace_start));
  const bool stamps = graphed == 2;
  if (stamps && !e->ktrace_buf) {
    DDPB_CUDA_CHECK(cudaMalloc(&e->ktrace_buf, sizeof(unsigned long long) * 2 * kTraceSlots));
    void* p = e->ktrace_buf;
    int bad = ddpb_ktrace_set_conv_tc(p) | ddpb_ktrace_set_wgrad_tc(p) | ddpb_ktrace_set_gemm_rows_tc(p) |
              ddpb_ktrace_set_bn_relu_pool(p) | ddpb_ktrace_set_fc_ce(p) | ddpb_ktrace_set_elementwise(p) |
              ddpb_ktrace_set_sync_p2p(p) | ddpb_ktrace_set_first_layer_bwd(p) | ddpb_ktrace_set_conv_tf32(p);
    DDPB_REQUIRE(!bad, "engine_trace_step: could not publish the trace buffer to the kernels");
  }
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  if (graphed) {
    // The step as the product path runs it: one CUDA graph, no host in the loop.  Mode 1 puts an
    // event-record node after every kernel (which also disables programmatic launches across it and costs
    // a few microseconds per node); mode 2 captures the unmodified step and lets every kernel stamp
    // %globaltimer itself, so the rows are the real timeline, prologue overlap included.
    DDPB_CUDA_CHECK(cudaStreamSynchronize(s));
    DDPB_CUDA_CHECK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
  }
  e->trace_on = true; e->trace_external = graphed == 1; e->trace_mode = graphed; e->trace_n = 0; e->trace_main = s;
  e->trace_id_mark = 0;
  e->counting = 0; e->counting_by_name.clear();
  if (stamps) ktrace_enable(1);
  else if (graphed) cudaEventRecordWithFlags(e->trace_start, s, cudaEventRecordExternal);
  else cudaEventRecord(e->trace_start, s);
  int rc = forward(e, batch, 1, s, e->zero_async);
  if (!rc) rc = loss_fwd(e, batch, 1, s);
  if (!rc) rc = backward(e, batch, s, e->zero_async);
  if (stamps) ktrace_enable(0);
  e->trace_on = false; e->trace_external = false; e->trace_mode = 0;
  if (graphed) {
    cudaError_t ce = cudaStreamEndCapture(s, &graph);
    if (rc || ce != cudaSuccess || !graph) {
      (void)cudaGetLastError();
      if (graph) cudaGraphDestroy(graph);
      if (rc) return rc;
      DDPB_REQUIRE(false, "engine_trace_step: capture failed (%s)", cudaGetErrorString(ce));
    }
    tune_graph(e, graph);
    ce = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    DDPB_REQUIRE(ce == cudaSuccess, "engine_trace_step: cudaGraphInstantiate failed (%s)", cudaGetErrorString(ce));
    for (int rep = 0; rep < 3; ++rep) {           // three real steps; the last one is read
      if (stamps) {
        DDPB_CUDA_CHECK(cudaMemsetAsync(e->ktrace_buf, 0xFF, sizeof(unsigned long long) * 2 * kTraceSlots, s));
        // (begin slots start at the maximum, end slots are taken with atomicMax: clear those to zero)
        DDPB_CUDA_CHECK(cudaMemset2DAsync(e->ktrace_buf + 1, 16, 0, 8, kTraceSlots, s));
      }
      DDPB_CUDA_CHECK(cudaGraphLaunch(exec, s));
    }
    e->launches += 3 * e->counting; add_counts(e->launches_by_name, e->counting_by_name, 3);
  } else {
    e->launches += e->counting; add_counts(e->launches_by_name, e->counting_by_name);
    if (rc) return rc;
  }
  DDPB_CUDA_CHECK(cudaStreamSynchronize(s));
  DDPB_CUDA_CHECK(cudaStreamSynchronize(e->comm));
  DDPB_CUDA_CHECK(cudaStreamSynchronize(e->comm2));
  DDPB_CUDA_CHECK(cudaStreamSynchronize(e->wg));
  if (exec) cudaGraphExecDestroy(exec);
  std::vector<unsigned long long> host;
  unsigned long long t0 = ~0ull;
  if (stamps) {
    host.resize(2 * kTraceSlots);
    DDPB_CUDA_CHECK(cudaMemcpy(host.data(), e->ktrace_buf, host.size() * 8, cudaMemcpyDeviceToHost));
    for (int i = 32; i < kTraceSlots; ++i) if (host[2 * i + 1] != 0 && host[2 * i] < t0) t0 = host[2 * i];
  }
  int n = 0;
  for (size_t i = 0; i < e->trace_n && n < max_rows; ++i, ++n) {
    const Engine::TraceRec& r = e->trace[i];
    memcpy(names + 48 * n, r.name, 48);
    block[n] = r.block; stream_id[n] = r.stream;
    float b_ms = -1.f, e_ms = -1.f;
    if (stamps) {
      unsigned long long lo = ~0ull, hi = 0;
      for (int id = r.id_lo; id <= r.id_hi && id < kTraceSlots / 32; ++id)
        for (int lane = 0; lane < 32; ++lane) {
          const size_t k = 2 * ((size_t)id * 32 + lane);
          if (host[k + 1] == 0) continue;           // lane unused / kernel not instrumented
          if (host[k] < lo) lo = host[k];
          if (host[k + 1] > hi) hi = host[k + 1];
        }
      if (hi) { b_ms = (float)((double)(lo - t0) * 1e-6); e_ms = (float)((double)(hi - t0) * 1e-6); }
    } else if (cudaEventElapsedTime(&e_ms, e->trace_start, r.ev) != cudaSuccess) {
      (void)cudaGetLastError();
      e_ms = -1.f;
    }
    if (begin_ms) begin_ms[n] = b_ms;
    end_ms[n] = e_ms;
  }
  *n_rows = n;
  return DDPB_OK;
}

// Eval forward: BN uses running statistics; eval_loss += mean CE of the batch; correct += hits.
extern "C" int ddpb_engine_eval_step(void* handle, int batch, cudaStream_t s) {
  Engine* e = (Engine*)handle;
  DDPB_REQUIRE(batch > 0 && batch <= e->cfg.max_batch, "engine_eval_step: batch %d out of range", batch);
  e->counting = 0; e->counting_by_name.clear();
  ECHECK(forward(e, batch, 0, s));
  ECHECK(loss_fwd(e, batch, 0, s));
  e->launches += e->counting; add_counts(e->launches_by_name, e->counting_by_name);
  return DDPB_OK;
}

extern "C" int ddpb_engine_eval_reset(void* handle, cudaStream_t s) {
  Engine* e = (Engine*)handle;
  DDPB_CUDA_CHECK(cudaMemsetAsync(e->eval_loss, 0, 64, s));
  return DDPB_OK;
}

// Collective over the symmetric arena outside the training step: what = 0 broadcast master weights,
// 1 broadcast BN running statistics (rank 0 -> all; DDP's init / buffer sync), 2 all-gather the
// momentum shards of the two-shot buckets, 3 the same for the sharded fp32 master weights (before a
// checkpoint / state_dict).
extern "C" int ddpb_engine_collective(void* handle, int what, cudaStream_t s) {
  Engine* e = (Engine*)handle;
  const auto& c = e->cfg;
  if (c.world == 1 || !e->peers_set) return DDPB_OK;
  void *buf[DDPB_MAX_RANKS], *f[DDPB_MAX_RANKS];
  peer_table(e, e->off_flags, f);
  const int slot = DDPB_P2P_MAX_SLOTS - 2;
  if (what == 0) {
    peer_table(e, e->off_master, buf);
    ECHECK(ddpb_p2p_sync(DDPB_P2P_BCAST, buf, nullptr, nullptr, f, nullptr, e->epochs, c.rank, c.world, slot, 0,
                         e->arena_elems, 0, 0, 0, 0, s));
    return ddpb_engine_refresh_shadow(handle, s);
  }
  if (what == 1) {
    peer_table(e, e->off_rstats, buf);
    return ddpb_p2p_sync(DDPB_P2P_BCAST, buf, nullptr, nullptr, f, nullptr, e->epochs, c.rank, c.world, slot, 0,
                         round_up(e->total_c * 2, 4), 0, 0, 0, 0, s);
  }
  if (what == 2) {
    peer_table(e, e->off_mom, buf);
    if (c.strategy == 3) {
      for (size_t i = 0; i < e->buckets.size(); ++i)
        if (e->buckets[i].mode == DDPB_P2P_TWOSHOT)
          ECHECK(ddpb_p2p_sync(DDPB_P2P_ALLGATHER, buf, nullptr, nullptr, f, nullptr, e->epochs, c.rank, c.world,
                               slot, e->buckets[i].off, e->buckets[i].numel, 0, 0, 0, 0, s));
      return DDPB_OK;
    }
    if (c.strategy == 2)
      return ddpb_p2p_sync(DDPB_P2P_ALLGATHER, buf, nullptr, nullptr, f, nullptr, e->epochs, c.rank, c.world, slot,
                           0, e->arena_elems, 0, 0, 0, 0, s);
    if (c.strategy == 1)   // root owns all momentum: broadcast it
      return ddpb_p2p_sync(DDPB_P2P_BCAST, buf, nullptr, nullptr, f, nullptr, e->epochs, c.rank, c.world, slot, 0,
                           e->arena_elems, 0, 0, 0, 0, s);
    return DDPB_OK;
  }
  if (what == 3) {
    // fp32 master weights of the sharded ranges (conv weights): every rank holds current values only for
    // the slices it owns; same ownership rule as the momentum above
    if (e->shard.empty()) return DDPB_OK;
    peer_table(e, e->off_master, buf);
    if (c.strategy == 3) {
      for (size_t i = 0; i < e->buckets.size(); ++i)
        if (e->buckets[i].mode == DDPB_P2P_TWOSHOT)
          ECHECK(ddpb_p2p_sync(DDPB_P2P_ALLGATHER, buf, nullptr, nullptr, f, nullptr, e->epochs, c.rank, c.world,
                               slot, e->buckets[i].off, e->buckets[i].numel, 0, 0, 0, 0, s));
      return DDPB_OK;
    }
    if (c.strategy == 2)
      return ddpb_p2p_sync(DDPB_P2P_ALLGATHER, buf, nullptr, nullptr, f, nullptr, e->epochs, c.rank, c.world, slot,
                           0, e->arena_elems, 0, 0, 0, 0, s);
    if (c.strategy == 1)
      return ddpb_p2p_sync(DDPB_P2P_BCAST, buf, nullptr, nullptr, f, nullptr, e->epochs, c.rank, c.world, slot, 0,
                           e->arena_elems, 0, 0, 0, 0, s);
    return DDPB_OK;
  }
  DDPB_REQUIRE(false, "engine_collective: unknown op %d", what);
}
