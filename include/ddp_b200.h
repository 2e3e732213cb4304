/* C ABI of libddpb200.so - the sm_100a engine of the B200 data-parallel framework.
 *
 * Conventions
 *   - every function returns 0 on success; on failure ddpb_last_error() describes the problem;
 *   - activations are NHWC, conv weights KRSC (= [Cout][3][3][Cin]); `dtype` is DDPB_F32/DDPB_BF16;
 *   - all work is enqueued on `stream`; nothing synchronises the device;
 *   - functions documented as "accumulates" add into their output (zero it first);
 *   - no function allocates device memory unless its comment says so.
 * Reference citations (ref ...) are relative to /root/reference/src/.
 */
#ifndef DDP_B200_H_
#define DDP_B200_H_

#include <cuda_runtime.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DDPB_F32 0
#define DDPB_BF16 1

#define DDPB_MAX_RANKS 8
#define DDPB_P2P_MAX_BLOCKS 320
#define DDPB_P2P_MAX_SHARD_RANGES 24
#define DDPB_P2P_MAX_SLOTS 64

#define DDPB_P2P_ONESHOT 0
#define DDPB_P2P_TWOSHOT 1
#define DDPB_P2P_ROOT 2
#define DDPB_P2P_BCAST 3
#define DDPB_P2P_ALLGATHER 4
/* OR into `mode`: skip the kernel's closing cross-rank barrier (a later launch on the same stream closes the step) */
#define DDPB_P2P_NO_EXIT_BARRIER 0x100

const char* ddpb_last_error(void);

/* ---- data / optimizer / layout (csrc/kernels/elementwise.cu) ------------------------------ */

/* Gather `batch` images (rows `index`, or 0..batch-1 when null) of a uint8 [N,32,32,3] dataset,
 * apply RandomCrop(32, pad 4) with offsets crop_xy[b] in [0,8] (null: none), horizontal flip
 * (null: none), /255 and per-channel normalisation; write NHWC.  ref Part 1/main.py:82-93. */
int ddpb_augment(const void* images_u8, const void* index_i64, const void* crop_xy_i32, const void* flip_u8,
                 void* out, int batch, int out_dtype, const float* mean3, const float* std3, cudaStream_t stream);
/* The same, and the batch's int64 targets are copied labels_src -> labels_dst by the same kernel (both nullable). */
int ddpb_augment_labels(const void* images_u8, const void* index_i64, const void* crop_xy_i32, const void* flip_u8,
                        void* out, const void* labels_src_i64, void* labels_dst_i64, int batch, int out_dtype,
                        const float* mean3, const float* std3, cudaStream_t stream);

/* SGD with momentum and weight decay over a flat range (torch.optim.SGD semantics, dampening 0,
 * no Nesterov; ref Part 1/main.py:114-115): g = grad*grad_scale + wd*p; m = momentum*m + g;
 * p -= lr*m.  Refreshes the bf16 copy when shadow_bf16 is non-null.  n must be a multiple of 4. */
int ddpb_sgd_step(float* param, float* mom, const float* grad, void* shadow_bf16, long long n, float lr,
                  float momentum, float wd, float grad_scale, cudaStream_t stream);

int ddpb_cast_f32_bf16(const float* src, void* dst_bf16, long long n, cudaStream_t stream);
int ddpb_fill_f32(float* dst, float value, long long n, cudaStream_t stream);
int ddpb_nchw_to_nhwc(const void* src, void* dst, int n, int c, int hw, int src_dtype, int dst_dtype,
                      cudaStream_t stream);

/* ---- BatchNorm + ReLU (+ 2x2 max-pool) (csrc/kernels/bn_relu_pool.cu) --------------------- */

/* ref Part 1/model.py:16,24-25.  y: conv output [batch,H,W,C]; stats: [2C] per-channel sum and
 * sum of squares of y (training); out: [batch,H(/2),W(/2),C].  Training mode also writes
 * save_mean/save_invstd and updates running stats (momentum, unbiased variance) and the
 * batch counter; eval mode normalises with the running statistics. */
int ddpb_bn_relu_pool_fwd(const void* y, const float* stats, const float* gamma, const float* beta, void* out,
                          float* save_mean, float* save_invstd, float* running_mean, float* running_var,
                          long long* num_batches_tracked, int batch, int H, int W, int C, int pool, int training,
                          float eps, float momentum, int dtype, cudaStream_t stream);
/* Same, and `winner` (nullable, shaped like out) receives the conv output that won each ReLU + max-pool
 * window (NaN where ReLU blocked the whole window): the input of ddpb_conv3x3_dgrad_bnred_tc. */
int ddpb_bn_relu_pool_fwd_winner(const void* y, const float* stats, const float* gamma, const float* beta, void* out,
                                 float* save_mean, float* save_invstd, float* running_mean, float* running_var,
                                 long long* num_batches_tracked, void* winner, int batch, int H, int W, int C,
                                 int pool, int training, float eps, float momentum, int dtype, cudaStream_t stream);

/* Backward of the above.  sums: zeroed [2C] scratch.  dgamma/dbeta are written; dbias (gradient of
 * the conv bias that precedes the BN, may be null) is accumulated. */
int ddpb_bn_relu_pool_bwd(const void* dout, const void* y, const float* save_mean, const float* save_invstd,
                          const float* gamma, const float* beta, float* sums, void* dy, float* dgamma,
                          float* dbeta, float* dbias, int batch, int H, int W, int C, int pool, int dtype,
                          cudaStream_t stream);
/* Second pass of the above only (sums already reduced, e.g. by ddpb_conv3x3_dgrad_bnred_tc). */
int ddpb_bn_relu_pool_bwd_apply(const void* dout, const void* y, const float* save_mean, const float* save_invstd,
                                const float* gamma, const float* beta, float* sums, void* dy, float* dgamma,
                                float* dbeta, float* dbias, int batch, int H, int W, int C, int pool, int dtype,
                                cudaStream_t stream);

/* Same computation as ddpb_bn_relu_pool_bwd for callers that cannot supply the [2*C] scratch: a
 * host that pre-committed all device memory to its own pool, or a plug-in interface that passes only
 * inputs and outputs.  The scratch is a fixed 16 KB device array inside the library (resident on
 * every GPU that loads it), so this call performs no allocation of any kind.  Limits: bf16 tensors
 * C <= DDPB_BN_NOALLOC_MAX_C (2048, what the array holds); fp32 tensors C <=
 * DDPB_BN_NOALLOC_MAX_C_F32 (1024: the BN kernels give each thread 16 bytes of channels and one
 * 256-thread block spans a pixel, so fp32 reaches half as far); C / 8 (bf16) or C / 4 (fp32) must be
 * a power of two; H and W even when pool != 0.  A call outside these limits is rejected with an error
 * that names the limit, before the shared array is touched.  All calls on a device share the
 * array, so calls that overlap on the GPU race - serialise them; the first call on each device
 * resolves the array's device address on the host and should happen before any stream capture.
 * This entry point has no counterpart in the reference. */
#define DDPB_BN_NOALLOC_MAX_C 2048
#define DDPB_BN_NOALLOC_MAX_C_F32 1024
int ddpb_bn_relu_pool_bwd_noalloc(const void* dout, const void* y, const float* save_mean,
                                  const float* save_invstd, const float* gamma, const float* beta, void* dy,
                                  float* dgamma, float* dbeta, float* dbias, int batch, int H, int W, int C,
                                  int pool, int dtype, cudaStream_t stream);

/* stats[0:C] += sum, stats[C:2C] += sum of squares over `pixels` rows of y (accumulates). */
int ddpb_channel_stats(const void* y, float* stats, long long pixels, int C, int dtype, cudaStream_t stream);

int ddpb_maxpool2x2_fwd(const void* x, void* out, int batch, int H, int W, int C, int dtype, cudaStream_t stream);
int ddpb_maxpool2x2_bwd(const void* x, const void* dout, void* dx, int batch, int H, int W, int C, int dtype,
                        cudaStream_t stream);

/* ---- classifier + cross-entropy (csrc/kernels/fc_ce.cu) ----------------------------------- */

/* ref Part 1/model.py:40,45 and Part 1/main.py:39,70-71.  act: [batch,512]; w: fp32 [10,512].
 * loss (1 float) accumulates sum_b CE_b / loss_divisor; correct (1 int, may be null) accumulates
 * the number of argmax hits; logits/dlogits ([batch,10] fp32) may be null. */
int ddpb_fc_ce_fwd(const void* act, const float* w, const float* bias, const void* target_i64, float* logits,
                   float* dlogits, float* loss, int* correct, int batch, int loss_divisor, int dtype,
                   cudaStream_t stream);

/* The same, and when dact ([batch,512], dtype of act) is non-null the kernel also writes the gradient of the
 * classifier's input (what ddpb_fc_bwd computes as dact), so that backward can start with the last block at once. */
int ddpb_fc_ce_fwd_dact(const void* act, const float* w, const float* bias, const void* target_i64, float* logits,
                        float* dlogits, float* loss, int* correct, void* dact, int batch, int loss_divisor, int dtype,
                        cudaStream_t stream);

/* dact [batch,512] is written (null: skipped - weight and bias gradients only); dw [10,512] and db [10] accumulate. */
int ddpb_fc_bwd(const void* act, const float* w, const float* dlogits, void* dact, float* dw, float* db,
                int batch, int dtype, cudaStream_t stream);

/* ---- convolutions ------------------------------------------------------------------------- */

/* SIMT fp32-accumulate kernels with fp32 weights (csrc/kernels/conv_direct.cu): the fp32 parity
 * path, the oracle for the tensor-core kernels, and the Cin=3 first layer (dedicated kernels).
 * ref Part 1/model.py:18-23.  stats (may be null) accumulates per-channel sum / sum of squares. */
int ddpb_conv3x3_fwd_simt(const void* x, const float* w, const float* bias, void* y, float* stats, int B, int H,
                          int W, int Cin, int Cout, int dtype, cudaStream_t stream);
int ddpb_conv3x3_dgrad_simt(const void* dy, const float* w, void* dx, int B, int H, int W, int Cin, int Cout,
                            int dtype, cudaStream_t stream);
/* dw accumulates. */
int ddpb_conv3x3_wgrad_simt(const void* dy, const void* x, float* dw, int B, int H, int W, int Cin, int Cout,
                            int dtype, cudaStream_t stream);

/* tcgen05 / TMEM / TMA implicit-GEMM kernels, bf16 operands, fp32 accumulation
 * (csrc/kernels/conv_tc.cu, wgrad_tc.cu).  Cin, Cout multiples of 64 (wgrad: Cout of 128);
 * H = W in {2,4,8,16,32}.  tile_n / split_k = 0 choose automatically. */
int ddpb_conv3x3_fwd_tc(const void* x, const void* w_bf16, const float* bias, void* y, float* stats, int B, int H,
                        int W, int Cin, int Cout, int tile_n, cudaStream_t stream);
int ddpb_conv3x3_dgrad_tc(const void* dy, const void* w_bf16, void* dx, int B, int H, int W, int Cin, int Cout,
                          int tile_n, cudaStream_t stream);
/* Either of the two with per-CTA phase time stamps (tuning aid, no reference counterpart): trace[cta][8] =
 * %globaltimer (ns) at CTA start, prologue done, first stage landed, last MMA issued, accumulator complete,
 * split-K partials exchanged, tile stored, CTA done.  cta = x + gridDim.x * (y + gridDim.y * z). */
int ddpb_conv3x3_trace(int dgrad, const void* in, const void* w_bf16, const float* bias, void* out, float* stats,
                       int B, int H, int W, int Cin, int Cout, int tile_n, unsigned long long* trace,
                       cudaStream_t stream);
/* Forward conv and, when the whole grid is resident at once (*fused = 1), this block's training-mode
 * BatchNorm + ReLU + optional 2x2 max-pool in the same kernel (grid-wide barrier after the statistics).
 * *fused = 0: only y / stats were produced, run ddpb_bn_relu_pool_fwd(_winner) next.
 * `barrier`: one zeroed uint32 per launch.  No reference counterpart (cuDNN runs conv and BN separately). */
int ddpb_conv3x3_fwd_bn_tc(const void* x, const void* w_bf16, const float* bias, void* y, float* stats, int B, int H,
                           int W, int Cin, int Cout, int tile_n, const float* gamma, const float* beta, void* act,
                           void* winner, float* save_mean, float* save_invstd, float* running_mean,
                           float* running_var, long long* num_batches_tracked, int pool, float eps, float momentum,
                           unsigned int* barrier, int* fused, cudaStream_t stream);
/* dgrad plus, in its epilogue, the BatchNorm-backward reduction of the block that produced this conv's
 * input: sums[0:Cin] += sum dz, sums[Cin:2*Cin] += sum dz*xhat, where dz is dx masked by that block's
 * ReLU / max-pool.  y_win (bf16 [B,H,W,Cin]) is the `winner` tensor of ddpb_bn_relu_pool_fwd_winner
 * for a pooling block, or the block's conv output y itself when it does not pool.
 * Follow with ddpb_bn_relu_pool_bwd_apply.  Fuses the first BN-backward kernel of every block into the
 * tensor-core kernel that precedes it (no reference counterpart: autograd runs them separately). */
int ddpb_conv3x3_dgrad_bnred_tc(const void* dy, const void* w_bf16, void* dx, int B, int H, int W, int Cin, int Cout,
                                int tile_n, const void* y_win, const float* mean, const float* invstd,
                                const float* gamma, const float* beta, float* sums, cudaStream_t stream);
/* dw (fp32 KRSC) accumulates. */
int ddpb_conv3x3_wgrad_tc(const void* dy, const void* x, float* dw, int B, int H, int W, int Cin, int Cout,
                          int split_k, cudaStream_t stream);
/* The same for the output channels [co_lo, co_lo + co_n) only (both multiples of 128): rows co_lo.. of dw.  Lets a
 * caller hand the first part of a layer's gradient to the communication stream while the rest is still being
 * computed (the engine's sub-bucket hand-off; DDP has no counterpart below bucket granularity). */
int ddpb_conv3x3_wgrad_tc_range(const void* dy, const void* x, float* dw, int B, int H, int W, int Cin, int Cout,
                                int co_lo, int co_n, int split_k, cudaStream_t stream);

/* First layer (Cin = 3) on the tensor cores: its 3x3x3 patches are laid out as a [B*H*W][64] bf16
 * matrix (27 real columns), after which forward is a 1x1 convolution and the weight gradient a plain
 * split-K GEMM.  im2col also pads the bf16 weights [Cout][27] -> [Cout][64]. */
int ddpb_im2col3x3_c3(const void* x_nhwc3, const void* w_bf16, void* cols, void* w_pad, int B, int H, int W, int Cout,
                      int zero_pad /* 0: columns 32..63 of cols were zeroed by the caller and are left alone */,
                      cudaStream_t stream);
int ddpb_conv1x1_fwd_tc(const void* x, const void* w_bf16, const float* bias, void* y, float* stats, int B, int H,
                        int W, int Cin, int Cout, cudaStream_t stream);
/* y[P][Cout] = x[P][64] * w[Cout][64]^T + bias (+ BN statistics): the first layer's forward over its patch
 * matrix; one CTA walks several 128-row tiles with double-buffered TMEM accumulators. */
int ddpb_gemm_rows_fwd_tc(const void* x, const void* w_bf16, const float* bias, void* y, float* stats, long long P,
                          int Cout, cudaStream_t stream);
/* dw (fp32 [Cout][ld], columns < cols) accumulates dy^T * x_cols. */
int ddpb_patch_wgrad_tc(const void* dy, const void* x_cols, float* dw, int B, int H, int W, int Cout, int ld, int cols,
                        int split_k, cudaStream_t stream);

/* fp32 convolutions on the tensor cores (csrc/kernels/conv_tf32.cu): tcgen05.mma kind::tf32, fp32 NHWC in / out,
 * fp32 KRSC weights, fp32 accumulation.  The engine's fp32 path (the reference is fp32 throughout,
 * ref Part 1/main.py:10; Part 1/model.py:18-23).  Channels multiples of 64 (wgrad: Cout multiple of 128). */
int ddpb_conv3x3_fwd_tf32(const float* x, const float* w, const float* bias, float* y, float* stats, int B, int H, int W,
                          int Cin, int Cout, cudaStream_t stream);
int ddpb_conv3x3_dgrad_tf32(const float* dy, const float* w, float* dx, int B, int H, int W, int Cin, int Cout,
                            cudaStream_t stream);
int ddpb_conv3x3_wgrad_tf32(const float* dy, const float* x, float* dw, int B, int H, int W, int Cin, int Cout,
                            int split_k, cudaStream_t stream);

/* Backward of block 0 (ref Part 1/model.py:16,18-25 for the first conv / BN / ReLU / pool) in one launch: the
 * BatchNorm + ReLU + max-pool backward apply pass fused with the 3 -> 64 convolution's weight gradient, so that
 * the 33.5 MB gradient of the conv output never goes through global memory.  dout [B,16,16,64] bf16, y
 * [B,32,32,64] bf16, x_cols [B*1024][64] bf16, sums [128] (BN-backward reduction).  dw (fp32 [64][ld], columns <
 * cols <= 32) and dbias (nullable) accumulate; dgamma / dbeta are written.  ctas = 0 picks. */
int ddpb_first_layer_bwd_fused(const void* dout, const void* y, const void* x_cols, const float* save_mean,
                               const float* save_invstd, const float* gamma, const float* beta, const float* sums,
                               float* dw, float* dgamma, float* dbeta, float* dbias, int B, int H, int W, int C, int ld,
                               int cols, int ctas, cudaStream_t stream);

/* Allow (1) / forbid (0) programmatic dependent launches (default: on unless DDPB_PDL=0); returns the
 * previous setting.  They are only ever used while the launching stream is being captured. */
int ddpb_set_pdl(int on);

/* Diagnostic (csrc/kernels/tma_probe.cu, tools/tma_probe.py): per-CTA L2 -> shared-memory ingest rate of
 * the conv load patterns.  No reference counterpart. */
int ddpb_probe_tma(int mode, int stages, int iters, int ctas, int smem_pad_kb, int cols, const void* act, const void* wgt,
                   const void* flat, long long flat_bytes, cudaStream_t stream);

/* ---- peer-to-peer gradient engine (csrc/kernels/sync_p2p.cu, csrc/runtime/symm_mem.cpp) --- */

/* One fused reduce(+SGD) / broadcast / all-gather step over elements [off, off+n) of the symmetric
 * arenas; pointer tables have `world` entries indexed by rank.  Replaces dist.gather/scatter
 * (ref Part 2a/main.py:117-127), dist.all_reduce + divide (ref Part 2b/main.py:116-119) and DDP's
 * bucket all-reduce / broadcasts (ref Part 3/main.py:61), each fused with optimizer.step()
 * (ref Part 2a/main.py:96).  flags == null disables the cross-rank barriers.  `blocks` = 0: the small-footprint
 * shape that runs next to the backward kernels; > 0: 512-thread CTAs on that grid; < 0: 512-thread CTAs, grid
 * sized by the data (a caller that owns the GPU). */
int ddpb_p2p_sync(int mode, void* const* grad, void* const* master, void* const* shadow, void* const* flags,
                  float* mom, void* epochs, int rank, int world, int slot, long long off, long long n, float lr,
                  float momentum, float wd, int blocks, cudaStream_t stream);

/* The same with "sharded" element ranges [lo, hi) (shard[2k], shard[2k+1]; 8-element aligned): ranges whose
 * fp32 master weights the owning rank keeps to itself, like the momentum, publishing only the bf16 copy
 * (the conv weights: nothing but the optimizer reads their fp32 form).  Two-shot / root modes with a bf16
 * copy only; n_shard = 0 is ddpb_p2p_sync.  Bring stale slices up to date with DDPB_P2P_ALLGATHER (two-shot)
 * or DDPB_P2P_BCAST (root) over the master buffers before a full state_dict is read. */
int ddpb_p2p_sync_sharded(int mode, void* const* grad, void* const* master, void* const* shadow, void* const* flags,
                          float* mom, void* epochs, int rank, int world, int slot, long long off, long long n,
                          float lr, float momentum, float wd, int blocks, const long long* shard, int n_shard,
                          cudaStream_t stream);

/* NVLS variant of the two-shot mode: mc_* are multicast mappings (runtime/vmm_arena.cpp); the switch
 * performs the reduction (multimem.ld_reduce) and the broadcast (multimem.st).  Replaces the same
 * reference calls as ddpb_p2p_sync. */
int ddpb_p2p_sync_nvls(void* mc_grad, void* mc_master, void* mc_shadow, float* master_local, void* const* flags,
                       float* mom, void* epochs, int rank, int world, int slot, long long off, long long n,
                       float lr, float momentum, float wd, int blocks, int skip_exit_barrier, cudaStream_t stream);
/* NVLS with sharded ranges (see ddpb_p2p_sync_sharded) and a choice of mode: DDPB_P2P_TWOSHOT as above, or
 * DDPB_P2P_ONESHOT where every rank pulls the switch-reduced sum of the whole range and updates its own
 * replica (master_local / shadow_local) without any store over NVLink - the small buckets; or DDPB_P2P_ROOT
 * where rank 0 alone reduces and updates the range and multicasts the result (Part 2a's gather -> mean ->
 * scatter, ref Part 2a/main.py:117-127, as one kernel). */
int ddpb_p2p_sync_nvls_sharded(int mode, void* mc_grad, void* mc_master, void* mc_shadow, float* master_local,
                               void* shadow_local, void* const* flags, float* mom, void* epochs, int rank, int world,
                               int slot, long long off, long long n, float lr, float momentum, float wd, int blocks,
                               int skip_exit_barrier, const long long* shard, int n_shard, cudaStream_t stream);
/* A cross-rank wait that times out (DDPB_SYNC_TIMEOUT_S, default 120) records rec8 = {1, rank, peer, slot,
 * phase, block, epoch, 0} in host memory before it traps the kernel; returns 1 when such a record exists.
 * No reference counterpart (the reference blocks forever inside gloo when a rank dies). */
int ddpb_p2p_last_timeout(unsigned int* rec8);

/* Diagnostic (tools/peer_read_probe.py): read bandwidth of `src` (local, peer or multicast memory) with a
 * given CTA shape; kind 0 plain loads, 1 ld.relaxed.sys, 2 multimem.ld_reduce, 3-6 with bf16 stores / local
 * optimizer traffic added (see sync_p2p.cu).  No reference counterpart. */
int ddpb_probe_peer_read(const void* src, void* sink, long long n_floats, int kind, int blocks, int threads, int unroll,
                         void* dst_bf16, float* local_rw, cudaStream_t stream);

/* Virtual-memory-management arena with an NVSwitch multicast mapping (csrc/runtime/vmm_arena.cpp).
 * ddpb_vmm_alloc ALLOCATES (cuMemCreate); fds are POSIX file descriptors to pass between processes. */
int ddpb_vmm_supported(int* multicast_supported);
int ddpb_vmm_padded_size(size_t bytes, int num_devices, size_t* padded);
int ddpb_vmm_alloc(size_t bytes, void** ptr, int* fd, unsigned long long* handle);
int ddpb_vmm_import(int fd, size_t bytes, void** ptr);
int ddpb_mc_create(size_t bytes, int num_devices, int* fd, unsigned long long* mc_handle);
int ddpb_mc_import(int fd, unsigned long long* mc_handle);
int ddpb_mc_add_device(unsigned long long mc_handle);
int ddpb_mc_bind_and_map(unsigned long long mc_handle, unsigned long long mem_handle, size_t bytes, void** mc_ptr);
int ddpb_vmm_unmap(void* ptr, size_t bytes);
int ddpb_vmm_release(unsigned long long handle);

/* Symmetric memory: ALLOCATES `bytes` of zeroed device memory with cudaMalloc on the current
 * device and returns it with a 64-byte IPC handle that peers open with ddpb_symm_open. */
int ddpb_symm_alloc(size_t bytes, void** ptr, void* handle64);
int ddpb_symm_open(const void* handle64, void** peer_ptr);
int ddpb_symm_close(void* peer_ptr);
int ddpb_symm_free(void* ptr);
/* Enable peer access from the current device to `peer_device` (same-process multi-GPU tests). */
int ddpb_enable_peer_access(int peer_device);

/* ---- training-step engine (csrc/runtime/engine.cpp) ---------------------------------------- */

#define DDPB_MAX_BLOCKS 24

/* Strategy: 0 none (ref Part 1), 1 gather_scatter (Part 2a), 2 allreduce (Part 2b), 3 bucketed (Part 3). */
typedef struct ddpb_engine_config {
  int dtype;            /* DDPB_F32 / DDPB_BF16 activations + conv operands */
  int max_batch;        /* per-rank batch the buffers are sized for */
  int world, rank;
  int strategy;
  int num_blocks;       /* conv+BN+ReLU(+pool) blocks */
  int cin[DDPB_MAX_BLOCKS], cout[DDPB_MAX_BLOCKS], hw[DDPB_MAX_BLOCKS], pool[DDPB_MAX_BLOCKS];
  float lr, momentum, wd;
  int bucket_plan;      /* 0 size-capped (DDP rule), 1 per layer, 2 single */
  float first_bucket_mb, bucket_cap_mb;
  int use_graph;        /* capture each step kind into a CUDA graph */
  int use_tc;           /* tcgen05 kernels for bf16 layers with Cin%64==0, Cout%128==0 */
  int external_sync;    /* 1: the engine neither syncs nor steps; the caller reduces `grad` and calls
                           ddpb_sgd_step (NCCL baseline path) */
  void* external_arena; /* non-null: caller-owned symmetric arena (e.g. cuMemCreate memory bound to a
                           multicast object, runtime/vmm_arena.cpp) of external_arena_bytes bytes */
  size_t external_arena_bytes;
} ddpb_engine_config;

typedef struct ddpb_engine_info_t {
  void* arena; size_t arena_bytes; unsigned char ipc_handle[64];
  long long arena_elems, total_channels;
  float *grad, *master, *mom; void* shadow;
  float *running_mean, *running_var; long long* num_batches_tracked;
  void* input; long long* target; float *loss, *eval_loss; int* correct; float* logits;
  int num_buckets;
  long long bucket_off[DDPB_P2P_MAX_SLOTS], bucket_numel[DDPB_P2P_MAX_SLOTS];
  int bucket_mode[DDPB_P2P_MAX_SLOTS], bucket_last_block[DDPB_P2P_MAX_SLOTS];
  long long fc_w_off, fc_b_off;
  long long conv_w_off[DDPB_MAX_BLOCKS], conv_b_off[DDPB_MAX_BLOCKS], bn_w_off[DDPB_MAX_BLOCKS],
      bn_b_off[DDPB_MAX_BLOCKS], ch_off[DDPB_MAX_BLOCKS];
  void *y[DDPB_MAX_BLOCKS], *act[DDPB_MAX_BLOCKS];
  long long launches;   /* kernels launched by the engine so far (graph replays included) */
  int graph_active;     /* 1 when steps are being replayed from captured CUDA graphs */
  int multicast;        /* 1 when the NVLS (multimem) reduce kernels are in use */
} ddpb_engine_info_t;

/* ALLOCATES all device buffers (cudaMalloc), including the symmetric arena. */
/* Bytes of symmetric arena this configuration needs (for cfg.external_arena); allocates nothing. */
int ddpb_engine_arena_bytes(const ddpb_engine_config* cfg, size_t* bytes);
int ddpb_engine_create(const ddpb_engine_config* cfg, void** handle);
int ddpb_engine_destroy(void* handle);
int ddpb_engine_info(void* handle, ddpb_engine_info_t* out);
int ddpb_engine_set_peers(void* handle, void* const* peer_arena);
int ddpb_engine_set_multicast(void* handle, void* mc_arena);
int ddpb_engine_set_lr(void* handle, float lr);
int ddpb_engine_refresh_shadow(void* handle, cudaStream_t stream);
/* One iteration on the batch already staged in info.input / info.target.  forward = conv stack +
 * classifier + loss (info.loss); backward = all gradients + gradient sync + optimizer step
 * (ref Part 2a/main.py:88-96); train_step = both. */
int ddpb_engine_forward(void* handle, int batch, cudaStream_t stream);
int ddpb_engine_backward(void* handle, int batch, cudaStream_t stream);
int ddpb_engine_train_step(void* handle, int batch, cudaStream_t stream);
/* One eager step instrumented with CUDA events (replaces the reference's host timers around the
 * loop body, ref Part 1/main.py:33-43, with a device timeline).  out[0] forward end, out[1] last
 * compute kernel of backward, out[2] step end, then (ready, done) per bucket; milliseconds since the
 * step start.  Exposed communication = out[2] - out[1].  Synchronises `stream`. */
int ddpb_engine_profile_step(void* handle, int batch, cudaStream_t stream, float* out, int max_out, int* n_out);
/* Same idea at kernel granularity: one eager step with an event after every launch on every stream.
 * Row i: names + 48 * i (entry point name), block[i] (conv block, -1 classifier / optimizer),
 * stream_id[i] (0 main, 1 weight-gradient, 2 / 3 communication), end_ms[i] (completion time since the
 * step start).  graphed = 1: the step runs as one CUDA graph with event-record nodes (the product path's
 * timing, minus programmatic-launch overlap); 0: eager launches (host-bound for the small kernels);
 * 2: the unmodified production graph with every kernel stamping %globaltimer itself (begin_ms = first CTA
 * start, end_ms = last warp end): the real timeline, programmatic launch overlap included.
 * The reference has only three host timers per iteration (ref Part 1/main.py:33-43). */
int ddpb_engine_trace_step(void* handle, int batch, cudaStream_t stream, int graphed, char* names, int* block,
                           int* stream_id, float* begin_ms, float* end_ms, int max_rows, int* n_rows);
/* "entry_point=count;..." of everything launched since the engine was created, graph replays included
 * (no reference counterpart; the reference's sync calls are visible in its Python source instead).
 * Returns the buffer length needed. */
int ddpb_engine_launch_counts(void* handle, char* buf, int cap);
/* ref Part 1/main.py:61-76: eval-mode forward; accumulates info.eval_loss / info.correct. */
int ddpb_engine_eval_step(void* handle, int batch, cudaStream_t stream);
int ddpb_engine_eval_reset(void* handle, cudaStream_t stream);
/* what: 0 broadcast weights from rank 0, 1 broadcast BN running stats, 2 gather momentum shards. */
int ddpb_engine_collective(void* handle, int what, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DDP_B200_H_ */
