"""Tensor-level wrappers around the native kernels (one Python function per C entry point).

Layouts: activations NHWC, conv weights KRSC ``[Cout,3,3,Cin]``; see ``include/ddp_b200.h``.
These are what the unit tests and the Python-driven engine call; the C++ step engine calls the
same kernels directly.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence, Tuple

import torch

from ..config import CIFAR_MEAN, CIFAR_STD
from . import native as nv
from .native import cf, cll, check, dtype_code, lib, ptr, stream_handle

BN_EPS = 1e-5
BN_MOMENTUM = 0.1


# ---- data / optimizer ---------------------------------------------------------------------
def augment(images_u8: torch.Tensor, index: Optional[torch.Tensor], crop_xy: Optional[torch.Tensor],
            flip: Optional[torch.Tensor], dtype: torch.dtype, out: Optional[torch.Tensor] = None,
            batch: Optional[int] = None, labels: Optional[Tuple[torch.Tensor, torch.Tensor]] = None) -> torch.Tensor:
    """``labels=(src, dst)``: the batch's int64 targets are copied by the same kernel."""
    b = batch if batch is not None else (index.numel() if index is not None else images_u8.shape[0])
    if out is None:
        out = torch.empty((b, 32, 32, 3), dtype=dtype, device=images_u8.device)
    mean = (ctypes.c_float * 3)(*CIFAR_MEAN)
    std = (ctypes.c_float * 3)(*CIFAR_STD)
    if labels is not None:
        src, dst = labels
        assert src.dtype == torch.int64 and dst.dtype == torch.int64 and src.numel() >= b and dst.numel() >= b
        check(lib().ddpb_augment_labels(ptr(images_u8), ptr(index), ptr(crop_xy), ptr(flip), ptr(out), ptr(src), ptr(dst), b,
                                        dtype_code(dtype), mean, std, stream_handle()), "augment_labels")
        return out
    check(lib().ddpb_augment(ptr(images_u8), ptr(index), ptr(crop_xy), ptr(flip), ptr(out), b,
                             dtype_code(dtype), mean, std, stream_handle()), "augment")
    return out


def sgd_step(param: torch.Tensor, mom: torch.Tensor, grad: torch.Tensor, shadow: Optional[torch.Tensor],
             lr: float, momentum: float, wd: float, grad_scale: float = 1.0) -> None:
    check(lib().ddpb_sgd_step(ptr(param), ptr(mom), ptr(grad), ptr(shadow), cll(param.numel()), cf(lr),
                              cf(momentum), cf(wd), cf(grad_scale), stream_handle()), "sgd_step")


def cast_bf16(src: torch.Tensor, dst: torch.Tensor) -> None:
    check(lib().ddpb_cast_f32_bf16(ptr(src), ptr(dst), cll(src.numel()), stream_handle()), "cast")


# ---- BN + ReLU + pool -----------------------------------------------------------------------
def channel_stats(y: torch.Tensor, stats: Optional[torch.Tensor] = None) -> torch.Tensor:
    c = y.shape[-1]
    if stats is None:
        stats = torch.zeros(2 * c, dtype=torch.float32, device=y.device)
    check(lib().ddpb_channel_stats(ptr(y), ptr(stats), cll(y.numel() // c), c, dtype_code(y.dtype),
                                   stream_handle()), "channel_stats")
    return stats


def bn_relu_pool_fwd(y: torch.Tensor, stats: Optional[torch.Tensor], gamma, beta, running_mean, running_var,
                     nbt: Optional[torch.Tensor], pool: bool, training: bool = True,
                     out: Optional[torch.Tensor] = None, want_winner: bool = False):
    """Returns (out, save_mean, save_invstd) and, with ``want_winner``, the tensor of window winners
    (the conv output that won each ReLU + pool window, NaN where ReLU blocked it)."""
    b, h, w, c = y.shape
    oh, ow = (h // 2, w // 2) if pool else (h, w)
    if out is None:
        out = torch.empty((b, oh, ow, c), dtype=y.dtype, device=y.device)
    save_mean = torch.empty(c, dtype=torch.float32, device=y.device)
    save_invstd = torch.empty(c, dtype=torch.float32, device=y.device)
    winner = torch.empty_like(out) if want_winner else None
    check(lib().ddpb_bn_relu_pool_fwd_winner(ptr(y), ptr(stats), ptr(gamma), ptr(beta), ptr(out), ptr(save_mean),
                                             ptr(save_invstd), ptr(running_mean), ptr(running_var), ptr(nbt),
                                             ptr(winner), b, h, w, c, int(pool), int(training), cf(BN_EPS),
                                             cf(BN_MOMENTUM), dtype_code(y.dtype), stream_handle()),
          "bn_relu_pool_fwd")
    if want_winner:
        return out, save_mean, save_invstd, winner
    return out, save_mean, save_invstd


def bn_relu_pool_bwd(dout: torch.Tensor, y: torch.Tensor, save_mean, save_invstd, gamma, beta, pool: bool,
                     dbias: Optional[torch.Tensor] = None):
    b, h, w, c = y.shape
    dy = torch.empty_like(y)
    sums = torch.zeros(2 * c, dtype=torch.float32, device=y.device)
    dgamma = torch.empty(c, dtype=torch.float32, device=y.device)
    dbeta = torch.empty(c, dtype=torch.float32, device=y.device)
    check(lib().ddpb_bn_relu_pool_bwd(ptr(dout), ptr(y), ptr(save_mean), ptr(save_invstd), ptr(gamma), ptr(beta),
                                      ptr(sums), ptr(dy), ptr(dgamma), ptr(dbeta), ptr(dbias), b, h, w, c,
                                      int(pool), dtype_code(y.dtype), stream_handle()), "bn_relu_pool_bwd")
    return dy, dgamma, dbeta


def bn_relu_pool_bwd_apply(dout: torch.Tensor, y: torch.Tensor, save_mean, save_invstd, gamma, beta, sums, pool: bool,
                           dbias: Optional[torch.Tensor] = None):
    """Second BN-backward pass only; ``sums`` comes from ``conv_dgrad_bnred_tc``."""
    b, h, w, c = y.shape
    dy = torch.empty_like(y)
    dgamma = torch.empty(c, dtype=torch.float32, device=y.device)
    dbeta = torch.empty(c, dtype=torch.float32, device=y.device)
    check(lib().ddpb_bn_relu_pool_bwd_apply(ptr(dout), ptr(y), ptr(save_mean), ptr(save_invstd), ptr(gamma),
                                            ptr(beta), ptr(sums), ptr(dy), ptr(dgamma), ptr(dbeta), ptr(dbias), b, h,
                                            w, c, int(pool), dtype_code(y.dtype), stream_handle()),
          "bn_relu_pool_bwd_apply")
    return dy, dgamma, dbeta


def maxpool_fwd(x: torch.Tensor) -> torch.Tensor:
    b, h, w, c = x.shape
    out = torch.empty((b, h // 2, w // 2, c), dtype=x.dtype, device=x.device)
    check(lib().ddpb_maxpool2x2_fwd(ptr(x), ptr(out), b, h, w, c, dtype_code(x.dtype), stream_handle()), "pool")
    return out


def maxpool_bwd(x: torch.Tensor, dout: torch.Tensor) -> torch.Tensor:
    b, h, w, c = x.shape
    dx = torch.empty_like(x)
    check(lib().ddpb_maxpool2x2_bwd(ptr(x), ptr(dout), ptr(dx), b, h, w, c, dtype_code(x.dtype),
                                    stream_handle()), "pool_bwd")
    return dx


# ---- classifier + CE ------------------------------------------------------------------------
def fc_ce_fwd(act: torch.Tensor, w: torch.Tensor, bias: torch.Tensor, target: torch.Tensor,
              loss_divisor: Optional[int] = None, want_grad: bool = True):
    b = act.shape[0]
    dev = act.device
    logits = torch.empty((b, 10), dtype=torch.float32, device=dev)
    dlogits = torch.empty((b, 10), dtype=torch.float32, device=dev) if want_grad else None
    loss = torch.zeros(1, dtype=torch.float32, device=dev)
    correct = torch.zeros(1, dtype=torch.int32, device=dev)
    check(lib().ddpb_fc_ce_fwd(ptr(act), ptr(w), ptr(bias), ptr(target), ptr(logits), ptr(dlogits), ptr(loss),
                               ptr(correct), b, loss_divisor or b, dtype_code(act.dtype), stream_handle()),
          "fc_ce_fwd")
    return logits, dlogits, loss, correct


def fc_ce_fwd_dact(act: torch.Tensor, w: torch.Tensor, bias: torch.Tensor, target: torch.Tensor):
    """Loss kernel that also writes the gradient of its input (what ``fc_bwd`` returns as ``dact``)."""
    b = act.shape[0]
    dev = act.device
    logits = torch.empty((b, 10), dtype=torch.float32, device=dev)
    dlogits = torch.empty((b, 10), dtype=torch.float32, device=dev)
    loss = torch.zeros(1, dtype=torch.float32, device=dev)
    dact = torch.empty_like(act)
    check(lib().ddpb_fc_ce_fwd_dact(ptr(act), ptr(w), ptr(bias), ptr(target), ptr(logits), ptr(dlogits), ptr(loss), None,
                                    ptr(dact), b, b, dtype_code(act.dtype), stream_handle()), "fc_ce_fwd_dact")
    return logits, dlogits, loss, dact


def fc_bwd(act: torch.Tensor, w: torch.Tensor, dlogits: torch.Tensor, want_dact: bool = True):
    b = act.shape[0]
    dact = torch.empty_like(act) if want_dact else None
    dw = torch.zeros((10, 512), dtype=torch.float32, device=act.device)
    db = torch.zeros(10, dtype=torch.float32, device=act.device)
    check(lib().ddpb_fc_bwd(ptr(act), ptr(w), ptr(dlogits), ptr(dact), ptr(dw), ptr(db), b,
                            dtype_code(act.dtype), stream_handle()), "fc_bwd")
    return dact, dw, db


# ---- convolutions -----------------------------------------------------------------------------
def conv_fwd_simt(x, w_f32, bias, want_stats: bool = False):
    b, h, w_, cin = x.shape
    cout = w_f32.shape[0]
    y = torch.empty((b, h, w_, cout), dtype=x.dtype, device=x.device)
    stats = torch.zeros(2 * cout, dtype=torch.float32, device=x.device) if want_stats else None
    check(lib().ddpb_conv3x3_fwd_simt(ptr(x), ptr(w_f32), ptr(bias), ptr(y), ptr(stats), b, h, w_, cin, cout,
                                      dtype_code(x.dtype), stream_handle()), "conv_fwd_simt")
    return (y, stats) if want_stats else y


def conv_dgrad_simt(dy, w_f32):
    b, h, w_, cout = dy.shape
    cin = w_f32.shape[3]
    dx = torch.empty((b, h, w_, cin), dtype=dy.dtype, device=dy.device)
    check(lib().ddpb_conv3x3_dgrad_simt(ptr(dy), ptr(w_f32), ptr(dx), b, h, w_, cin, cout, dtype_code(dy.dtype),
                                        stream_handle()), "conv_dgrad_simt")
    return dx


def conv_wgrad_simt(dy, x):
    b, h, w_, cout = dy.shape
    cin = x.shape[3]
    dw = torch.zeros((cout, 3, 3, cin), dtype=torch.float32, device=dy.device)
    check(lib().ddpb_conv3x3_wgrad_simt(ptr(dy), ptr(x), ptr(dw), b, h, w_, cin, cout, dtype_code(dy.dtype),
                                        stream_handle()), "conv_wgrad_simt")
    return dw


def conv_fwd_tc(x, w_bf16, bias, want_stats: bool = False, tile_n: int = 0):
    b, h, w_, cin = x.shape
    cout = w_bf16.shape[0]
    y = torch.empty((b, h, w_, cout), dtype=torch.bfloat16, device=x.device)
    stats = torch.zeros(2 * cout, dtype=torch.float32, device=x.device) if want_stats else None
    check(lib().ddpb_conv3x3_fwd_tc(ptr(x), ptr(w_bf16), ptr(bias), ptr(y), ptr(stats), b, h, w_, cin, cout,
                                    tile_n, stream_handle()), "conv_fwd_tc")
    return (y, stats) if want_stats else y


def conv_fwd_bn_tc(x, w_bf16, bias, gamma, beta, running_mean, running_var, nbt, pool: bool, tile_n: int = 0):
    """Forward conv with this block's BatchNorm + ReLU + pool fused behind a grid barrier.
    Returns dict(y, stats, act, winner, save_mean, save_invstd, fused)."""
    b, h, w_, cin = x.shape
    cout = w_bf16.shape[0]
    oh, ow = (h // 2, w_ // 2) if pool else (h, w_)
    dev = x.device
    y = torch.empty((b, h, w_, cout), dtype=torch.bfloat16, device=dev)
    stats = torch.zeros(2 * cout, dtype=torch.float32, device=dev)
    act = torch.empty((b, oh, ow, cout), dtype=torch.bfloat16, device=dev)
    winner = torch.empty_like(act)
    save_mean = torch.empty(cout, dtype=torch.float32, device=dev)
    save_invstd = torch.empty(cout, dtype=torch.float32, device=dev)
    barrier = torch.zeros(1, dtype=torch.int32, device=dev)
    fused = ctypes.c_int(0)
    check(lib().ddpb_conv3x3_fwd_bn_tc(ptr(x), ptr(w_bf16), ptr(bias), ptr(y), ptr(stats), b, h, w_, cin, cout, tile_n,
                                       ptr(gamma), ptr(beta), ptr(act), ptr(winner), ptr(save_mean), ptr(save_invstd),
                                       ptr(running_mean), ptr(running_var), ptr(nbt), int(pool), cf(BN_EPS),
                                       cf(BN_MOMENTUM), ptr(barrier), ctypes.byref(fused), stream_handle()),
          "conv_fwd_bn_tc")
    return dict(y=y, stats=stats, act=act, winner=winner, save_mean=save_mean, save_invstd=save_invstd,
                fused=bool(fused.value))


def conv_dgrad_tc(dy, w_bf16, tile_n: int = 0):
    b, h, w_, cout = dy.shape
    cin = w_bf16.shape[3]
    dx = torch.empty((b, h, w_, cin), dtype=torch.bfloat16, device=dy.device)
    check(lib().ddpb_conv3x3_dgrad_tc(ptr(dy), ptr(w_bf16), ptr(dx), b, h, w_, cin, cout, tile_n,
                                      stream_handle()), "conv_dgrad_tc")
    return dx


def conv_dgrad_bnred_tc(dy, w_bf16, y_win, save_mean, save_invstd, gamma, beta, tile_n: int = 0):
    """dgrad whose epilogue also reduces the previous block's BN backward sums.  ``y_win``: the previous
    block's window winners (``bn_relu_pool_fwd(..., want_winner=True)``), or its conv output when it
    does not pool.  Returns (dx, sums[2*Cin])."""
    b, h, w_, cout = dy.shape
    cin = w_bf16.shape[3]
    assert tuple(y_win.shape) == (b, h, w_, cin)
    dx = torch.empty((b, h, w_, cin), dtype=torch.bfloat16, device=dy.device)
    sums = torch.zeros(2 * cin, dtype=torch.float32, device=dy.device)
    check(lib().ddpb_conv3x3_dgrad_bnred_tc(ptr(dy), ptr(w_bf16), ptr(dx), b, h, w_, cin, cout, tile_n, ptr(y_win),
                                            ptr(save_mean), ptr(save_invstd), ptr(gamma), ptr(beta), ptr(sums),
                                            stream_handle()), "conv_dgrad_bnred_tc")
    return dx, sums


def conv_wgrad_tc(dy, x, split_k: int = 0, out: Optional[torch.Tensor] = None):
    b, h, w_, cout = dy.shape
    cin = x.shape[3]
    dw = out if out is not None else torch.zeros((cout, 3, 3, cin), dtype=torch.float32, device=dy.device)
    check(lib().ddpb_conv3x3_wgrad_tc(ptr(dy), ptr(x), ptr(dw), b, h, w_, cin, cout, split_k, stream_handle()),
          "conv_wgrad_tc")
    return dw


def conv_wgrad_tc_range(dy, x, co_lo: int, co_n: int, out: torch.Tensor, split_k: int = 0):
    """Weight gradient of output channels [co_lo, co_lo + co_n) only, accumulated into rows co_lo.. of ``out``."""
    b, h, w_, cout = dy.shape
    cin = x.shape[3]
    check(lib().ddpb_conv3x3_wgrad_tc_range(ptr(dy), ptr(x), ptr(out), b, h, w_, cin, cout, co_lo, co_n, split_k,
                                            stream_handle()), "conv_wgrad_tc_range")
    return out


def first_layer_fwd_tc(x, w_bf16, bias, want_stats: bool = False, generic_kernel: bool = False):
    """conv3x3(3 -> Cout) on the tensor cores: im2col to [pixels][64], then a 1x1 convolution.
    Returns (y, stats | None, cols) - ``cols`` is what the weight gradient needs later."""
    b, h, w_, cin = x.shape
    assert cin == 3 and x.dtype == torch.bfloat16
    cout = w_bf16.shape[0]
    cols = torch.empty((b, h, w_, 64), dtype=torch.bfloat16, device=x.device)
    w_pad = torch.empty((cout, 64), dtype=torch.bfloat16, device=x.device)
    y = torch.empty((b, h, w_, cout), dtype=torch.bfloat16, device=x.device)
    stats = torch.zeros(2 * cout, dtype=torch.float32, device=x.device) if want_stats else None
    check(lib().ddpb_im2col3x3_c3(ptr(x), ptr(w_bf16), ptr(cols), ptr(w_pad), b, h, w_, cout, 1, stream_handle()),
          "im2col3x3_c3")
    if generic_kernel:       # the conv kernel with a single tap (one CTA per 128-pixel tile)
        check(lib().ddpb_conv1x1_fwd_tc(ptr(cols), ptr(w_pad), ptr(bias), ptr(y), ptr(stats), b, h, w_, 64, cout,
                                        stream_handle()), "conv1x1_fwd_tc")
    else:                    # the multi-tile GEMM kernel the engine uses
        check(lib().ddpb_gemm_rows_fwd_tc(ptr(cols), ptr(w_pad), ptr(bias), ptr(y), ptr(stats),
                                          ctypes.c_longlong(b * h * w_), cout, stream_handle()), "gemm_rows_fwd_tc")
    return y, stats, cols


def im2col3x3_c3(x, w_bf16, zero_pad: bool = True, cols: Optional[torch.Tensor] = None):
    """Patch matrix [pixels][64] (27 real columns) of a 3-channel NHWC image and the zero-padded [Cout][64] weight
    matrix.  ``zero_pad=False`` leaves columns 32..63 alone (the engine's mode: it clears them once; a 32 x 32 image
    then takes the eight-rows-per-CTA kernel) - pass a ``cols`` buffer whose upper half is already zero."""
    b, h, w_, cin = x.shape
    assert cin == 3 and x.dtype == torch.bfloat16
    cout = w_bf16.shape[0]
    if cols is None:
        cols = (torch.empty if zero_pad else torch.zeros)((b, h, w_, 64), dtype=torch.bfloat16, device=x.device)
    w_pad = torch.empty((cout, 64), dtype=torch.bfloat16, device=x.device)
    check(lib().ddpb_im2col3x3_c3(ptr(x), ptr(w_bf16), ptr(cols), ptr(w_pad), b, h, w_, cout, 1 if zero_pad else 0,
                                  stream_handle()), "im2col3x3_c3")
    return cols, w_pad


def first_layer_wgrad_tc(dy, cols, split_k: int = 0):
    b, h, w_, cout = dy.shape
    dw = torch.zeros((cout, 3, 3, 3), dtype=torch.float32, device=dy.device)
    check(lib().ddpb_patch_wgrad_tc(ptr(dy), ptr(cols), ptr(dw), b, h, w_, cout, 27, 27, split_k, stream_handle()),
          "patch_wgrad_tc")
    return dw


# ---- P2P sync ---------------------------------------------------------------------------------
def _ptr_table(ptrs: Optional[Sequence[int]]):
    if ptrs is None:
        return None
    arr = (ctypes.c_void_p * len(ptrs))(*[ctypes.c_void_p(int(p)) for p in ptrs])
    return arr


def p2p_sync(mode: int, grad_ptrs: Sequence[int], master_ptrs: Optional[Sequence[int]],
             shadow_ptrs: Optional[Sequence[int]], flag_ptrs: Optional[Sequence[int]], mom: Optional[torch.Tensor],
             epochs: Optional[torch.Tensor], rank: int, world: int, slot: int, off: int, n: int, lr: float = 0.0,
             momentum: float = 0.0, wd: float = 0.0, blocks: int = 0, stream=None,
             shard: Optional[Sequence[int]] = None) -> None:
    """``shard``: flat [lo0, hi0, lo1, hi1, ...] element ranges whose fp32 master stays with its owner."""
    n_sh = len(shard) // 2 if shard else 0
    tab = (ctypes.c_longlong * (2 * n_sh))(*shard) if n_sh else None
    check(lib().ddpb_p2p_sync_sharded(mode, _ptr_table(grad_ptrs), _ptr_table(master_ptrs), _ptr_table(shadow_ptrs),
                                      _ptr_table(flag_ptrs), ptr(mom), ptr(epochs), rank, world, slot, cll(off), cll(n),
                                      cf(lr), cf(momentum), cf(wd), blocks, tab, n_sh, stream_handle(stream)),
          "p2p_sync")


def first_layer_bwd_fused(dout, y, x_cols, save_mean, save_invstd, gamma, beta, sums, ctas: int = 0):
    """Block 0 backward in one kernel (BN + ReLU + pool backward apply fused with the 3 -> 64 weight gradient).
    Returns (dw [64, 27] fp32, dgamma, dbeta, dbias)."""
    b, h, w_, c = y.shape
    dev = y.device
    dw = torch.zeros((c, 27), dtype=torch.float32, device=dev)
    dgamma = torch.empty(c, dtype=torch.float32, device=dev)
    dbeta = torch.empty(c, dtype=torch.float32, device=dev)
    dbias = torch.zeros(c, dtype=torch.float32, device=dev)
    check(lib().ddpb_first_layer_bwd_fused(ptr(dout), ptr(y), ptr(x_cols), ptr(save_mean), ptr(save_invstd), ptr(gamma),
                                           ptr(beta), ptr(sums), ptr(dw), ptr(dgamma), ptr(dbeta), ptr(dbias), b, h, w_, c,
                                           27, 27, ctas, stream_handle()), "first_layer_bwd_fused")
    return dw, dgamma, dbeta, dbias


# ---- fp32 convolutions on the tensor cores (tcgen05 kind::tf32) ---------------------------------------------
def conv_fwd_tf32(x, w_krsc, bias, want_stats: bool = False):
    b, h, w_, cin = x.shape
    cout = w_krsc.shape[0]
    y = torch.empty((b, h, w_, cout), dtype=torch.float32, device=x.device)
    stats = torch.zeros(2 * cout, dtype=torch.float32, device=x.device) if want_stats else None
    check(lib().ddpb_conv3x3_fwd_tf32(ptr(x), ptr(w_krsc), ptr(bias), ptr(y), ptr(stats), b, h, w_, cin, cout,
                                      stream_handle()), "conv_fwd_tf32")
    return (y, stats) if want_stats else y


def conv_dgrad_tf32(dy, w_krsc):
    b, h, w_, cout = dy.shape
    cin = w_krsc.shape[3]
    dx = torch.empty((b, h, w_, cin), dtype=torch.float32, device=dy.device)
    check(lib().ddpb_conv3x3_dgrad_tf32(ptr(dy), ptr(w_krsc), ptr(dx), b, h, w_, cin, cout, stream_handle()),
          "conv_dgrad_tf32")
    return dx


def conv_wgrad_tf32(dy, x, split_k: int = 0):
    b, h, w_, cout = dy.shape
    cin = x.shape[3]
    dw = torch.zeros((cout, 3, 3, cin), dtype=torch.float32, device=dy.device)
    check(lib().ddpb_conv3x3_wgrad_tf32(ptr(dy), ptr(x), ptr(dw), b, h, w_, cin, cout, split_k, stream_handle()),
          "conv_wgrad_tf32")
    return dw
