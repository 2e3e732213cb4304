"""The allocation-free BN-backward entry point, exercised through the raw C ABI.

`ddpb_bn_relu_pool_bwd_noalloc` exists for callers that cannot bring their own scratch buffer (a
host whose memory pool owns the whole GPU, a plug-in slot with inputs and outputs only): its [2*C]
reduction scratch is a fixed device array inside the library.  That array is shared by every call
and keeps whatever the previous call left in it, so the checks run TWICE in a row with different
data: the second call only matches if the entry point really clears the part it uses.

This file sorts last on purpose: if it fails, a `pytest -x` run has already executed every other
test.  It must also pass on its own in a fresh process (nothing here relies on earlier tests).
"""
import ctypes
import pathlib

import pytest
import torch

pytestmark = pytest.mark.gpu

LIB = pathlib.Path(__file__).resolve().parent.parent / "cs744-distributed-data-parallel_b200" / "_native" / "libddpb200.so"
BF16, F32 = 1, 0
MAX_C = 2048          # DDPB_BN_NOALLOC_MAX_C in include/ddp_b200.h (bf16)
MAX_C_F32 = 1024      # DDPB_BN_NOALLOC_MAX_C_F32: a BN-kernel thread owns 16 bytes of channels


def _lib():
    if not LIB.exists():
        from cs744_ddp_b200.ops import native
        native.build()
    lib = ctypes.CDLL(str(LIB))
    lib.ddpb_last_error.restype = ctypes.c_char_p
    return lib


def _p(t):
    return ctypes.c_void_p(t.data_ptr() if t is not None else 0)


def _case(seed, c, hw, pool, dtype, b=6):
    g = torch.Generator(device="cuda").manual_seed(seed)
    y = (torch.randn(b, hw, hw, c, device="cuda", generator=g) * 1.5 + 0.3).to(dtype)
    ohw = hw // 2 if pool else hw
    dout = torch.randn(b, ohw, ohw, c, device="cuda", generator=g).to(dtype)
    gamma = torch.rand(c, device="cuda", generator=g) + 0.5
    beta = torch.randn(c, device="cuda", generator=g) * 0.2
    yf = y.float().reshape(-1, c)
    mean = yf.mean(0)
    invstd = (yf.var(0, unbiased=False) + 1e-5).rsqrt()
    return y, dout, gamma, beta, mean.contiguous(), invstd.contiguous()


def _run(lib, entry, case, pool, dtype, with_scratch):
    y, dout, gamma, beta, mean, invstd = case
    b, h, w, c = y.shape
    dy = torch.empty_like(y)
    dgamma = torch.empty(c, device="cuda")
    dbeta = torch.empty(c, device="cuda")
    dbias = torch.zeros(c, device="cuda")
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    code = BF16 if dtype == torch.bfloat16 else F32
    if with_scratch:
        sums = torch.zeros(2 * c, device="cuda")
        rc = entry(_p(dout), _p(y), _p(mean), _p(invstd), _p(gamma), _p(beta), _p(sums), _p(dy), _p(dgamma),
                   _p(dbeta), _p(dbias), b, h, w, c, int(pool), code, stream)
    else:
        rc = entry(_p(dout), _p(y), _p(mean), _p(invstd), _p(gamma), _p(beta), _p(dy), _p(dgamma), _p(dbeta),
                   _p(dbias), b, h, w, c, int(pool), code, stream)
    assert rc == 0, lib.ddpb_last_error()
    torch.cuda.synchronize()
    return dy, dgamma, dbeta, dbias


def _rel(a, b):
    return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-12))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_noalloc_matches_workspace_entry_point_twice_in_a_row(dtype):
    lib = _lib()
    # two different problems back to back: the second one sees the first one's leftovers in the
    # built-in scratch unless the entry point clears it
    for seed, (c, hw, pool) in enumerate([(512, 4, True), (128, 16, False)]):
        case = _case(100 + seed, c, hw, pool, dtype)
        ref = _run(lib, lib.ddpb_bn_relu_pool_bwd, case, pool, dtype, with_scratch=True)
        got = _run(lib, lib.ddpb_bn_relu_pool_bwd_noalloc, case, pool, dtype, with_scratch=False)
        tol = 2e-4 if dtype == torch.float32 else 1.5e-2          # same tolerances as the BN kernel tests
        for name, a, b in zip(("dy", "dgamma", "dbeta"), got, ref):
            assert _rel(a, b) < tol, (seed, name, _rel(a, b))
        scale = max(1.0, float(ref[0].float().abs().sum(dim=(0, 1, 2)).max()))
        assert float((got[3] - ref[3]).abs().max()) < 1e-2 * scale      # conv-bias gradient (numerical noise)


def test_noalloc_rejects_more_channels_than_the_builtin_scratch_holds():
    lib = _lib()
    c = MAX_C + 8
    t = torch.zeros(8, device="cuda")
    rc = lib.ddpb_bn_relu_pool_bwd_noalloc(_p(t), _p(t), _p(t), _p(t), _p(t), _p(t), _p(t), _p(t), _p(t), _p(t),
                                           1, 2, 2, c, 0, F32, ctypes.c_void_p(0))
    assert rc != 0
    assert b"exceeds the built-in scratch capacity" in lib.ddpb_last_error()


def test_noalloc_names_the_fp32_limit_and_leaves_the_scratch_alone():
    """fp32 tensors stop at 1024 channels (half of what the array could hold).  The call must say so,
    and - like every rejected call - must not clear the shared scratch on its way out."""
    lib = _lib()
    t = torch.zeros(8, device="cuda")
    for c, dtype, needle in ((MAX_C, F32, b"exceeds the fp32 limit of 1024"),
                             (MAX_C_F32 + 4, F32, b"exceeds the fp32 limit of 1024"),
                             (24, BF16, b"must be a power of two"),
                             (64, 7, b"unknown dtype")):
        rc = lib.ddpb_bn_relu_pool_bwd_noalloc(_p(t), _p(t), _p(t), _p(t), _p(t), _p(t), _p(t), _p(t), _p(t), _p(t),
                                               1, 2, 2, c, 0, dtype, ctypes.c_void_p(0))
        assert rc != 0, (c, dtype)
        assert needle in lib.ddpb_last_error(), (c, dtype, lib.ddpb_last_error())
    # the largest sizes that ARE accepted, checked against torch's own BatchNorm backward (fp32 oracle)
    for c, dtype in ((MAX_C_F32, torch.float32), (MAX_C, torch.bfloat16)):
        case = _case(7, c, 2, False, dtype, b=8)
        y, dout, gamma, beta, mean, invstd = case
        dy, dgamma, dbeta, _ = _run(lib, lib.ddpb_bn_relu_pool_bwd_noalloc, case, False, dtype, with_scratch=False)
        yf = y.float().clone().requires_grad_(True)
        g = gamma.clone().requires_grad_(True)
        bt = beta.clone().requires_grad_(True)
        x = yf.permute(0, 3, 1, 2)
        out = torch.relu(torch.nn.functional.batch_norm(x, None, None, g, bt, True, 0.1, 1e-5))
        out.backward(dout.float().permute(0, 3, 1, 2))
        tol = 2e-4 if dtype == torch.float32 else 1.5e-2
        assert _rel(dy, yf.grad) < tol and _rel(dgamma, g.grad) < tol and _rel(dbeta, bt.grad) < tol
