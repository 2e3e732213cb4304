"""Numerics of every native kernel against a plain PyTorch fp32 reference of the same op."""
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda"


@pytest.fixture(scope="module")
def F():
    from cs744_ddp_b200.ops import functional
    return functional


@pytest.fixture(scope="module")
def R():
    from cs744_ddp_b200.ops import reference
    return reference


def _rel_err(a, b):
    a, b = a.float(), b.float()
    return float((a - b).norm() / (b.norm() + 1e-12))


def _report(name, got, ref, tol):
    err = _rel_err(got, ref)
    maxabs = float((got.float() - ref.float()).abs().max())
    print(f"[num] {name}: rel_l2={err:.3e} max_abs={maxabs:.3e} ref_max={float(ref.abs().max()):.3e}")
    assert err < tol, f"{name}: rel l2 error {err:.3e} >= {tol}"


# ----------------------------------------------------------------------------------------------
def test_native_library_is_loaded(F):
    from cs744_ddp_b200.ops import native
    assert native.available()
    maps = open("/proc/self/maps").read()
    assert "libddpb200.so" in maps


def test_augment_matches_torch_pipeline(F):
    from cs744_ddp_b200.data import synthetic_cifar10, transform_batch
    from cs744_ddp_b200.data.pipeline import RawBatch
    split = synthetic_cifar10(64, seed=1).to(DEV)
    g = torch.Generator().manual_seed(0)
    idx = torch.randperm(64, generator=g)[:40].to(DEV)
    crop = torch.randint(0, 9, (40, 2), generator=g, dtype=torch.int32).to(DEV)
    flip = (torch.rand(40, generator=g) < 0.5).to(torch.uint8).to(DEV)
    ref = transform_batch(RawBatch(split.images[idx], split.labels[idx], crop, flip), layout="nhwc")
    out = F.augment(split.images, idx, crop, flip, torch.float32)
    assert torch.allclose(out, ref, atol=1e-6)
    out16 = F.augment(split.images, idx, crop, flip, torch.bfloat16)
    assert torch.allclose(out16.float(), ref, atol=2e-2, rtol=1e-2)
    plain = F.augment(split.images, None, None, None, torch.float32, batch=64)
    ref2 = transform_batch(RawBatch(split.images, split.labels, None, None), layout="nhwc")
    assert torch.allclose(plain, ref2, atol=1e-6)
    # the targets of the batch ride along in the same launch (the engine's staged-batch path)
    dst = torch.full((80,), -1, dtype=torch.int64, device=DEV)
    lab = split.labels[idx].contiguous()
    out2 = F.augment(split.images, idx, crop, flip, torch.float32, labels=(lab, dst))
    assert torch.equal(out2, out) and torch.equal(dst[:40], lab) and bool((dst[40:] == -1).all())


def test_sgd_matches_torch_optim(F):
    n = 64 * 100
    p0 = torch.randn(n, device=DEV)
    p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.SGD([p], lr=0.1, momentum=0.9, weight_decay=1e-4)
    mine, mom = p0.clone(), torch.zeros(n, device=DEV)
    shadow = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    for step in range(4):
        g = torch.randn(n, device=DEV)
        p.grad = g.clone()
        opt.step()
        F.sgd_step(mine, mom, g, shadow, 0.1, 0.9, 1e-4)
        assert torch.allclose(mine, p.data, rtol=1e-5, atol=1e-6), step
    assert torch.equal(shadow, mine.to(torch.bfloat16))
    assert torch.allclose(mom, opt.state[p]["momentum_buffer"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("c,hw,pool", [(64, 32, True), (256, 8, False), (512, 2, True), (128, 16, True)])
def test_bn_relu_pool_forward_backward(F, R, dtype, c, hw, pool):
    torch.manual_seed(0)
    b = 6
    y = (torch.randn(b, hw, hw, c, device=DEV) * 1.5 + 0.3).to(dtype)
    gamma = torch.rand(c, device=DEV) + 0.5
    beta = torch.randn(c, device=DEV) * 0.2
    rm, rv = torch.randn(c, device=DEV) * 0.1, torch.rand(c, device=DEV) + 0.5
    nbt = torch.zeros(1, dtype=torch.int64, device=DEV)
    stats = F.channel_stats(y)
    _report("channel_stats", stats, R.channel_stats(y), 1e-5)
    rm_n, rv_n = rm.clone(), rv.clone()
    out, mean, invstd = F.bn_relu_pool_fwd(y, stats, gamma, beta, rm_n, rv_n, nbt, pool, True)
    ref_out, ref_rm, ref_rv = R.bn_relu_pool_fwd(y, gamma, beta, rm, rv, pool, True)
    tol = 1e-5 if dtype == torch.float32 else 6e-3
    _report("bn_fwd", out, ref_out, tol)
    _report("running_mean", rm_n, ref_rm, 1e-5)
    _report("running_var", rv_n, ref_rv, 1e-5)
    assert int(nbt) == 1
    # eval mode uses the running statistics
    out_e, _, _ = F.bn_relu_pool_fwd(y, None, gamma, beta, rm_n, rv_n, None, pool, False)
    ref_e, _, _ = R.bn_relu_pool_fwd(y, gamma, beta, rm_n, rv_n, pool, False)
    _report("bn_eval", out_e, ref_e, tol)
    # backward
    dout = torch.randn_like(out.float()).to(dtype)
    dbias = torch.zeros(c, device=DEV)
    dy, dgamma, dbeta = F.bn_relu_pool_bwd(dout, y, mean, invstd, gamma, beta, pool, dbias)
    rdy, rdg, rdb = R.bn_relu_pool_bwd(dout, y, gamma, beta, pool)
    btol = 2e-4 if dtype == torch.float32 else 1.5e-2
    _report("bn_bwd_dy", dy, rdy, btol)
    _report("bn_bwd_dgamma", dgamma, rdg, btol)
    _report("bn_bwd_dbeta", dbeta, rdb, btol)
    # conv-bias gradient = column sums of dy (mathematically ~0)
    assert float(dbias.abs().max()) < 1e-2 * max(1.0, float(dy.float().abs().sum(dim=(0, 1, 2)).max()))


def test_maxpool_standalone(F):
    x = torch.randn(4, 8, 8, 64, device=DEV)
    out = F.maxpool_fwd(x)
    ref = torch.nn.functional.max_pool2d(x.permute(0, 3, 1, 2), 2, 2).permute(0, 2, 3, 1)
    assert torch.equal(out, ref.contiguous())
    dout = torch.randn_like(out)
    xr = x.clone().requires_grad_(True)
    torch.nn.functional.max_pool2d(xr.permute(0, 3, 1, 2), 2, 2).backward(dout.permute(0, 3, 1, 2))
    assert torch.equal(F.maxpool_bwd(x, dout), xr.grad)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("b", [256, 10])
def test_fc_cross_entropy(F, R, dtype, b):
    torch.manual_seed(1)
    act = torch.randn(b, 512, device=DEV).to(dtype)
    w = torch.randn(10, 512, device=DEV) * 0.05
    bias = torch.randn(10, device=DEV) * 0.1
    tgt = torch.randint(0, 10, (b,), device=DEV)
    logits, dlogits, loss, correct = F.fc_ce_fwd(act, w, bias, tgt)
    rl, rloss, rdl, rda, rdw, rdb, rc = R.fc_ce(act, w, bias, tgt)
    _report("fc_logits", logits, rl, 1e-5)
    assert abs(float(loss) - float(rloss)) < 1e-4 * max(1.0, abs(float(rloss)))
    _report("fc_dlogits", dlogits, rdl, 1e-4)
    assert int(correct) == int(rc)
    dact, dw, db = F.fc_bwd(act, w, dlogits)
    tol = 1e-4 if dtype == torch.float32 else 6e-3
    _report("fc_dact", dact, rda, tol)
    _report("fc_dw", dw, rdw, 1e-4)
    _report("fc_db", db, rdb, 1e-4)
    # the loss kernel can produce dact itself (the engine's path: backward then starts with the last conv block):
    # same values as fc_bwd's, and fc_bwd without dact leaves the weight / bias gradients unchanged
    logits2, dlogits2, loss2, dact2 = F.fc_ce_fwd_dact(act, w, bias, tgt)
    assert torch.equal(logits2, logits) and torch.equal(dlogits2, dlogits)
    assert torch.equal(dact2, dact)
    none, dw2, db2 = F.fc_bwd(act, w, dlogits, want_dact=False)
    assert none is None
    _report("fc_dw (no dact)", dw2, rdw, 1e-4)
    _report("fc_db (no dact)", db2, rdb, 1e-4)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("b,hw,cin,cout", [(4, 32, 3, 64), (3, 8, 16, 32), (5, 4, 32, 16)])
def test_conv_simt(F, R, dtype, b, hw, cin, cout):
    torch.manual_seed(2)
    x = torch.randn(b, hw, hw, cin, device=DEV).to(dtype)
    w = torch.randn(cout, 3, 3, cin, device=DEV) * 0.1
    bias = torch.randn(cout, device=DEV) * 0.1
    tol = 1e-5 if dtype == torch.float32 else 6e-3
    y, stats = F.conv_fwd_simt(x, w, bias, want_stats=True)
    _report("simt_stats", stats, R.channel_stats(y), 1e-4)
    _report("simt_fwd", y, R.conv_fwd(x, w, bias), tol)
    dy = torch.randn(b, hw, hw, cout, device=DEV).to(dtype)
    _report("simt_dgrad", F.conv_dgrad_simt(dy, w), R.conv_dgrad(dy, w), tol)
    _report("simt_wgrad", F.conv_wgrad_simt(dy, x), R.conv_wgrad(dy, x), 1e-4)


def _tile_ok(code, channels, hw):
    """Can this tile code run a layer with `channels` output channels on hw x hw images?"""
    if code < 1000:
        return channels % code == 0
    if code > 60000:          # CTA pairs: 6<splits><tile>
        return channels % (code % 1000) == 0
    if code > 4000:           # cluster split-K: 4<splits><tile>
        return channels % (64 if code % 100 == 64 else 128) == 0
    if code > 2000:
        return channels % (code - 2000) == 0
    nc, tile = (code // 100) % 10, (64 if code % 100 == 64 else 128)
    if channels % tile or (channels // tile) % nc:
        return False
    bh, bn = (128 // hw, 1) if hw * hw >= 128 else (hw, 128 // (hw * hw))
    return bn % nc == 0 or (bn == 1 and bh % nc == 0)


TILE_CODES = [64, 128, 1264, 1464, 1864, 1228, 1428, 2128, 2256, 4264, 4464, 4228, 4428, 61128, 61256, 62128, 62256]

TC_SHAPES = [  # (hw, cin, cout) of VGG-11 blocks 1..7
    (16, 64, 128), (8, 128, 256), (8, 256, 256), (4, 256, 512), (4, 512, 512), (2, 512, 512)]


@pytest.mark.parametrize("b", [16, 10])
@pytest.mark.parametrize("hw,cin,cout", TC_SHAPES)
@pytest.mark.parametrize("tile_n", TILE_CODES)
def test_conv_tc_forward(F, R, b, hw, cin, cout, tile_n):
    if not _tile_ok(tile_n, cout, hw):
        pytest.skip("tile code does not fit this layer")
    torch.manual_seed(3)
    x = torch.randn(b, hw, hw, cin, device=DEV).to(torch.bfloat16)
    w = (torch.randn(cout, 3, 3, cin, device=DEV) * 0.05).to(torch.bfloat16)
    bias = torch.randn(cout, device=DEV) * 0.1
    y, stats = F.conv_fwd_tc(x, w, bias, want_stats=True, tile_n=tile_n)
    torch.cuda.synchronize()
    ref = R.conv_fwd(x, w, bias)
    _report(f"tc_fwd hw{hw} {cin}->{cout} b{b} n{tile_n}", y, ref, 6e-3)
    _report("tc_fwd_stats", stats, R.channel_stats(y), 1e-3)


@pytest.mark.parametrize("b", [16, 10])
@pytest.mark.parametrize("hw,cin,cout", TC_SHAPES)
@pytest.mark.parametrize("tile_n", TILE_CODES)
def test_conv_tc_dgrad(F, R, b, hw, cin, cout, tile_n):
    if not _tile_ok(tile_n, cin, hw):
        pytest.skip("tile code does not fit this layer")
    torch.manual_seed(4)
    dy = torch.randn(b, hw, hw, cout, device=DEV).to(torch.bfloat16)
    w = (torch.randn(cout, 3, 3, cin, device=DEV) * 0.05).to(torch.bfloat16)
    dx = F.conv_dgrad_tc(dy, w, tile_n=tile_n)
    torch.cuda.synchronize()
    _report(f"tc_dgrad hw{hw} {cin}->{cout} b{b} n{tile_n}", dx, R.conv_dgrad(dy, w), 6e-3)


@pytest.mark.parametrize("b", [16, 10])
@pytest.mark.parametrize("pool", [False, True])
@pytest.mark.parametrize("hw,cin,cout", TC_SHAPES)
@pytest.mark.parametrize("tile_n", [0, 64, 128, 1264, 4264, 4464, 4228, 4428])
def test_conv_tc_forward_fused_bn(F, R, b, hw, cin, cout, tile_n, pool):
    """conv + grid barrier + BatchNorm/ReLU/pool in one kernel == the two separate kernels."""
    if tile_n and not _tile_ok(tile_n, cout, hw):
        pytest.skip("tile code does not fit this layer")
    torch.manual_seed(22)
    x = torch.randn(b, hw, hw, cin, device=DEV).to(torch.bfloat16)
    w = (torch.randn(cout, 3, 3, cin, device=DEV) * 0.05).to(torch.bfloat16)
    bias = torch.randn(cout, device=DEV) * 0.1
    gamma, beta = torch.rand(cout, device=DEV) + 0.5, torch.randn(cout, device=DEV) * 0.2
    rm1, rv1 = torch.zeros(cout, device=DEV), torch.ones(cout, device=DEV)
    rm2, rv2 = rm1.clone(), rv1.clone()
    nbt1 = torch.zeros(1, dtype=torch.int64, device=DEV)
    nbt2 = nbt1.clone()
    got = F.conv_fwd_bn_tc(x, w, bias, gamma, beta, rm1, rv1, nbt1, pool, tile_n=tile_n)
    if not got["fused"]:
        pytest.skip("grid does not fit on the GPU at once: the launcher kept the kernels separate")
    y, stats = F.conv_fwd_tc(x, w, bias, want_stats=True, tile_n=tile_n)
    act, mean, invstd, win = F.bn_relu_pool_fwd(y, stats, gamma, beta, rm2, rv2, nbt2, pool, want_winner=True)
    torch.cuda.synchronize()
    assert torch.equal(got["y"], y)
    _report(f"fusedbn stats hw{hw} {cin}->{cout} b{b} n{tile_n}", got["stats"], stats, 1e-5)
    _report("fusedbn mean", got["save_mean"], mean, 1e-4)
    _report("fusedbn invstd", got["save_invstd"], invstd, 1e-4)
    _report("fusedbn act", got["act"], act, 4e-3)
    same = torch.isnan(got["winner"].float()) == torch.isnan(win.float())
    assert same.float().mean().item() > 0.999       # ReLU decisions can flip on values that round to zero
    _report("fusedbn running_mean", rm1, rm2, 1e-4)
    _report("fusedbn running_var", rv1, rv2, 1e-4)
    assert int(nbt1) == int(nbt2) == 1


@pytest.mark.parametrize("b", [16, 10])
@pytest.mark.parametrize("pool", [False, True])
@pytest.mark.parametrize("hw,cin,cout", TC_SHAPES)
@pytest.mark.parametrize("tile_n", [0, 64, 128, 1264, 4264, 4464, 4228, 4428, 61128, 61256, 62128, 62256])
def test_conv_tc_dgrad_fused_bn_reduce(F, R, b, hw, cin, cout, tile_n, pool):
    """dgrad epilogue also produces the previous block's BN-backward sums; the apply pass finishes the job."""
    if tile_n and not _tile_ok(tile_n, cin, hw):
        pytest.skip("tile code does not fit this layer")
    torch.manual_seed(21)
    dy = torch.randn(b, hw, hw, cout, device=DEV).to(torch.bfloat16)
    w = (torch.randn(cout, 3, 3, cin, device=DEV) * 0.05).to(torch.bfloat16)
    hy = hw * 2 if pool else hw                                  # the previous block's conv output
    y_prev = torch.randn(b, hy, hy, cin, device=DEV).to(torch.bfloat16)
    gamma, beta = torch.rand(cin, device=DEV) + 0.5, torch.randn(cin, device=DEV) * 0.2
    rm, rv = torch.zeros(cin, device=DEV), torch.ones(cin, device=DEV)
    _, mean, invstd, win = F.bn_relu_pool_fwd(y_prev, R.channel_stats(y_prev), gamma, beta, rm, rv, None, pool,
                                              want_winner=True)
    dx, sums = F.conv_dgrad_bnred_tc(dy, w, win if pool else y_prev, mean, invstd, gamma, beta, tile_n=tile_n)
    dyp, dgamma, dbeta = F.bn_relu_pool_bwd_apply(dx, y_prev, mean, invstd, gamma, beta, sums, pool)
    torch.cuda.synchronize()
    _report(f"bnred dx hw{hw} {cin}->{cout} b{b} n{tile_n}", dx, R.conv_dgrad(dy, w), 6e-3)
    dy2, dgamma2, dbeta2 = F.bn_relu_pool_bwd(dx, y_prev, mean, invstd, gamma, beta, pool)      # two-pass path
    torch.cuda.synchronize()
    _report("bnred dbeta", dbeta, dbeta2, 1e-4)
    _report("bnred dgamma", dgamma, dgamma2, 1e-4)
    _report("bnred dy", dyp, dy2, 2e-3)


@pytest.mark.parametrize("b", [16, 10])
@pytest.mark.parametrize("hw,cin,cout", TC_SHAPES)
@pytest.mark.parametrize("split_k", [0, 1, 3])
def test_conv_tc_wgrad(F, R, b, hw, cin, cout, split_k):
    torch.manual_seed(5)
    dy = torch.randn(b, hw, hw, cout, device=DEV).to(torch.bfloat16)
    x = torch.randn(b, hw, hw, cin, device=DEV).to(torch.bfloat16)
    dw = F.conv_wgrad_tc(dy, x, split_k=split_k)
    torch.cuda.synchronize()
    _report(f"tc_wgrad hw{hw} {cin}->{cout} b{b} s{split_k}", dw, R.conv_wgrad(dy, x), 2e-3)


@pytest.mark.parametrize("b", [1, 5, 16, 37])
def test_im2col_eight_rows_per_cta_matches_the_one_row_kernel(F, b):
    """The engine's im2col (columns 32..63 pre-cleared, eight image rows per CTA) against the zero-padding one-row
    kernel: bit-identical patch matrix and weight matrix, and against an unfold of the image."""
    torch.manual_seed(21)
    x = torch.randn(b, 32, 32, 3, device=DEV).to(torch.bfloat16)
    w = (torch.randn(64, 3, 3, 3, device=DEV) * 0.2).to(torch.bfloat16)
    ref_cols, ref_w = F.im2col3x3_c3(x, w, zero_pad=True)
    cols, w_pad = F.im2col3x3_c3(x, w, zero_pad=False)
    torch.cuda.synchronize()
    assert torch.equal(cols, ref_cols) and torch.equal(w_pad, ref_w)
    # independent oracle: column k = (r * 3 + s) * 3 + c holds x[n, h + r - 1, w + s - 1, c]
    pad = torch.nn.functional.pad(x.float().permute(0, 3, 1, 2), (1, 1, 1, 1))
    for r in range(3):
        for s_ in range(3):
            want = pad[:, :, r:r + 32, s_:s_ + 32].permute(0, 2, 3, 1)
            k = (r * 3 + s_) * 3
            assert torch.equal(cols[..., k:k + 3].float(), want), (r, s_)
    assert cols[..., 27:].abs().max().item() == 0


@pytest.mark.parametrize("hw,cin,cout", [(4, 256, 512), (2, 512, 512), (8, 128, 256)])
def test_conv_tc_wgrad_by_channel_range(F, R, hw, cin, cout):
    """The engine's sub-bucket hand-off computes a layer's weight gradient as output-channel halves: the halves
    must fill disjoint rows of dW and together equal the one-launch result (fp64 reference)."""
    torch.manual_seed(6)
    b = 16
    dy = torch.randn(b, hw, hw, cout, device=DEV).to(torch.bfloat16)
    x = torch.randn(b, hw, hw, cin, device=DEV).to(torch.bfloat16)
    dw = torch.zeros((cout, 3, 3, cin), dtype=torch.float32, device=DEV)
    half = cout // 2
    F.conv_wgrad_tc_range(dy, x, 0, half, dw)
    torch.cuda.synchronize()
    assert dw[half:].abs().max().item() == 0                      # the second half's rows are untouched so far
    _report(f"tc_wgrad range lo hw{hw} {cin}->{cout}", dw[:half], R.conv_wgrad(dy, x)[:half], 2e-3)
    F.conv_wgrad_tc_range(dy, x, half, half, dw)
    torch.cuda.synchronize()
    _report(f"tc_wgrad range all hw{hw} {cin}->{cout}", dw, R.conv_wgrad(dy, x), 2e-3)
    with pytest.raises(Exception):
        F.conv_wgrad_tc_range(dy, x, 64, half, dw)                # not a multiple of 128


@pytest.mark.parametrize("b", [16, 5, 37])
@pytest.mark.parametrize("cout", [64, 128])
@pytest.mark.parametrize("generic", [False, True])
def test_first_layer_tc(F, R, b, cout, generic):
    """3-channel first layer as im2col + tensor-core GEMM: forward, BN statistics and weight gradient."""
    torch.manual_seed(15)
    x = torch.randn(b, 32, 32, 3, device=DEV).to(torch.bfloat16)
    w = (torch.randn(cout, 3, 3, 3, device=DEV) * 0.2).to(torch.bfloat16)
    bias = torch.randn(cout, device=DEV) * 0.1
    y, stats, cols = F.first_layer_fwd_tc(x, w, bias, want_stats=True, generic_kernel=generic)
    torch.cuda.synchronize()
    assert cols[..., 27:].abs().max().item() == 0
    _report(f"first_tc_fwd b{b} c{cout}", y, R.conv_fwd(x, w, bias), 6e-3)
    _report("first_tc_stats", stats, R.channel_stats(y), 1e-3)
    dy = torch.randn(b, 32, 32, cout, device=DEV).to(torch.bfloat16)
    for split_k in (0, 1, 7):
        dw = F.first_layer_wgrad_tc(dy, cols, split_k=split_k)
        torch.cuda.synchronize()
        _report(f"first_tc_wgrad b{b} c{cout} s{split_k}", dw, R.conv_wgrad(dy, x), 2e-3)


@pytest.mark.parametrize("mode_name", ["oneshot", "twoshot", "root"])
@pytest.mark.parametrize("world", [1, 2, 4])
def test_p2p_reduce_sgd_single_gpu_emulation(F, mode_name, world):
    """All 'ranks' live on one GPU and run one after the other with the barriers disabled: checks
    the reduction order, slicing, SGD epilogue and the push of weights to every replica."""
    from cs744_ddp_b200.ops import native as nv
    mode = {"oneshot": nv.P2P_ONESHOT, "twoshot": nv.P2P_TWOSHOT, "root": nv.P2P_ROOT}[mode_name]
    torch.manual_seed(6)
    n, off = 4 * 1000 + 64, 64
    total = off + n + 64
    grads = [torch.randn(total, device=DEV) for _ in range(world)]
    p0 = torch.randn(total, device=DEV)
    masters = [p0.clone() for _ in range(world)]
    shadows = [torch.zeros(total, device=DEV, dtype=torch.bfloat16) for _ in range(world)]
    moms = [torch.zeros(total, device=DEV) for _ in range(world)]
    for r in range(world):
        F.p2p_sync(mode, [g.data_ptr() for g in grads], [m.data_ptr() for m in masters],
                   [s.data_ptr() for s in shadows], None, moms[r], None, r, world, 0, off, n, 0.1, 0.9, 1e-4)
    torch.cuda.synchronize()
    mean_g = torch.stack(grads).sum(0) / world if mode_name != "root" else torch.stack(grads).mean(0)
    g = mean_g + 1e-4 * p0
    ref = p0 - 0.1 * g
    for r in range(world):
        assert torch.allclose(masters[r][off:off + n], ref[off:off + n], rtol=1e-5, atol=1e-6), (mode_name, r)
        assert torch.equal(masters[r][:off], p0[:off]) and torch.equal(masters[r][off + n:], p0[off + n:])
        assert torch.equal(shadows[r][off:off + n], masters[r][off:off + n].to(torch.bfloat16))
    # every replica must hold bit-identical weights
    for r in range(1, world):
        assert torch.equal(masters[0], masters[r])


@pytest.mark.parametrize("mode_name", ["twoshot", "root"])
@pytest.mark.parametrize("world", [2, 4])
def test_p2p_sharded_master_single_gpu_emulation(F, mode_name, world):
    """Ranges marked as sharded (the conv weights) publish only their bf16 copy; the fp32 master stays with
    the owner of the slice until an all-gather (two-shot) / broadcast (root) brings it everywhere."""
    from cs744_ddp_b200.ops import native as nv
    mode = {"twoshot": nv.P2P_TWOSHOT, "root": nv.P2P_ROOT}[mode_name]
    torch.manual_seed(8)
    n, off = 8 * 600, 64
    total = off + n + 64
    shard = [off + 8 * 100, off + 8 * 500]                 # the middle of the range
    grads = [torch.randn(total, device=DEV) for _ in range(world)]
    p0 = torch.randn(total, device=DEV)
    masters = [p0.clone() for _ in range(world)]
    shadows = [torch.zeros(total, device=DEV, dtype=torch.bfloat16) for _ in range(world)]
    moms = [torch.zeros(total, device=DEV) for _ in range(world)]
    tabs = lambda ts: [t.data_ptr() for t in ts]
    for r in range(world):
        F.p2p_sync(mode, tabs(grads), tabs(masters), tabs(shadows), None, moms[r], None, r, world, 0, off, n, 0.1, 0.9,
                   1e-4, shard=shard)
    torch.cuda.synchronize()
    ref = p0 - 0.1 * (torch.stack(grads).mean(0) + 1e-4 * p0)
    lo, hi = shard
    for r in range(world):
        # the bf16 copy is complete everywhere, and so is the fp32 master outside the sharded range
        assert torch.equal(shadows[r][off:off + n], ref[off:off + n].to(torch.bfloat16)) or \
            torch.allclose(shadows[r][off:off + n].float(), ref[off:off + n], rtol=1e-2, atol=1e-2)
        assert torch.allclose(masters[r][off:lo], ref[off:lo], rtol=1e-5, atol=1e-6)
        assert torch.allclose(masters[r][hi:off + n], ref[hi:off + n], rtol=1e-5, atol=1e-6)
    # inside it, non-owners still hold the old values somewhere
    stale = [int((masters[r][lo:hi] == p0[lo:hi]).sum()) for r in range(world)]
    assert any(s > 0 for s in stale), stale
    # bring the shards together the way the engine does before a state_dict is read
    for r in range(world):
        F.p2p_sync(nv.P2P_ALLGATHER if mode_name == "twoshot" else nv.P2P_BCAST, tabs(masters), None, None, None, None,
                   None, r, world, 0, off, n)
    torch.cuda.synchronize()
    for r in range(world):
        assert torch.allclose(masters[r][off:off + n], ref[off:off + n], rtol=1e-5, atol=1e-6), (mode_name, r)
        assert torch.equal(masters[r], masters[0])


@pytest.mark.parametrize("b", [16, 5, 37])
def test_first_layer_backward_fused(F, R, b):
    """BN/ReLU/pool backward apply + 3 -> 64 weight gradient in one kernel == the two separate kernels, and both
    close to a plain fp32 torch autograd of conv -> BN -> ReLU -> max-pool (ref Part 1/model.py:16,18-25)."""
    torch.manual_seed(31)
    x = torch.randn(b, 32, 32, 3, device=DEV).to(torch.bfloat16)
    w = (torch.randn(64, 3, 3, 3, device=DEV) * 0.2).to(torch.bfloat16)
    bias = torch.zeros(64, device=DEV)
    gamma = torch.rand(64, device=DEV) + 0.5
    beta = torch.randn(64, device=DEV) * 0.2
    y, stats, cols = F.first_layer_fwd_tc(x, w, bias, True)
    rm, rv = torch.zeros(64, device=DEV), torch.ones(64, device=DEV)
    out, mean, invstd = F.bn_relu_pool_fwd(y, stats, gamma, beta, rm, rv, None, True, True)
    dout = torch.randn_like(out)
    # the two-kernel path (its own reduction pass produces the sums the fused kernel is given)
    dbias_ref = torch.zeros(64, device=DEV)
    dy, dgamma_ref, dbeta_ref = F.bn_relu_pool_bwd(dout, y, mean, invstd, gamma, beta, True, dbias_ref)
    dw_ref = F.first_layer_wgrad_tc(dy, cols).reshape(64, 27)
    sums = torch.cat([dbeta_ref, dgamma_ref]).contiguous()
    for ctas in (0, 1, 7):
        dw, dgamma, dbeta, dbias = F.first_layer_bwd_fused(dout, y, cols, mean, invstd, gamma, beta, sums, ctas)
        torch.cuda.synchronize()
        _report(f"first_bwd_fused b{b} ctas{ctas} dw", dw, dw_ref, 2e-3)
        assert torch.equal(dgamma, dgamma_ref) and torch.equal(dbeta, dbeta_ref)
        scale = max(1.0, float(dy.float().abs().sum(dim=(0, 1, 2)).max()))
        assert float((dbias - dbias_ref).abs().max()) < 1e-2 * scale
    # fp32 torch oracle of the whole block
    xt = x.float().permute(0, 3, 1, 2).requires_grad_(False)
    wt = w.float().permute(0, 3, 1, 2).clone().requires_grad_(True)       # KRSC -> OIHW
    yt = torch.nn.functional.conv2d(xt, wt, None, 1, 1)
    ot = torch.nn.functional.max_pool2d(torch.relu(torch.nn.functional.batch_norm(yt, None, None, gamma, beta, True, 0.1, 1e-5)), 2)
    ot.backward(dout.float().permute(0, 3, 1, 2))
    # (y, dY and the patch matrix are bf16 in the kernel: ~2^-8 relative rounding on each factor of a long sum)
    _report(f"first_bwd_fused b{b} dw vs torch", dw, wt.grad.permute(0, 2, 3, 1).reshape(64, 27), 8e-2)


@pytest.mark.parametrize("b", [16, 10])
@pytest.mark.parametrize("hw,cin,cout", TC_SHAPES)
def test_conv_tf32(F, R, b, hw, cin, cout):
    """fp32 convolutions on tcgen05 kind::tf32 (10-bit mantissa inputs, fp32 accumulation) against the fp64-based
    reference: forward (+ BatchNorm statistics), data gradient, weight gradient."""
    torch.manual_seed(5)
    x = torch.randn(b, hw, hw, cin, device=DEV)
    dy = torch.randn(b, hw, hw, cout, device=DEV)
    w = torch.randn(cout, 3, 3, cin, device=DEV) * 0.05
    bias = torch.randn(cout, device=DEV) * 0.1
    y, stats = F.conv_fwd_tf32(x, w, bias, want_stats=True)
    dx = F.conv_dgrad_tf32(dy, w)
    torch.cuda.synchronize()
    _report(f"tf32_fwd hw{hw} {cin}->{cout} b{b}", y, R.conv_fwd(x, w, bias), 2e-3)
    _report("tf32_fwd_stats", stats, R.channel_stats(y), 1e-4)
    _report(f"tf32_dgrad hw{hw} {cin}->{cout} b{b}", dx, R.conv_dgrad(dy, w), 2e-3)
    if cout % 128 == 0:
        for sk in (0, 1, 3):
            dw = F.conv_wgrad_tf32(dy, x, sk)
            torch.cuda.synchronize()
            _report(f"tf32_wgrad hw{hw} {cin}->{cout} b{b} split{sk}", dw, R.conv_wgrad(dy, x), 2e-3)
