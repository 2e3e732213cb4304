#!/usr/bin/env python
"""Per-kernel timings at the headline shapes (batch 256 per GPU), CUDA events, L2 flushed between
timed launches.  Prints one line per kernel with achieved TFLOP/s or GB/s and the fraction of the
measured peaks in MEASURED_PEAKS.json (fallback 6.65 TB/s, 1.59 PFLOP/s)."""
import json
import os
import sys

import torch


def tile_ok(code, channels, hw):
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

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cs744_ddp_b200.models.vgg import conv_blocks  # noqa: E402
from cs744_ddp_b200.ops import functional as F  # noqa: E402

B = int(os.environ.get("KB_BATCH", "256"))
DEV = "cuda"
peaks = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
if os.path.exists(p):
    peaks.update(json.load(open(p)))
flush_buf = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=DEV)


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush_buf.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def line(name, ms, flops=None, nbytes=None):
    s = f"{name:44s} {ms*1e3:9.1f} us"
    if flops:
        tf = flops / ms / 1e9
        s += f"  {tf:8.1f} TFLOP/s ({tf / peaks['bf16_tflops'] * 100:5.1f}% of measured bf16 peak)"
    if nbytes:
        gb = nbytes / ms / 1e6
        s += f"  {gb:8.1f} GB/s ({gb / peaks['hbm_gbs'] * 100:5.1f}% of measured HBM copy)"
    print(s, flush=True)
    return ms


def main():
    torch.manual_seed(0)
    total = {"fwd": 0.0, "dgrad": 0.0, "wgrad": 0.0, "bn_fwd": 0.0, "bn_bwd": 0.0, "cudnn_fwd": 0.0}
    for blk in conv_blocks("VGG11"):
        hw, cin, cout = blk.hw, blk.cin, blk.cout
        flops = 2.0 * B * hw * hw * cout * 9 * cin
        x = torch.randn(B, hw, hw, cin, device=DEV).to(torch.bfloat16)
        dy = torch.randn(B, hw, hw, cout, device=DEV).to(torch.bfloat16)
        w32 = torch.randn(cout, 3, 3, cin, device=DEV) * 0.05
        w16 = w32.to(torch.bfloat16)
        bias = torch.zeros(cout, device=DEV)
        tag = f"L{blk.index} {hw}x{hw} {cin}->{cout}"
        if cin == 3:
            line(f"first_layer_fwd (simt) {tag}", timeit(lambda: F.conv_fwd_simt(x, w32, bias, True)), flops)
            line(f"first_layer_wgrad (simt) {tag}", timeit(lambda: F.conv_wgrad_simt(dy, x)), flops)
            line(f"first_layer_fwd im2col+conv1x1 {tag}", timeit(lambda: F.first_layer_fwd_tc(x, w16, bias, True, True)), flops)
            total["fwd"] += line(f"first_layer_fwd im2col+gemm_rows {tag}", timeit(lambda: F.first_layer_fwd_tc(x, w16, bias, True)), flops)
            cols = F.first_layer_fwd_tc(x, w16, bias, True)[2]
            total["wgrad"] += line(f"first_layer_wgrad tc {tag}", timeit(lambda: F.first_layer_wgrad_tc(dy, cols)), flops)
        else:
            for tn in TILE_CODES:
                if tile_ok(tn, cout, hw):
                    t = line(f"conv_fwd_tc n{tn}   {tag}", timeit(lambda: F.conv_fwd_tc(x, w16, bias, True, tn)), flops)
            total["fwd"] += timeit(lambda: F.conv_fwd_tc(x, w16, bias, True, 0))
            for tn in TILE_CODES:
                if tile_ok(tn, cin, hw):
                    line(f"conv_dgrad_tc n{tn} {tag}", timeit(lambda: F.conv_dgrad_tc(dy, w16, tn)), flops)
            total["dgrad"] += timeit(lambda: F.conv_dgrad_tc(dy, w16, 0))
            dw = torch.zeros(cout, 3, 3, cin, device=DEV)
            total["wgrad"] += line(f"conv_wgrad_tc     {tag}", timeit(lambda: F.conv_wgrad_tc(dy, x, 0, dw)), flops)
        # cuDNN for context (bf16 channels_last)
        xc = x.permute(0, 3, 1, 2)
        wc = w16.permute(0, 3, 1, 2)
        total["cudnn_fwd"] += line(f"cudnn conv fwd    {tag}", timeit(lambda: torch.nn.functional.conv2d(xc, wc, None, 1, 1)), flops)
        # BN
        y = torch.randn(B, hw, hw, cout, device=DEV).to(torch.bfloat16)
        gamma, beta = torch.ones(cout, device=DEV), torch.zeros(cout, device=DEV)
        rm, rv = torch.zeros(cout, device=DEV), torch.ones(cout, device=DEV)
        stats = F.channel_stats(y)
        nb = y.numel() * 2
        ob = nb // 4 if blk.pool else nb
        out, mean, invstd = F.bn_relu_pool_fwd(y, stats, gamma, beta, rm, rv, None, blk.pool, True)
        total["bn_fwd"] += line(f"bn_relu_pool_fwd  {tag}", timeit(lambda: F.bn_relu_pool_fwd(y, stats, gamma, beta, rm, rv, None, blk.pool, True, out)), nbytes=nb + ob)
        dout = torch.randn_like(out)
        total["bn_bwd"] += line(f"bn_relu_pool_bwd  {tag}", timeit(lambda: F.bn_relu_pool_bwd(dout, y, mean, invstd, gamma, beta, blk.pool)), nbytes=3 * nb + 2 * ob)
    n = 9_232_384
    p_, m_, g_ = (torch.randn(n, device=DEV) for _ in range(3))
    sh = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    line("sgd_step (whole arena)", timeit(lambda: F.sgd_step(p_, m_, g_, sh, 0.1, 0.9, 1e-4)), nbytes=n * (12 + 8 + 2))
    print("TOTALS (us): " + ", ".join(f"{k}={v*1e3:.0f}" for k, v in total.items()))
    print(f"sum of our kernels: {sum(v for k, v in total.items() if k != 'cudnn_fwd')*1e3:.0f} us per step (excl. fc/sgd)")


if __name__ == "__main__":
    main()
