#!/usr/bin/env python
"""Where does a convolution kernel's time go?  Per-CTA %globaltimer stamps (ddpb_conv3x3_trace) for every
VGG-11 layer, forward and dgrad, at the headline batch.  Printed per phase as median / max over CTAs, in
microseconds since the first CTA started.

    python tools/conv_trace.py [--tiles 0,128,4228] [--out profiles/conv_trace.txt]
"""
import argparse
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from cs744_ddp_b200.models.vgg import conv_blocks  # noqa: E402
from cs744_ddp_b200.ops.native import check, lib, ptr, stream_handle  # noqa: E402

PHASES = ["start", "prologue", "1st stage", "last mma", "accum", "split-K", "stored", "done"]


def run(dgrad, x, w, bias, B, hw, cin, cout, tile, reps=5):
    out = torch.empty(B, hw, hw, cin if dgrad else cout, device="cuda", dtype=torch.bfloat16)
    stats = None if dgrad else torch.zeros(2 * cout, device="cuda")
    trace = torch.zeros(4096 * 8, device="cuda", dtype=torch.int64)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    best = None
    for r in range(reps):
        trace.zero_()
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        check(lib().ddpb_conv3x3_trace(int(dgrad), ptr(x), ptr(w), ptr(bias), ptr(out), ptr(stats), B, hw, hw, cin, cout,
                                       tile, ctypes.c_void_p(trace.data_ptr()), stream_handle()), "conv_trace")
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        t = trace.view(-1, 8)
        t = t[t[:, 0] > 0].double()
        if best is None or ms < best[0]:
            best = (ms, t.clone())
    ms, t = best
    t0 = t[:, 0].min()
    rel = (t - t0) / 1e3
    rel[t == 0] = float("nan")
    cols = []
    for k in range(8):
        c = rel[:, k]
        c = c[~torch.isnan(c)]
        cols.append((float(c.median()), float(c.max())) if c.numel() else (float("nan"), float("nan")))
    return ms * 1e3, t.shape[0], cols


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tiles", default="0")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    B = a.batch
    lines = ["# per-CTA phase stamps, us since the first CTA started: median/max over CTAs; event = CUDA-event time of the launch",
             f"{'kernel':30s} {'ctas':>5s} {'event':>7s} " + " ".join(f"{p:>13s}" for p in PHASES[1:])]
    for blk in conv_blocks("VGG11"):
        if blk.cin == 3:
            continue
        hw, cin, cout = blk.hw, blk.cin, blk.cout
        x = torch.randn(B, hw, hw, cin, device="cuda").to(torch.bfloat16)
        dy = torch.randn(B, hw, hw, cout, device="cuda").to(torch.bfloat16)
        w = (torch.randn(cout, 3, 3, cin, device="cuda") * 0.05).to(torch.bfloat16)
        bias = torch.zeros(cout, device="cuda")
        for tile in [int(v) for v in a.tiles.split(",")]:
            for dgrad in (0, 1):
                try:
                    us, n, cols = run(dgrad, dy if dgrad else x, w, bias, B, hw, cin, cout, tile)
                except Exception as e:      # tile code does not fit this layer
                    continue
                name = f"L{blk.index} {'dgrad' if dgrad else 'fwd  '} {hw}x{hw} {cin}->{cout} n{tile}"
                lines.append(f"{name:30s} {n:5d} {us:7.1f} " + " ".join(f"{m:6.1f}/{x_:6.1f}" for m, x_ in cols[1:]))
                print(lines[-1], flush=True)
    text = "\n".join(lines)
    if a.out:
        with open(a.out, "w") as f:
            f.write(text + "\n")


if __name__ == "__main__":
    main()
