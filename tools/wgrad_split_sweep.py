#!/usr/bin/env python
"""Weight-gradient kernel time per VGG-11 layer as a function of the split-K factor (CUDA events, L2 flushed)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from cs744_ddp_b200.models.vgg import conv_blocks  # noqa: E402
from cs744_ddp_b200.ops import functional as F  # noqa: E402

B = int(os.environ.get("KB_BATCH", "256"))
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timeit(fn, iters=7):
    for _ in range(2):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2] * 1e3


for blk in conv_blocks("VGG11"):
    if blk.cin == 3:
        continue
    hw, cin, cout = blk.hw, blk.cin, blk.cout
    x = torch.randn(B, hw, hw, cin, device="cuda").to(torch.bfloat16)
    dy = torch.randn(B, hw, hw, cout, device="cuda").to(torch.bfloat16)
    dw = torch.zeros(cout, 3, 3, cin, device="cuda")
    chunks = max(1, B * hw * hw // 64)
    row = []
    for sp in (0, 1, 2, 3, 4, 6, 8, 12, 16, 24, 32, 48, 64):
        if sp > chunks:
            continue
        row.append(f"{sp}:{timeit(lambda: F.conv_wgrad_tc(dy, x, sp, dw)):.1f}")
    print(f"L{blk.index} {hw}x{hw} {cin}->{cout} base_ctas {(cout // 128) * (cin // 64) * 3:3d} chunks {chunks:5d} | " + "  ".join(row), flush=True)
