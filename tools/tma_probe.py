"""Per-CTA L2 -> shared-memory ingest rate of the conv kernels' load patterns (csrc/kernels/tma_probe.cu).

Run on a B200: ``python tools/tma_probe.py``.  Prints bytes per clock per CTA for each load flavour,
pipeline depth, CTAs per SM and CTA count, so the limiter of the conv main loops can be identified.
"""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cs744_ddp_b200.ops import native as nv  # noqa: E402

MODES = {0: "conv A(4-D)+B", 1: "conv A(4-D) only", 2: "2-D box", 3: "1-D bulk 16 KB", 4: "A+B + tcgen05.mma", 5: "mma, 2 producer warps", 6: "mma, A halves by 2 warps"}
CLK_MHZ = 1965.0
ITERS = 2000


def main():
    dev = torch.device("cuda")
    lib = nv.lib()
    act = torch.randn(256, 4, 4, 512, device=dev).to(torch.bfloat16)
    wgt = torch.randn(512, 4608, device=dev).to(torch.bfloat16)
    flat = torch.empty(64 << 20, dtype=torch.uint8, device=dev)
    # (mode, stages, smem pad KB, CTAs, cols); pad 0 lets several CTAs share an SM, pad 40/130 forces one
    cases = [(m, st, pad, c, 512) for m in (4, 5, 6) for st, pad in ((4, 0), (8, 0)) for c in (64, 148, 296)]
    cases += [(m, st, pad, c, 512) for m in (0, 1) for st, pad in ((4, 0),) for c in (148, 296)]
    cases += [(4, 2, 0, c, 512) for c in (148, 296, 444, 592)]          # 2 stages: up to 4 CTAs per SM
    cases += [(4, 2, 130, 148, 512)]
    cases += [(2, 4, 40, 148, cols) for cols in (64, 128, 256, 512, 1024, 2048, 4608)]   # row stride = 2 * cols bytes
    cases += [(3, 4, 40, 148, 512)]
    for mode, stages, pad, ctas, cols in cases:
        name = MODES[mode] + (f" cols={cols}" if mode == 2 else "")

        def run():
            nv.check(lib.ddpb_probe_tma(mode, stages, ITERS, ctas, pad, cols, nv.ptr(act), nv.ptr(wgt), nv.ptr(flat),
                                        ctypes.c_longlong(flat.numel()), nv.stream_handle()), "probe_tma")
        run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        nbytes = ctas * ITERS * (16384 + (8192 if mode in (0, 4, 5, 6) else 0))
        print(f"mode {mode} {name:24s} stages {stages} pad {pad:3d} KB ctas {ctas:3d}: {ms * 1e3:8.1f} us  "
              f"{nbytes / ms / 1e9:6.2f} TB/s  {nbytes / (ms * 1e-3) / (CLK_MHZ * 1e6) / ctas:6.1f} B/clk/CTA", flush=True)


if __name__ == "__main__":
    main()
