"""Per-CTA L2 -> shared-memory ingest rate of the conv kernels' load patterns (csrc/kernels/tma_probe.cu).

Run on a B200: ``python tools/tma_probe.py``.  Prints bytes per clock per SM for each load flavour,
pipeline depth and CTA count, so the limiter of the conv main loops can be identified.
"""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cs744_ddp_b200.ops import native as nv  # noqa: E402

MODES = {0: "conv A(4-D)+B", 1: "conv A(4-D) only", 2: "2-D matrix box", 3: "1-D bulk 16 KB", 4: "A+B + tcgen05.mma"}


def main():
    dev = torch.device("cuda")
    lib = nv.lib()
    act = torch.randn(256, 4, 4, 512, device=dev).to(torch.bfloat16)
    wgt = torch.randn(512, 4608, device=dev).to(torch.bfloat16)
    flat = torch.empty(64 << 20, dtype=torch.uint8, device=dev)
    clk_mhz = torch.cuda.clock_rate() / 1e3 if hasattr(torch.cuda, "clock_rate") else 1965.0
    iters = 2000
    print(f"SM clock assumed {clk_mhz:.0f} MHz; {iters} stages per CTA")
    for mode, name in MODES.items():
        for stages, pad in ((4, 0), (4, 40), (8, 0)):
            for ctas in (64, 128, 148, 256, 296):
                per_sm = 2 if (stages == 4 and pad == 0) else 1
                def run():
                    nv.check(lib.ddpb_probe_tma(mode, stages, iters, ctas, pad, nv.ptr(act), nv.ptr(wgt), nv.ptr(flat),
                                                ctypes.c_longlong(flat.numel()), nv.stream_handle()), "probe_tma")
                run()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                run()
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1)
                nbytes = ctas * iters * (16384 + (8192 if mode in (0, 4) else 0))
                sms = min(148, ctas if per_sm == 1 else (ctas + 1) // 2 if ctas > 148 else ctas)
                tbs = nbytes / ms / 1e9
                print(f"mode {mode} {name:20s} stages {stages} max CTA/SM {per_sm} ctas {ctas:3d}: {ms * 1e3:8.1f} us  "
                      f"{tbs:6.2f} TB/s  {nbytes / (ms * 1e-3) / (clk_mhz * 1e6) / ctas:6.1f} B/clk/CTA", flush=True)


if __name__ == "__main__":
    main()
