#!/usr/bin/env python
"""Gradient all-reduce sweep, 4 KB - 512 MB (BASELINE.json config 5).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/allreduce_sweep.py

For each message size S (fp32 gradient bytes) it times, with CUDA events, max over ranks:
  * p2p fused   - this repo's peer-to-peer kernel: reduce + 1/W + SGD momentum/weight decay + bf16 copy
                  (one-shot below 256 KB, two-shot above; sync_p2p.cu)
  * nccl        - torch.distributed.all_reduce (the reference's backend, ref Part 2b/main.py:118)
  * nccl + sgd  - all_reduce followed by the standalone fused SGD kernel (what the P2P kernel replaces)
Reported: time, algorithm bandwidth S/t, and for the P2P path the fraction of the NVLink roofline
(two-shot moves 2(W-1)/W*S per direction per GPU; measured peer copy 770 GB/s per direction).
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cs744_ddp_b200.ops import functional as F  # noqa: E402
from cs744_ddp_b200.ops import native as nv  # noqa: E402
from cs744_ddp_b200.parallel.p2p import SymmetricArena  # noqa: E402

LINK_GBS = 770.0


def timed(fn, iters, world):
    for _ in range(3):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t) * 1e3       # us


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    max_bytes = int(os.environ.get("SWEEP_MAX_MB", "512")) << 20
    arena = SymmetricArena(max_bytes // 4, torch.device("cuda", local), nvls=os.environ.get("SWEEP_NVLS", "1") == "1")
    has_nvls = bool(arena.mc_base)
    arena.grad.normal_()
    flat = torch.randn(max_bytes // 4, device="cuda")
    p_, m_ = torch.zeros_like(flat), torch.zeros_like(flat)
    sh = torch.empty(flat.numel(), device="cuda", dtype=torch.bfloat16)
    stream = torch.cuda.Stream()
    rows = []
    # 4 KB .. 512 MB in powers of four, plus the VGG-11 gradient itself (36.9 MB) and the 512 MB end point
    sizes = sorted({4 << (10 + 2 * k) for k in range(9)} | {36_924_456 // 32 * 32, 512 << 20})
    sizes = [v for v in sizes if v <= max_bytes]
    # the sweep owns the GPU: as many CTAs as the flag table allows (in a training step the same kernels run
    # next to backward with a small grid)
    big_grid = int(os.environ.get("SWEEP_BLOCKS", str(min(nv.P2P_MAX_BLOCKS, 148))))
    with torch.cuda.stream(stream):
        for size in sizes:
            n = size // 4
            iters = 50 if size <= (8 << 20) else 10
            two = size >= 256 * 1024
            # conv weights (99.85 % of a VGG gradient) publish their bf16 copy only: the swept range is "sharded"
            shard = [0, n] if two else None
            blocks = (big_grid if "SWEEP_BLOCKS" in os.environ else -1) if size >= (4 << 20) else 0     # -1: fat CTAs, grid by size
            t_p2p = timed(lambda: arena.sync(nv.P2P_TWOSHOT if two else nv.P2P_ONESHOT, 0, n, 0.1, 0.9, 1e-4,
                                             blocks=blocks, shard=shard), iters, world)
            t_nvls = timed(lambda: arena.sync_nvls(0, n, 0.1, 0.9, 1e-4, blocks=blocks, shard=shard, one_shot=not two),
                           iters, world) if has_nvls else float("nan")
            t_nccl = timed(lambda: dist.all_reduce(flat[:n]), iters, world)

            def nccl_sgd():
                dist.all_reduce(flat[:n])
                F.sgd_step(p_[:n], m_[:n], flat[:n], sh[:n], 0.1, 0.9, 1e-4, 1.0 / world)
            t_ns = timed(nccl_sgd, iters, world)
            # bytes per direction per GPU: unicast two-shot pulls (W-1)/W*S of fp32 and pushes (W-1)/W*S/2 of bf16;
            # NVLS: (W-1)/W*S of gradients leave each GPU for the switch (the slower direction)
            wire = ((world - 1) / world * size) if two else ((world - 1) * size)
            roof_us = wire / (LINK_GBS * 1e3)
            nvls_wire = (world - 1) / world * size + size / (2.0 * world)      # + the bf16 copy of the own slice
            rows.append({"bytes": size, "mode": "two-shot" if two else "one-shot", "p2p_fused_us": t_p2p,
                         "nvls_fused_us": t_nvls, "nvls_frac_of_nvlink_roofline": (nvls_wire / (LINK_GBS * 1e3)) / t_nvls,
                         "nvls_beats_nccl": bool(t_nvls <= t_nccl),
                         "nccl_us": t_nccl, "nccl_plus_sgd_us": t_ns, "p2p_algbw_gbs": size / t_p2p / 1e3,
                         "nccl_algbw_gbs": size / t_nccl / 1e3, "p2p_frac_of_nvlink_roofline": roof_us / t_p2p})
            if rank == 0:
                r = rows[-1]
                print(f"{size:>11d} B {r['mode']:8s} p2p+sgd {t_p2p:9.1f} us | nvls+sgd {t_nvls:9.1f} us | nccl {t_nccl:9.1f} us | nccl+sgd {t_ns:9.1f} us"
                      f" | p2p algbw {r['p2p_algbw_gbs']:7.1f} GB/s | roofline frac {r['p2p_frac_of_nvlink_roofline']:.2f}",
                      flush=True)
    if rank == 0:
        out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
        os.makedirs(out, exist_ok=True)
        with open(os.path.join(out, f"allreduce_sweep_w{world}.json"), "w") as fh:
            json.dump({"world": world, "link_gbs": LINK_GBS, "rows": rows}, fh, indent=1)
    dist.barrier()
    arena.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
