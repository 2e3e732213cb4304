#!/usr/bin/env python
"""Read bandwidth over NVLink as a function of the CTA shape (torchrun, >= 2 GPUs).

Rank 0 reads 64 MB of rank 1's arena with plain loads / ld.relaxed.sys, and all ranks read the switch-reduced
multicast view (multimem.ld_reduce) of their slice, for several grids, block sizes and unroll depths.
"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from cs744_ddp_b200.ops.native import check, lib, stream_handle  # noqa: E402
from cs744_ddp_b200.parallel.p2p import SymmetricArena  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    n = (64 << 20) // 4
    arena = SymmetricArena(n, torch.device("cuda", local), nvls=True)
    arena.grad.normal_()
    sink = torch.zeros(160 * 512, device="cuda")
    local_rw = torch.zeros(2 * n, device="cuda")
    torch.cuda.synchronize()
    dist.barrier()
    peer = arena.peer_base[(rank + 1) % world] + arena.off_grad
    mc = arena.mc_base + arena.off_grad if arena.mc_base else 0
    L = lib()

    peer_sh = arena.peer_base[(rank + 1) % world] + arena.off_shadow
    mc_sh = arena.mc_base + arena.off_shadow if arena.mc_base else 0

    def run(src, count, kind, blocks, threads, unroll, collective, dst=0):
        def once():
            check(L.ddpb_probe_peer_read(ctypes.c_void_p(src), ctypes.c_void_p(sink.data_ptr()), ctypes.c_longlong(count), kind,
                                         blocks, threads, unroll, ctypes.c_void_p(dst), ctypes.c_void_p(local_rw.data_ptr()),
                                         stream_handle()), "probe")
        for _ in range(2):
            once()
        torch.cuda.synchronize()
        if collective:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            once()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / 5 * 1e3

    shapes = [(32, 256, 2), (148, 128, 1), (148, 128, 2), (148, 128, 4), (148, 256, 2), (148, 512, 2), (148, 512, 4)]
    kinds = ((1, "ld.relaxed.sys"), (4, "ld.sys + st.sys bf16"), (2, "multimem.ld_reduce"), (3, "mm.ld_reduce + mm.st"),
             (5, "kind 3 + local r/w"), (6, "mm.st only"))
    for kind, name in kinds:
        for blocks, threads, unroll in shapes:
            if kind in (2, 3, 5, 6):
                if not mc:
                    continue
                per = n // world
                us = run(mc + rank * per * 4, per, kind, blocks, threads, unroll, True, mc_sh + rank * per * 2)
                t = torch.tensor([us], device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                if rank == 0:
                    print(f"{name:20s} {blocks:4d} x {threads:4d} U{unroll:<2d}: {float(t):8.1f} us for S/W = {per * 4 >> 20} MB per rank"
                          f"  -> S / t = {n * 4 / float(t) / 1e3:7.1f} GB/s", flush=True)
            else:
                if rank == 0:
                    us = run(peer, n, kind, blocks, threads, unroll, False, peer_sh)
                    print(f"{name:20s} {blocks:4d} x {threads:4d} U{unroll:<2d}: {us:8.1f} us for 64 MB  -> {n * 4 / us / 1e3:7.1f} GB/s", flush=True)
                dist.barrier()
    dist.barrier()
    arena.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
