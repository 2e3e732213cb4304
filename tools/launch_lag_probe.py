#!/usr/bin/env python
"""Does NVLink traffic of one kernel delay the START of unrelated kernels? (torchrun, >= 2 GPUs)

The step tracer shows backward kernels starting 15-20 us after their dependencies resolved whenever a bucket
kernel (reduce + SGD over NVLink) is running, and never on one GPU.  This probe isolates the effect: a CUDA graph
of CHAIN tiny dependent kernels (one stream, or alternating between two streams) is timed alone and while a
background kernel on another stream moves data with one kind of access:
  local ld | peer ld.relaxed.sys | multimem.ld_reduce | multimem.st only | peer ld + peer st | mm.ld_reduce + mm.st
The chain does no memory traffic of its own, so any slowdown is launch / dependency-resolution latency.
"""
import argparse
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from cs744_ddp_b200.ops.native import check, lib  # noqa: E402
from cs744_ddp_b200.parallel.p2p import SymmetricArena  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--chain", type=int, default=200)
    ap.add_argument("--mb", type=int, default=256, help="arena size: the background kernel must outlast the chain")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    n = (a.mb << 20) // 4
    arena = SymmetricArena(n, torch.device("cuda", local), nvls=True)
    arena.grad.normal_()
    sink = torch.zeros(160 * 512, device="cuda")
    local_rw = torch.zeros(16, device="cuda")
    L = lib()
    nxt = (rank + 1) % world
    peer, peer_sh = arena.peer_base[nxt] + arena.off_grad, arena.peer_base[nxt] + arena.off_shadow
    own = arena.peer_base[rank] + arena.off_grad
    mc = arena.mc_base + arena.off_grad if arena.mc_base else 0
    mc_sh = arena.mc_base + arena.off_shadow if arena.mc_base else 0
    per = n // world

    # the chains: tiny kernels, each depending on the previous one
    x = torch.zeros(256, device="cuda")
    xw = torch.zeros(148 * 4 * 1024, device="cuda")        # a chain kernel with CTAs on every SM
    s_main, s_side, s_bg = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()

    def capture(two_streams, x=x):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(s_main):
            x.add_(1.0)
            torch.cuda.synchronize()
            with torch.cuda.graph(g, stream=s_main):
                for i in range(a.chain):
                    if two_streams and i % 2:
                        s_side.wait_stream(s_main)
                        with torch.cuda.stream(s_side):
                            x.add_(1.0)
                        s_main.wait_stream(s_side)
                    else:
                        x.add_(1.0)
        return g

    graphs = {"one stream": capture(False), "two streams": capture(True)}

    def background(kind, blocks=148, threads=128):
        if kind is None:
            return
        src, cnt, dst = {0: (own, n, 0), 1: (peer, n, 0), 2: (mc + rank * per * 4, per, 0),
                         6: (mc + rank * per * 4, per, mc_sh + rank * per * 2), 4: (peer, n, peer_sh),
                         3: (mc + rank * per * 4, per, mc_sh + rank * per * 2)}[kind]
        for _ in range(4 if cnt == n else 4 * world):
            check(L.ddpb_probe_peer_read(ctypes.c_void_p(src), ctypes.c_void_p(sink.data_ptr()), ctypes.c_longlong(cnt), kind,
                                         blocks, threads, 4, ctypes.c_void_p(dst), ctypes.c_void_p(local_rw.data_ptr()),
                                         ctypes.c_void_p(s_bg.cuda_stream)), "probe")

    kinds = [(None, "nothing"), (0, "local ld"), (1, "peer ld.relaxed.sys"), (2, "multimem.ld_reduce"), (6, "multimem.st only"),
             (4, "peer ld + peer st"), (3, "mm.ld_reduce + mm.st")]
    lines = [f"# launch-lag probe, {world} GPUs, chain of {a.chain} dependent tiny kernels in a CUDA graph; us per kernel of the chain"]
    for kind, name in kinds:
        if kind in (2, 3, 6) and not mc:
            continue
        row = f"{name:24s}"
        for gname, g in graphs.items():
            best = []
            for rep in range(5):
                torch.cuda.synchronize()
                dist.barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                b0.record(s_bg)
                background(kind)
                b1.record(s_bg)
                with torch.cuda.stream(s_main):
                    e0.record(s_main)
                    g.replay()
                    e1.record(s_main)
                torch.cuda.synchronize()
                best.append((e0.elapsed_time(e1) * 1e3 / a.chain, b0.elapsed_time(b1) * 1e3))
            best.sort()
            med = best[len(best) // 2]
            t = torch.tensor([med[0]], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            row += f"  {gname}: {float(t):6.2f} us/kernel (background ran {med[1]:7.0f} us)"
        lines.append(row)
        if rank == 0:
            print(row, flush=True)
    # is the cost per SM or per GPU?  background confined to a few SMs (fat CTAs) vs a chain kernel that is tiny
    # (lands on an SM of its own) or wide (CTAs on every SM)
    graphs2 = {"tiny chain": graphs["one stream"], "wide chain": capture(False, xw)}
    lines.append("# background confined to `blocks` SMs")
    for kind, name in ((1, "peer ld.relaxed.sys"), (6, "multimem.st only"), (4, "peer ld + peer st"), (3, "mm.ld_reduce + mm.st")):
        if kind in (3, 6) and not mc:
            continue
        for blocks, threads in ((148, 128), (32, 512), (8, 512), (2, 512)):
            row = f"{name:22s} {blocks:3d} x {threads:3d}"
            for gname, g in graphs2.items():
                best = []
                for rep in range(5):
                    torch.cuda.synchronize()
                    dist.barrier()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    b0.record(s_bg)
                    background(kind, blocks, threads)
                    b1.record(s_bg)
                    with torch.cuda.stream(s_main):
                        e0.record(s_main)
                        g.replay()
                        e1.record(s_main)
                    torch.cuda.synchronize()
                    best.append((e0.elapsed_time(e1) * 1e3 / a.chain, b0.elapsed_time(b1) * 1e3))
                best.sort()
                med = best[len(best) // 2]
                t = torch.tensor([med[0]], device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                row += f"  {gname}: {float(t):6.2f} us/kernel (bg {med[1]:7.0f} us)"
            lines.append(row)
            if rank == 0:
                print(row, flush=True)
    if rank == 0 and a.out:
        with open(a.out, "w") as f:
            f.write("\n".join(lines) + "\n")
    dist.barrier()
    arena.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
