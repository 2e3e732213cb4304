#!/usr/bin/env python
"""Per-kernel timeline of one training step.

    python tools/step_timeline.py [--batch 256] [--mode stamps|graph|eager] [--out profiles/step_timeline.txt]
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/step_timeline.py --part 3

`stamps` (default) replays the PRODUCTION CUDA graph; every kernel records %globaltimer at its first CTA's start
and its last warp's end, so the rows show the step as it really runs: programmatic-launch overlap, the
weight-gradient and communication streams next to the critical path, and the gaps between kernels.
"""
import argparse
import collections
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256, help="per-GPU batch")
    ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--mode", default="stamps", choices=["stamps", "graph", "eager"])
    ap.add_argument("--part", default="3", choices=["1", "2a", "2b", "3"])
    ap.add_argument("--bucket-plan", default="per_layer")
    ap.add_argument("--out", default=None)
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    from cs744_ddp_b200.config import TrainConfig
    from cs744_ddp_b200.data.pipeline import RawBatch
    from cs744_ddp_b200.engine import NativeBackend
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    sync = {"1": "none", "2a": "gather_scatter", "2b": "allreduce", "3": "bucketed"}[a.part] if world > 1 else "none"
    cfg = TrainConfig(num_nodes=world, rank=rank, synthetic=True, global_batch=a.batch * world, device="cuda",
                      backend="native", dtype=a.dtype, sync=sync, bucket_plan=a.bucket_plan)
    be = NativeBackend(cfg, dev)
    g = torch.Generator().manual_seed(rank)
    raw = RawBatch(torch.randint(0, 256, (a.batch, 32, 32, 3), generator=g, dtype=torch.uint8).cuda(),
                   torch.randint(0, 10, (a.batch,), generator=g).cuda(), None, None)

    class T:
        def mark(self, i):
            pass

    with torch.cuda.stream(be.stream):
        data, target = be.prepare(raw)
        for _ in range(5):
            be.train_step(data, target, T())
        runs = [be.trace_step(data, target, mode=a.mode) for _ in range(a.reps)]
    # median over repetitions, row by row
    rows = runs[0]
    med = lambda vals: sorted(vals)[len(vals) // 2]
    for i, r in enumerate(rows):
        for k in ("begin_ms", "end_ms", "ms"):
            r[k] = med([x[i][k] for x in runs])
    if rank == 0:
        lines = [f"# step timeline, {world} GPU(s), part {a.part}, comm {be.comm_name()}, per-GPU batch {a.batch}, {a.dtype}, "
                 f"mode {a.mode}, median of {a.reps}",
                 f"{'kernel':34s} {'blk':>3s} {'stream':>6s} {'begin_us':>9s} {'end_us':>9s} {'us':>8s} {'gap_us':>7s}"]
        per_kind = collections.defaultdict(float)
        prev_end = {}
        gaps = 0.0
        for r in rows:
            gap = r["begin_ms"] - prev_end.get(r["stream"], 0.0) if a.mode == "stamps" and r["begin_ms"] >= 0 else 0.0
            lines.append(f"{r['kernel']:34s} {r['block']:3d} {r['stream']:>6s} {r['begin_ms'] * 1e3:9.1f} {r['end_ms'] * 1e3:9.1f} "
                         f"{r['ms'] * 1e3:8.1f} {gap * 1e3:7.1f}")
            if r["stream"] == "main":
                per_kind[r["kernel"]] += r["ms"] * 1e3
                gaps += max(gap, 0.0) * 1e3 if prev_end.get("main") is not None else 0.0
            prev_end[r["stream"]] = r["end_ms"]
        lines.append("# main-stream (critical path) totals by entry point, us:")
        for k, v in sorted(per_kind.items(), key=lambda kv: -kv[1]):
            lines.append(f"#   {k:34s} {v:8.1f}")
        lines.append(f"#   {'gaps between main-stream kernels':34s} {gaps:8.1f}   (negative gaps = overlap through programmatic launch)")
        lines.append(f"#   {'STEP (last end on any stream)':34s} {max(r['end_ms'] for r in rows) * 1e3:8.1f}")
        text = "\n".join(lines)
        print(text)
        if a.out:
            os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
            with open(a.out, "w") as f:
                f.write(text + "\n")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
