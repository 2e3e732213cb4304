#!/usr/bin/env python
"""Per-kernel timeline of one training step (eager, warm caches, all streams concurrent).

    python tools/step_timeline.py [--batch 256] [--dtype bf16] [--out profiles/step_timeline.txt]

Complements the ncu launch list (serialised, cold caches): this is what the step's critical path looks
like while the weight-gradient and optimizer streams run next to it.  `ms` of a row is the time since
the previous completion on the same stream, i.e. an upper bound of that kernel's duration.
"""
import argparse
import collections
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--out", default=None)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--eager", action="store_true", help="eager launches instead of a CUDA graph with event nodes")
    a = ap.parse_args()
    from cs744_ddp_b200.config import TrainConfig
    from cs744_ddp_b200.data.pipeline import RawBatch
    from cs744_ddp_b200.engine import NativeBackend
    dev = torch.device("cuda", 0)
    cfg = TrainConfig(synthetic=True, global_batch=a.batch, device="cuda", backend="native", dtype=a.dtype)
    be = NativeBackend(cfg, dev)
    g = torch.Generator().manual_seed(0)
    raw = RawBatch(torch.randint(0, 256, (a.batch, 32, 32, 3), generator=g, dtype=torch.uint8).cuda(),
                   torch.randint(0, 10, (a.batch,), generator=g).cuda(), None, None)

    class T:
        def mark(self, i):
            pass

    with torch.cuda.stream(be.stream):
        data, target = be.prepare(raw)
        for _ in range(5):
            be.train_step(data, target, T())
        runs = [be.trace_step(data, target, graphed=not a.eager) for _ in range(a.reps)]
    # median over repetitions, row by row
    rows = runs[0]
    for i, r in enumerate(rows):
        r["ms"] = sorted(x[i]["ms"] for x in runs)[len(runs) // 2]
        r["end_ms"] = sorted(x[i]["end_ms"] for x in runs)[len(runs) // 2]
    lines = [f"# step timeline, batch {a.batch}, {a.dtype}, {'eager' if a.eager else 'CUDA graph'} (one event per launch), median of {a.reps}",
             f"{'kernel':34s} {'blk':>3s} {'stream':>6s} {'end_us':>9s} {'us':>8s}"]
    per_kind = collections.defaultdict(float)
    for r in rows:
        lines.append(f"{r['kernel']:34s} {r['block']:3d} {r['stream']:>6s} {r['end_ms'] * 1e3:9.1f} {r['ms'] * 1e3:8.1f}")
        if r["stream"] == "main":
            per_kind[r["kernel"]] += r["ms"] * 1e3
    lines.append("# main-stream (critical path) totals by entry point, us:")
    for k, v in sorted(per_kind.items(), key=lambda kv: -kv[1]):
        lines.append(f"#   {k:34s} {v:8.1f}")
    lines.append(f"#   {'TOTAL (end of last main-stream kernel)':34s} {max(r['end_ms'] for r in rows if r['stream'] == 'main') * 1e3:8.1f}")
    text = "\n".join(lines)
    print(text)
    if a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        with open(a.out, "w") as f:
            f.write(text + "\n")


if __name__ == "__main__":
    main()
