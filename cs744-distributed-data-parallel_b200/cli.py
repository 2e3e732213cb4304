"""Command line: the reference's four flags, unchanged, plus optional framework flags.

Reference flags (ref ``Part 2a/main.py:156-165``): ``--master`` (required),
``--num-nodes`` (required), ``--rank`` (required), ``--epochs`` (default 1).  ``--master-ip`` is
accepted as an alias of ``--master`` (the wording BASELINE.json uses).  Part 1 takes no flags
(ref ``Part 1/main.py:124-125``), so nothing is required there.

Every added flag defaults to the reference's hard-coded value, so the same command line gives
the same run.
"""
from __future__ import annotations

import argparse
from typing import Optional, Sequence

from .config import SYNC_STRATEGIES, TrainConfig


def build_parser(part: str = "2a") -> argparse.ArgumentParser:
    distributed = part != "1"
    p = argparse.ArgumentParser(description="Process arguments for training")
    p.add_argument("--master", "--master-ip", dest="master", metavar="master-address",
                   required=distributed, default="127.0.0.1", action="store",
                   help="The IP address of Master")
    p.add_argument("--num-nodes", metavar="total-nodes", required=distributed, default=1,
                   action="store", type=int, help="Total number of nodes")
    p.add_argument("--rank", metavar="rank", required=distributed, default=0, action="store",
                   type=int, help="Rank of this node")
    p.add_argument("--epochs", metavar="epochs", required=False, default=1, action="store",
                   type=int, help="Number of epochs")
    g = p.add_argument_group("framework options (defaults reproduce the reference)")
    g.add_argument("--port", default="6585")
    g.add_argument("--batch-size", type=int, default=256, help="GLOBAL batch (split across ranks)")
    g.add_argument("--model", default="VGG11", choices=["VGG11", "VGG13", "VGG16", "VGG19"])
    g.add_argument("--sync", default=None, choices=list(SYNC_STRATEGIES),
                   help="override the part's gradient-sync strategy")
    g.add_argument("--device", default="auto", choices=["auto", "cpu", "cuda"])
    g.add_argument("--backend", default="auto", choices=["auto", "torch", "native"])
    g.add_argument("--comm", default="auto", choices=["auto", "gloo", "nccl", "p2p", "nvls"])
    g.add_argument("--dtype", default="fp32", choices=["fp32", "bf16"])
    g.add_argument("--synthetic", action="store_true", help="CIFAR-shaped synthetic data")
    g.add_argument("--train-size", type=int, default=50_000)
    g.add_argument("--test-size", type=int, default=10_000)
    g.add_argument("--iters", type=int, default=None, help="stop each epoch after N iterations")
    g.add_argument("--eval-batches", type=int, default=None)
    g.add_argument("--bucket-plan", default="ddp", choices=["ddp", "per_layer", "single"])
    g.add_argument("--bucket-cap-mb", type=float, default=25.0)
    g.add_argument("--checkpoint", default=None, help="save a training checkpoint here at the end")
    g.add_argument("--resume", default=None, help="load a checkpoint before training")
    g.add_argument("--data-root", default="./data")
    g.add_argument("--seed", type=int, default=0)
    g.add_argument("--json-log", action="store_true")
    g.add_argument("--no-cuda-graph", action="store_true")
    return p


_PART_SYNC = {"1": "none", "2a": "gather_scatter", "2b": "allreduce", "3": "bucketed"}


def config_from_args(part: str, argv: Optional[Sequence[str]] = None) -> TrainConfig:
    a = build_parser(part).parse_args(argv)
    return TrainConfig(
        master=a.master, num_nodes=a.num_nodes, rank=a.rank, epochs=a.epochs, port=str(a.port),
        global_batch=a.batch_size, model=a.model, sync=a.sync or _PART_SYNC[part],
        device=a.device, backend=a.backend, comm=a.comm, dtype=a.dtype, synthetic=a.synthetic,
        train_size=a.train_size, test_size=a.test_size, max_iters=a.iters,
        eval_batches=a.eval_batches, bucket_plan=a.bucket_plan, bucket_cap_mb=a.bucket_cap_mb,
        checkpoint=a.checkpoint, resume=a.resume, data_root=a.data_root, seed=a.seed,
        json_log=a.json_log, use_cuda_graph=not a.no_cuda_graph)
