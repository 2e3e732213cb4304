"""Shared driver for the four ``main.py``-style entry points."""
from __future__ import annotations

from typing import Optional, Sequence

import torch

from ..cli import config_from_args
from ..config import TrainConfig
from ..parallel.launch import init_process, pick_backend, shutdown
from ..trainer import Trainer


def run(rank: int, size: int, cfg: TrainConfig):
    """The reference's ``run(rank, size, epochs, batch_size)`` (ref ``Part 2a/main.py:19-68``)."""
    return Trainer(cfg).fit()


def main(part: str, argv: Optional[Sequence[str]] = None):
    cfg = config_from_args(part, argv)
    if part == "1" and cfg.num_nodes == 1:
        return run(0, 1, cfg)                      # Part 1 never touches torch.distributed
    device = cfg.device
    if device == "auto":
        device = "cuda" if torch.cuda.is_available() else "cpu"
    comm = cfg.comm if cfg.comm in ("gloo", "nccl") else "auto"
    backend = pick_backend(device, comm)
    try:
        return init_process(cfg.master, cfg.port, cfg.rank, cfg.num_nodes, run, cfg,
                            backend=backend, device=device)
    finally:
        shutdown()
