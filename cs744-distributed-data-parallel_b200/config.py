"""Run configuration.

Every constant the reference hard-codes is a field here, with the reference value as the
default, so ``TrainConfig()`` reproduces the reference run exactly:

* global batch 256, split as ``int(256 / world)``    (ref ``Part 2a/main.py:22,173``)
* SGD lr 0.1, momentum 0.9, weight decay 1e-4          (ref ``Part 1/main.py:114-115``)
* seed 0 on every rank                                 (ref ``Part 2a/main.py:20-21``)
* rendezvous port 6585                                 (ref ``Part 2a/main.py:172``)
* 4 intra-op CPU threads                               (ref ``Part 2a/main.py:15``)
* log window of 20 iterations                          (ref ``Part 1/main.py:47``)
"""
from __future__ import annotations

import dataclasses
from typing import Optional

SYNC_STRATEGIES = ("none", "gather_scatter", "allreduce", "bucketed")

# Which reference script each strategy reproduces; used for log wording.
PART_OF_STRATEGY = {
    "none": "1",
    "gather_scatter": "2a",
    "allreduce": "2b",
    "bucketed": "3",
}

CIFAR_MEAN = tuple(x / 255.0 for x in (125.3, 123.0, 113.9))
CIFAR_STD = tuple(x / 255.0 for x in (63.0, 62.1, 66.7))


@dataclasses.dataclass
class TrainConfig:
    # --- the reference's four CLI flags -------------------------------------------------
    master: str = "127.0.0.1"
    num_nodes: int = 1
    rank: int = 0
    epochs: int = 1
    # --- constants the reference hard-codes ---------------------------------------------
    port: str = "6585"
    global_batch: int = 256
    lr: float = 0.1
    momentum: float = 0.9
    weight_decay: float = 1e-4
    seed: int = 0
    cpu_threads: int = 4
    loader_workers: int = 2
    log_every: int = 20
    data_root: str = "./data"
    # --- framework options (defaults reproduce the reference) ---------------------------
    model: str = "VGG11"
    sync: str = "none"
    device: str = "auto"           # auto | cpu | cuda
    backend: str = "auto"          # auto | torch | native
    comm: str = "auto"             # auto | gloo | nccl | p2p
    dtype: str = "fp32"            # fp32 | bf16
    synthetic: bool = False        # synthetic CIFAR-shaped data instead of ./data
    train_size: int = 50_000
    test_size: int = 10_000
    max_iters: Optional[int] = None    # stop an epoch early (benchmarks / smoke tests)
    eval_batches: Optional[int] = None
    bucket_cap_mb: float = 25.0
    first_bucket_mb: float = 1.0
    bucket_plan: str = "ddp"       # ddp | per_layer | single
    checkpoint: Optional[str] = None   # path to save after training
    resume: Optional[str] = None       # path to load before training
    reproduce_eval_mode_quirk: bool = False  # ref never calls model.train() again (A.3 #1)
    use_cuda_graph: bool = True
    json_log: bool = False
    log_prefix_rank: bool = False

    def __post_init__(self) -> None:
        if self.sync not in SYNC_STRATEGIES:
            raise ValueError(f"sync must be one of {SYNC_STRATEGIES}, got {self.sync!r}")
        if self.dtype not in ("fp32", "bf16"):
            raise ValueError(f"dtype must be fp32 or bf16, got {self.dtype!r}")
        if self.num_nodes < 1 or not (0 <= self.rank < self.num_nodes):
            raise ValueError(f"bad rank/num_nodes: {self.rank}/{self.num_nodes}")

    @property
    def world_size(self) -> int:
        return self.num_nodes

    @property
    def per_rank_batch(self) -> int:
        """``int(256 / float(world))`` – truncating, as the reference does (A.3 #6)."""
        return int(self.global_batch / float(self.world_size))

    @property
    def part(self) -> str:
        return PART_OF_STRATEGY[self.sync]

    def replace(self, **kw) -> "TrainConfig":
        return dataclasses.replace(self, **kw)
