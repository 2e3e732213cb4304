"""Execution back-ends behind the trainer.

``TorchBackend`` is the reference's computation expressed once: torch modules, autograd,
``optim.SGD(lr .1, momentum .9, weight_decay 1e-4)`` and ``CrossEntropyLoss`` (ref
``Part 1/main.py:110-115``), with a pluggable gradient-sync strategy between ``backward()`` and
``step()`` (ref ``Part 2a/main.py:94-96``).  It runs on CPU (gloo) – config 1 of BASELINE.json –
and on CUDA (NCCL) where it doubles as the numerical oracle for the native engine.

``NativeBackend`` (``engine.py``) has the same surface and runs the sm_100a engine.
"""
from __future__ import annotations

from typing import Any, Dict, Optional, Tuple

import torch

from .config import TrainConfig
from .data.pipeline import RawBatch, transform_batch
from .models.vgg import build_model, strip_module_prefix
from .parallel.strategies import GradSync, make_strategy


class TorchBackend:
    name = "torch"

    def __init__(self, cfg: TrainConfig, device: torch.device) -> None:
        self.cfg = cfg
        self.device = device
        self.model = build_model(cfg.model).to(device)
        self.criterion = torch.nn.CrossEntropyLoss().to(device)
        self.optimizer = torch.optim.SGD(self.model.parameters(), lr=cfg.lr,
                                         momentum=cfg.momentum, weight_decay=cfg.weight_decay)
        kw = {}
        if cfg.sync == "bucketed":
            kw = dict(plan=cfg.bucket_plan, first_cap_mb=cfg.first_bucket_mb, cap_mb=cfg.bucket_cap_mb)
        self.strategy: GradSync = make_strategy(cfg.sync, **kw)
        self.strategy.attach(self.model)
        self.autocast = cfg.dtype == "bf16"
        self.launches_per_step = 0      # no kernels of ours on this path

    # ---- data --------------------------------------------------------------------------
    def prepare(self, batch: RawBatch) -> Tuple[torch.Tensor, torch.Tensor]:
        b = batch.to(self.device)
        return transform_batch(b, "nchw", torch.float32), b.labels

    # ---- one training iteration (regions as the reference times them) -----------------
    def train_step(self, data: torch.Tensor, target: torch.Tensor, timer) -> torch.Tensor:
        timer.mark(0)
        # set_to_none=False keeps bucket views alive for the bucketed reducer
        self.optimizer.zero_grad(set_to_none=not self.strategy.overlaps_backward)
        self.strategy.before_forward(True)
        with torch.autocast(self.device.type, dtype=torch.bfloat16, enabled=self.autocast):
            pred = self.model(data)
        timer.mark(1)
        with torch.autocast(self.device.type, dtype=torch.bfloat16, enabled=self.autocast):
            loss = self.criterion(pred, target)
        loss.backward()
        self.strategy.sync()
        self.optimizer.step()
        timer.mark(2)
        return loss.detach()

    def set_train(self, flag: bool) -> None:
        self.model.train(flag)

    @torch.no_grad()
    def eval_step(self, data: torch.Tensor, target: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        with torch.autocast(self.device.type, dtype=torch.bfloat16, enabled=self.autocast):
            out = self.model(data)
            loss = self.criterion(out, target)
        pred = out.max(1, keepdim=True)[1]
        return loss.float(), pred.eq(target.view_as(pred)).sum()

    def before_eval(self) -> None:
        self.strategy.before_forward(False)

    # ---- state -------------------------------------------------------------------------
    def state_dict(self) -> Dict[str, torch.Tensor]:
        return self.model.state_dict()

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        self.model.load_state_dict(strip_module_prefix(sd), strict=True)

    def optimizer_state_dict(self) -> Dict[str, Any]:
        return self.optimizer.state_dict()

    def load_optimizer_state_dict(self, sd: Optional[Dict[str, Any]]) -> None:
        if sd:
            self.optimizer.load_state_dict(sd)

    def grads(self) -> Dict[str, torch.Tensor]:
        return {n: p.grad.detach().clone() for n, p in self.model.named_parameters()
                if p.grad is not None}

    def synchronize(self) -> None:
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)


def make_backend(cfg: TrainConfig, device: torch.device):
    kind = cfg.backend
    if kind == "auto":
        kind = "native" if device.type == "cuda" else "torch"
    if kind == "torch":
        return TorchBackend(cfg, device)
    if kind == "native":
        if device.type != "cuda":
            raise RuntimeError("the native sm_100a back-end needs a CUDA device; use --backend torch")
        from .engine import NativeBackend
        return NativeBackend(cfg, device)
    raise ValueError(f"unknown backend {kind!r}")
