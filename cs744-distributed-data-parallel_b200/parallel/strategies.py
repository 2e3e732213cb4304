"""Gradient-synchronisation strategies over ``torch.distributed`` (the ``torch`` back-end).

One trainer, four pluggable strategies ("the only seam that changes"):

=================  ======================================================================
``none``            Part 1: single process, nothing to do.
``gather_scatter``  Part 2a: per parameter, rank 0 gathers every rank's gradient, averages,
                    scatters the mean back            (ref ``Part 2a/main.py:117-127``).
``allreduce``       Part 2b: per parameter all-reduce(SUM) then divide by the world size
                                                      (ref ``Part 2b/main.py:116-119``).
``bucketed``        Part 3: flat gradient buckets, an asynchronous all-reduce launched from
                    an autograd hook the moment a bucket's last gradient lands, overlapping
                    the rest of backward; parameters/buffers broadcast from rank 0 at
                    attach time and BN buffers before every training forward
                    (what the reference gets from ``DDP(model)``, ref ``Part 3/main.py:61``;
                    SURVEY C23 a-f).  This is this framework's own reducer, not torch DDP.
=================  ======================================================================

These run on any process group (gloo on CPU, NCCL on GPUs) and are the oracle for the
peer-to-peer sm_100a engine in ``parallel/p2p.py`` + ``csrc/kernels/sync_*.cu``.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch
import torch.distributed as dist

from .buckets import Bucket, plan_buckets


def _world() -> int:
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def _rank() -> int:
    return dist.get_rank() if dist.is_available() and dist.is_initialized() else 0


class GradSync:
    """Interface.  ``sync`` runs strictly after backward and strictly before the optimizer
    step, exactly where the reference calls its averaging function (``Part 2a/main.py:94-96``)."""
    name = "base"
    overlaps_backward = False

    def attach(self, model: torch.nn.Module) -> None:
        self.model = model
        self.params: List[torch.nn.Parameter] = list(model.parameters())

    def before_forward(self, training: bool = True) -> None:
        pass

    def sync(self) -> None:
        pass

    def collectives_per_iteration(self) -> int:
        return 0


class NoSync(GradSync):
    name = "none"


class GatherScatterSync(GradSync):
    name = "gather_scatter"

    def sync(self) -> None:
        world, rank = _world(), _rank()
        if world == 1:
            return
        for p in self.params:
            g = p.grad
            if rank == 0:
                inbox = [torch.empty_like(g) for _ in range(world)]
                dist.gather(g, inbox, dst=0)
                mean = torch.stack(inbox).mean(dim=0)
                dist.scatter(g, [mean for _ in range(world)], src=0)
            else:
                dist.gather(g, dst=0)
                dist.scatter(g, src=0)

    def collectives_per_iteration(self) -> int:
        return 2 * len(self.params)


class AllReduceSync(GradSync):
    name = "allreduce"

    def sync(self) -> None:
        world = _world()
        if world == 1:
            return
        for p in self.params:
            dist.all_reduce(p.grad, op=dist.ReduceOp.SUM)
            p.grad /= float(world)

    def collectives_per_iteration(self) -> int:
        return len(self.params)


class BucketedOverlapSync(GradSync):
    name = "bucketed"
    overlaps_backward = True

    def __init__(self, plan: str = "ddp", first_cap_mb: float = 1.0, cap_mb: float = 25.0,
                 broadcast_buffers: bool = True) -> None:
        self.plan = plan
        self.first_cap_mb = first_cap_mb
        self.cap_mb = cap_mb
        self.broadcast_buffers = broadcast_buffers
        self.launch_log: List[int] = []       # bucket indices in launch order (last iteration)

    # -- setup --------------------------------------------------------------------------
    def attach(self, model: torch.nn.Module) -> None:
        super().attach(model)
        named = list(model.named_parameters())
        self._verify_shapes(named)
        self._broadcast_state()
        # static ready order: reverse declaration order is what autograd produces for a
        # feed-forward stack; DDP starts from the same assumption before its rebuild.
        order = list(reversed(named))
        names = [n for n, _ in order]
        numels = [p.numel() for _, p in order]
        layer_of = _layer_groups(model, names)
        self.buckets: List[Bucket] = plan_buckets(names, numels, self.plan, self.first_cap_mb,
                                                  self.cap_mb, 4, layer_of)
        by_name = dict(named)
        self._flat: List[torch.Tensor] = []
        self._views: Dict[str, torch.Tensor] = {}
        self._bucket_of: Dict[str, int] = {}
        for b in self.buckets:
            total = sum(by_name[n].numel() for n in b.names)
            ref = by_name[b.names[0]]
            flat = torch.zeros(total, dtype=ref.dtype, device=ref.device)
            off = 0
            for n in b.names:
                p = by_name[n]
                self._views[n] = flat[off: off + p.numel()].view_as(p)
                self._bucket_of[n] = b.index
                off += p.numel()
            self._flat.append(flat)
        self._pending = [len(b.names) for b in self.buckets]
        self._work: List[Optional[dist.Work]] = [None] * len(self.buckets)
        self._hooks = []
        for n, p in named:
            self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(n)))
        self._by_name = by_name

    def _verify_shapes(self, named) -> None:
        if _world() == 1:
            return
        sig = torch.tensor([p.numel() for _, p in named], dtype=torch.int64,
                           device=named[0][1].device)
        ref = sig.clone()
        dist.broadcast(ref, src=0)
        if not torch.equal(sig, ref):
            raise RuntimeError("parameter shapes differ across ranks")

    def _broadcast_state(self) -> None:
        if _world() == 1:
            return
        for t in list(self.model.parameters()) + list(self.model.buffers()):
            dist.broadcast(t.data, src=0)

    # -- per iteration --------------------------------------------------------------------
    def before_forward(self, training: bool = True) -> None:
        """DDP broadcasts module buffers (BN running stats) from rank 0 ahead of the forward."""
        self._pending = [len(b.names) for b in self.buckets]
        self._work = [None] * len(self.buckets)
        self.launch_log = []
        if self.broadcast_buffers and _world() > 1:
            for buf in self.model.buffers():
                dist.broadcast(buf, src=0)

    def _make_hook(self, name: str):
        def hook(param: torch.nn.Parameter) -> None:
            view = self._views[name]
            if param.grad.data_ptr() != view.data_ptr():
                view.copy_(param.grad)
                param.grad = view          # later steps accumulate straight into the bucket
            b = self._bucket_of[name]
            self._pending[b] -= 1
            if self._pending[b] == 0:
                self._launch(b)
        return hook

    def _launch(self, b: int) -> None:
        self.launch_log.append(b)
        world = _world()
        if world == 1:
            return
        flat = self._flat[b]
        flat /= float(world)                                   # DDP divides before reducing
        self._work[b] = dist.all_reduce(flat, op=dist.ReduceOp.SUM, async_op=True)

    def sync(self) -> None:
        """Finalize: wait for every in-flight bucket (the end-of-backward callback in DDP)."""
        for w in self._work:
            if w is not None:
                w.wait()
        # a parameter whose hook never fired this step (unused) still has to agree
        for b, left in enumerate(self._pending):
            if left and left != len(self.buckets[b].names):
                raise RuntimeError(f"bucket {b} only partially ready ({left} gradients missing)")

    def collectives_per_iteration(self) -> int:
        return len(self.buckets) + (sum(1 for _ in self.model.buffers())
                                    if self.broadcast_buffers else 0)


def _layer_groups(model: torch.nn.Module, names: Sequence[str]) -> Dict[str, int]:
    """Group key per parameter: a conv and the BN that follows it share a group; the classifier
    gets its own (it is the first group to become ready)."""
    mods = dict(model.named_modules())
    out: Dict[str, int] = {}
    for n in names:
        mod_name = n.rsplit(".", 1)[0]
        m = mods.get(mod_name)
        if isinstance(m, torch.nn.BatchNorm2d):
            out[n] = int(mod_name.split(".")[-1]) - 1
        elif isinstance(m, torch.nn.Conv2d):
            out[n] = int(mod_name.split(".")[-1])
        else:
            out[n] = 1_000_000
    return out


def make_strategy(name: str, **kw) -> GradSync:
    if name == "none":
        return NoSync()
    if name == "gather_scatter":
        return GatherScatterSync()
    if name == "allreduce":
        return AllReduceSync()
    if name == "bucketed":
        return BucketedOverlapSync(**kw)
    raise ValueError(f"unknown sync strategy {name!r}")
