"""``NativeBackend``: the trainer-facing wrapper of the C++/CUDA step engine (csrc/runtime/engine.cpp).

Same surface as ``TorchBackend`` (prepare / train_step / eval_step / state_dict / ...), but every
FLOP and every byte of the training step runs in this repository's sm_100a kernels:

* inputs: ``ddpb_augment`` writes the normalised NHWC batch straight into the engine's static input
  buffer (CUDA-graph friendly);
* step: ``ddpb_engine_forward`` + ``ddpb_engine_backward`` (two graph launches, so the reference's
  forward / backward timing regions stay observable) or ``ddpb_engine_train_step`` (one launch);
* gradient sync: the engine's bucket scheduler launches the fused peer-to-peer reduce+SGD kernels
  over NVLink (``comm = p2p``).  ``comm = nccl`` keeps the kernels but reduces the flat gradient
  arena with NCCL and then runs the fused SGD kernel - the baseline the P2P path is measured
  against, and the fallback when peer mappings cannot be established.

Checkpoint contract: ``state_dict()`` returns the reference's 58 keys (fp32, OIHW) no matter how the
engine stores its weights internally (flat arena in gradient-ready order, conv weights KRSC, plus a
bf16 compute copy).
"""
from __future__ import annotations

import contextlib
import ctypes
import os
from typing import Any, Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

from .config import TrainConfig
from .data.pipeline import RawBatch
from .models.vgg import (ConvBlock, build_model, conv_blocks, krsc_to_oihw, oihw_to_krsc, param_specs,
                         state_dict_keys, strip_module_prefix)
from .ops import functional as F
from .ops import native as nv
from .ops.native import check, lib, stream_handle

MAX_BLOCKS, MAX_SLOTS = 24, 64
# comm="auto" tries the NVSwitch-multicast (NVLS) path first on 3+ GPUs; between two GPUs the unicast kernels win
# (36.9 MB reduce + SGD: 78 us against 114 us; bucketed step 0.575 ms against 0.595 ms).  DDPB_NVLS=0 / 1 overrides.
def _nvls_default(world):
    return "1" if world > 2 else "0"
STRATEGY_CODE = {"none": 0, "gather_scatter": 1, "allreduce": 2, "bucketed": 3}
PLAN_CODE = {"ddp": 0, "per_layer": 1, "single": 2}


class EngineConfig(ctypes.Structure):
    _fields_ = [
        ("dtype", ctypes.c_int), ("max_batch", ctypes.c_int), ("world", ctypes.c_int), ("rank", ctypes.c_int),
        ("strategy", ctypes.c_int), ("num_blocks", ctypes.c_int),
        ("cin", ctypes.c_int * MAX_BLOCKS), ("cout", ctypes.c_int * MAX_BLOCKS),
        ("hw", ctypes.c_int * MAX_BLOCKS), ("pool", ctypes.c_int * MAX_BLOCKS),
        ("lr", ctypes.c_float), ("momentum", ctypes.c_float), ("wd", ctypes.c_float),
        ("bucket_plan", ctypes.c_int), ("first_bucket_mb", ctypes.c_float), ("bucket_cap_mb", ctypes.c_float),
        ("use_graph", ctypes.c_int), ("use_tc", ctypes.c_int), ("external_sync", ctypes.c_int),
        ("external_arena", ctypes.c_void_p), ("external_arena_bytes", ctypes.c_size_t),
    ]


class EngineInfo(ctypes.Structure):
    _fields_ = [
        ("arena", ctypes.c_void_p), ("arena_bytes", ctypes.c_size_t), ("ipc_handle", ctypes.c_ubyte * 64),
        ("arena_elems", ctypes.c_longlong), ("total_channels", ctypes.c_longlong),
        ("grad", ctypes.c_void_p), ("master", ctypes.c_void_p), ("mom", ctypes.c_void_p),
        ("shadow", ctypes.c_void_p), ("running_mean", ctypes.c_void_p), ("running_var", ctypes.c_void_p),
        ("num_batches_tracked", ctypes.c_void_p), ("input", ctypes.c_void_p), ("target", ctypes.c_void_p),
        ("loss", ctypes.c_void_p), ("eval_loss", ctypes.c_void_p), ("correct", ctypes.c_void_p),
        ("logits", ctypes.c_void_p), ("num_buckets", ctypes.c_int),
        ("bucket_off", ctypes.c_longlong * MAX_SLOTS), ("bucket_numel", ctypes.c_longlong * MAX_SLOTS),
        ("bucket_mode", ctypes.c_int * MAX_SLOTS), ("bucket_last_block", ctypes.c_int * MAX_SLOTS),
        ("fc_w_off", ctypes.c_longlong), ("fc_b_off", ctypes.c_longlong),
        ("conv_w_off", ctypes.c_longlong * MAX_BLOCKS), ("conv_b_off", ctypes.c_longlong * MAX_BLOCKS),
        ("bn_w_off", ctypes.c_longlong * MAX_BLOCKS), ("bn_b_off", ctypes.c_longlong * MAX_BLOCKS),
        ("ch_off", ctypes.c_longlong * MAX_BLOCKS),
        ("y", ctypes.c_void_p * MAX_BLOCKS), ("act", ctypes.c_void_p * MAX_BLOCKS),
        ("launches", ctypes.c_longlong), ("graph_active", ctypes.c_int), ("multicast", ctypes.c_int),
    ]


def _view(address: int, numel: int, dtype: torch.dtype, device: torch.device) -> torch.Tensor:
    nbytes = numel * torch.empty((), dtype=dtype).element_size()
    return nv.DeviceBuffer(address, nbytes).as_tensor(device).view(dtype)


class HostScalar:
    """A scalar on its way to pinned host memory (``NativeBackend.loss_to_host``)."""

    def __init__(self, buf: torch.Tensor, index: int, event: "torch.cuda.Event") -> None:
        self._buf, self._index, self._event = buf, index, event

    def item(self) -> float:
        self._event.synchronize()
        return float(self._buf[self._index])

    __float__ = item


class NativeBackend:
    name = "native"

    def __init__(self, cfg: TrainConfig, device: torch.device, init_state: Optional[Dict[str, torch.Tensor]] = None,
                 max_batch: Optional[int] = None) -> None:
        if not nv.available():
            raise nv.NativeError("native backend requested but no CUDA device / library")
        self.cfg = cfg
        self.device = device
        self.blocks: List[ConvBlock] = conv_blocks(cfg.model)
        self.dtype = torch.bfloat16 if cfg.dtype == "bf16" else torch.float32
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.max_batch = int(max_batch or max(cfg.per_rank_batch, 1))
        self.training = True
        # The engine runs on its own stream: the legacy default stream cannot be captured into a
        # CUDA graph.  ``on_stream()`` orders it after/before the caller's current stream.
        self.stream = torch.cuda.Stream(device, priority=int(os.environ.get("DDPB_PRIO_MAIN", "0")))
        self._lib = lib()
        self._handle = ctypes.c_void_p()
        self._comm = "none"
        want_p2p = self.world > 1 and cfg.comm in ("auto", "p2p", "nvls")
        self._vmm = None
        if want_p2p and (cfg.comm == "nvls" or (cfg.comm == "auto" and os.environ.get("DDPB_NVLS", _nvls_default(self.world)) == "1")):
            self._vmm = self._try_vmm_arena()           # NVSwitch multicast (NVLS) when the box exposes it
            if self._vmm is None and cfg.comm == "nvls":
                raise nv.NativeError("comm=nvls requested but the multicast arena could not be set up")
        if self._vmm is not None:
            self._create(external_sync=False, arena=self._vmm)
            table = (ctypes.c_void_p * self.world)(*[ctypes.c_void_p(p) for p in self._vmm.peers])
            check(self._lib.ddpb_engine_set_peers(self._handle, table), "engine_set_peers")
            check(self._lib.ddpb_engine_set_multicast(self._handle, ctypes.c_void_p(self._vmm.mc)), "set_multicast")
            self._comm = "nvls"
            dist.barrier()
        else:
            self._create(external_sync=self.world > 1 and not want_p2p)
            if want_p2p:
                ok = self._connect_peers()
                if not ok:
                    if cfg.comm == "p2p":
                        raise nv.NativeError("comm=p2p requested but peer mappings could not be established")
                    self._destroy()
                    self._create(external_sync=True)
        if self.world > 1 and self._comm not in ("p2p", "nvls"):
            self._comm = "nccl" if dist.get_backend() == "nccl" else dist.get_backend()
        # weights: same construction order and RNG consumption as the reference (seeded by the trainer)
        sd = init_state if init_state is not None else build_model(cfg.model).state_dict()
        self.load_state_dict(sd)
        if self.world > 1:
            self.verify_plan_across_ranks()             # DDP checks parameter shapes before it broadcasts
        if cfg.sync == "bucketed" and self.world > 1:
            self.broadcast_from_rank0()                 # DDP init: params + buffers from rank 0
        torch.cuda.synchronize(device)

    # ---- engine life cycle ---------------------------------------------------------------------
    def _engine_config(self, external_sync: bool) -> "EngineConfig":
        cfg = self.cfg
        ec = EngineConfig()
        ec.dtype = nv.BF16 if cfg.dtype == "bf16" else nv.F32
        ec.max_batch = self.max_batch
        ec.world, ec.rank = self.world, self.rank
        ec.strategy = STRATEGY_CODE[cfg.sync]
        ec.num_blocks = len(self.blocks)
        for i, b in enumerate(self.blocks):
            ec.cin[i], ec.cout[i], ec.hw[i], ec.pool[i] = b.cin, b.cout, b.hw, int(b.pool)
        ec.lr, ec.momentum, ec.wd = cfg.lr, cfg.momentum, cfg.weight_decay
        ec.bucket_plan = PLAN_CODE[cfg.bucket_plan]
        ec.first_bucket_mb, ec.bucket_cap_mb = cfg.first_bucket_mb, cfg.bucket_cap_mb
        ec.use_graph = int(cfg.use_cuda_graph)
        ec.use_tc = 1
        ec.external_sync = int(external_sync)
        return ec

    def _try_vmm_arena(self):
        """Allocate the symmetric arena with the VMM API and bind it to a multicast object; None when
        the box does not expose NVLS (every rank takes the same decision)."""
        from .parallel.vmm import VmmArena
        need = ctypes.c_size_t(0)
        ec = self._engine_config(False)
        check(self._lib.ddpb_engine_arena_bytes(ctypes.byref(ec), ctypes.byref(need)), "engine_arena_bytes")
        return VmmArena.create(int(need.value), self.device)

    def _create(self, external_sync: bool, arena=None) -> None:
        ec = self._engine_config(external_sync)
        if arena is not None:
            ec.external_arena = ctypes.c_void_p(arena.local)
            ec.external_arena_bytes = arena.bytes
        self._external_sync = external_sync
        torch.cuda.set_device(self.device)
        check(self._lib.ddpb_engine_create(ctypes.byref(ec), ctypes.byref(self._handle)), "engine_create")
        self._refresh_info()
        info, dev = self.info, self.device
        n = int(info.arena_elems)
        self.master = _view(info.master, n, torch.float32, dev)
        self.grad = _view(info.grad, n, torch.float32, dev)
        self.mom = _view(info.mom, n, torch.float32, dev)
        self.shadow = _view(info.shadow, n, torch.bfloat16, dev)
        c = int(info.total_channels)
        self.running_mean = _view(info.running_mean, c, torch.float32, dev)
        self.running_var = _view(info.running_var, c, torch.float32, dev)
        self.nbt = _view(info.num_batches_tracked, len(self.blocks), torch.int64, dev)
        self.input = _view(info.input, self.max_batch * 32 * 32 * 3, self.dtype, dev).view(self.max_batch, 32, 32, 3)
        self.target = _view(info.target, self.max_batch, torch.int64, dev)
        self.loss = _view(info.loss, 1, torch.float32, dev)
        self.eval_loss = _view(info.eval_loss, 1, torch.float32, dev)
        self.correct = _view(info.correct, 1, torch.int32, dev)
        self.logits = _view(info.logits, self.max_batch * 10, torch.float32, dev).view(self.max_batch, 10)
        self.running_var.fill_(1.0)

    def _refresh_info(self) -> None:
        self.info = EngineInfo()
        check(self._lib.ddpb_engine_info(self._handle, ctypes.byref(self.info)), "engine_info")

    def _destroy(self) -> None:
        if self._handle:
            self._lib.ddpb_engine_destroy(self._handle)
            self._handle = ctypes.c_void_p()

    def __del__(self) -> None:
        try:
            self._destroy()
        except Exception:
            pass

    def _connect_peers(self) -> bool:
        """Exchange CUDA IPC handles over the control plane and map every peer's arena."""
        mine = bytes(self.info.ipc_handle)
        handles: List[Any] = [None] * self.world
        dist.all_gather_object(handles, mine)
        table = (ctypes.c_void_p * self.world)()
        ok = 1
        try:
            for r, h in enumerate(handles):
                if r == self.rank:
                    table[r] = self.info.arena
                    continue
                p = ctypes.c_void_p()
                buf = (ctypes.c_ubyte * 64).from_buffer_copy(h)
                check(self._lib.ddpb_symm_open(buf, ctypes.byref(p)), f"symm_open(rank {r})")
                table[r] = p
        except nv.NativeError as e:
            print(f"[ddpb] rank {self.rank}: peer mapping failed: {e}", flush=True)
            ok = 0
        flag = torch.tensor([ok], device=self.device, dtype=torch.int32)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag) == 0:
            return False
        check(self._lib.ddpb_engine_set_peers(self._handle, table), "engine_set_peers")
        self._comm = "p2p"
        dist.barrier()
        return True

    @contextlib.contextmanager
    def on_stream(self):
        """Run the enclosed work on the engine stream, ordered after everything already queued on
        the caller's stream and before whatever the caller queues next.  Free when the caller
        already made the engine stream current (the trainer and bench.py do)."""
        cur = torch.cuda.current_stream(self.device)
        if cur.cuda_stream == self.stream.cuda_stream:
            yield
            return
        self.stream.wait_stream(cur)
        with torch.cuda.stream(self.stream):
            yield
        cur.wait_stream(self.stream)

    # ---- data --------------------------------------------------------------------------------------
    def stage(self, batch: RawBatch) -> RawBatch:
        """Start the host -> device copy of a (pinned) host batch on the backend's copy stream and return
        the device batch; ``prepare`` orders itself after the copy.  Two staging slots alternate, so the
        copy of batch i+1 runs while step i computes (the loader's prefetch; the reference gets the
        same effect from ``DataLoader`` workers, ref ``Part 1/main.py:53-58``)."""
        n = batch.size
        if n > self.max_batch:
            raise nv.NativeError(f"batch {n} exceeds the engine's max batch {self.max_batch}")
        if not hasattr(self, "_stage_slots"):
            dev, mb = self.device, self.max_batch
            self._copy_stream = torch.cuda.Stream(dev)
            self._stage_slots = [dict(images=torch.empty((mb, 32, 32, 3), dtype=torch.uint8, device=dev),
                                labels=torch.empty(mb, dtype=torch.int64, device=dev),
                                crop=torch.empty((mb, 2), dtype=torch.int32, device=dev),
                                flip=torch.empty(mb, dtype=torch.uint8, device=dev),
                                ready=torch.cuda.Event(), consumed=None) for _ in range(2)]
            self._slot_next = 0
        slot = self._stage_slots[self._slot_next]
        self._slot_next ^= 1
        cs = self._copy_stream
        if slot["consumed"] is not None:
            cs.wait_event(slot["consumed"])           # the augmentation kernel that read this slot has run
        with torch.cuda.stream(cs):
            slot["images"][:n].copy_(batch.images, non_blocking=True)
            slot["labels"][:n].copy_(batch.labels, non_blocking=True)
            if batch.crop_xy is not None:
                slot["crop"][:n].copy_(batch.crop_xy, non_blocking=True)
            if batch.flip is not None:
                slot["flip"][:n].copy_(batch.flip, non_blocking=True)
            slot["ready"].record(cs)
        out = RawBatch(slot["images"][:n], slot["labels"][:n],
                       None if batch.crop_xy is None else slot["crop"][:n],
                       None if batch.flip is None else slot["flip"][:n])
        out._slot = slot                                # prepare() waits for `ready` and marks `consumed`
        return out

    def loss_to_host(self, loss: torch.Tensor) -> "HostScalar":
        """Queue a device -> host copy of a step's loss into pinned memory; ``.item()`` on the result
        waits for that copy only, so the host can queue the next step before it reads this one."""
        if not hasattr(self, "_loss_host"):
            self._loss_host = torch.empty(4, dtype=torch.float32).pin_memory()
            self._loss_ev = [torch.cuda.Event() for _ in range(4)]
            self._loss_next = 0
        k = self._loss_next
        self._loss_next = (k + 1) % 4
        with self.on_stream():
            self._loss_host[k:k + 1].copy_(loss.reshape(1), non_blocking=True)
            self._loss_ev[k].record(torch.cuda.current_stream(self.device))
        return HostScalar(self._loss_host, k, self._loss_ev[k])

    def prepare(self, batch: RawBatch) -> Tuple[torch.Tensor, torch.Tensor]:
        n = batch.size
        if n > self.max_batch:
            raise nv.NativeError(f"batch {n} exceeds the engine's max batch {self.max_batch}")
        slot = getattr(batch, "_slot", None)
        if slot is not None:                            # staged batch: order after its copy, release the slot after use
            with self.on_stream():
                cur = torch.cuda.current_stream(self.device)
                cur.wait_event(slot["ready"])
                F.augment(batch.images, None, batch.crop_xy, batch.flip, self.dtype, out=self.input, batch=n,
                          labels=(batch.labels, self.target))           # images and targets in one launch
                slot["consumed"] = torch.cuda.Event()
                slot["consumed"].record(cur)
            return self.input[:n], self.target[:n]
        with self.on_stream():
            b = batch.to(self.device)
            F.augment(b.images, None, b.crop_xy, b.flip, self.dtype, out=self.input, batch=n,
                      labels=(b.labels.contiguous(), self.target) if b.labels.dtype == torch.int64 else None)
            if b.labels.dtype != torch.int64:
                self.target[:n].copy_(b.labels, non_blocking=True)
        return self.input[:n], self.target[:n]

    def _stage(self, data: torch.Tensor, target: torch.Tensor) -> int:
        n = int(data.shape[0])
        if data.data_ptr() != self.input.data_ptr():
            if data.dim() == 4 and data.shape[1] == 3 and data.shape[-1] != 3:      # NCHW given
                data = data.permute(0, 2, 3, 1)
            self.input[:n].copy_(data.to(self.dtype))
        if target.data_ptr() != self.target.data_ptr():
            self.target[:n].copy_(target)
        return n

    # ---- training ----------------------------------------------------------------------------------
    def set_train(self, flag: bool) -> None:
        self.training = bool(flag)

    def train_step(self, data: torch.Tensor, target: torch.Tensor, timer) -> torch.Tensor:
        split = hasattr(timer, "is_async")            # region timers want the forward/backward seam
        timer.mark(0)
        if split or self._external_sync:
            with self.on_stream():
                n = self._stage(data, target)
                check(self._lib.ddpb_engine_forward(self._handle, n, stream_handle()), "engine_forward")
            timer.mark(1)
            with self.on_stream():
                check(self._lib.ddpb_engine_backward(self._handle, n, stream_handle()), "engine_backward")
                if self._external_sync:
                    self._nccl_sync_and_step()
        else:
            with self.on_stream():
                n = self._stage(data, target)
                check(self._lib.ddpb_engine_train_step(self._handle, n, stream_handle()), "engine_train_step")
            timer.mark(1)
        timer.mark(2)
        return self.loss[0]

    def _nccl_sync_and_step(self) -> None:
        """Baseline / fallback: reduce the flat gradient arena with torch.distributed, then the fused
        SGD kernel.  Strategy semantics follow the reference (ref Part 2a/main.py:117-127, 2b:116-119)."""
        cfg, w = self.cfg, self.world
        g = self.grad
        if cfg.sync == "gather_scatter":
            if self.rank == 0:
                inbox = [torch.empty_like(g) for _ in range(w)]
                dist.gather(g, inbox, dst=0)
                mean = torch.stack(inbox).mean(dim=0)
                dist.scatter(g, [mean for _ in range(w)], src=0)
            else:
                dist.gather(g, dst=0)
                dist.scatter(g, src=0)
            scale = 1.0
        else:
            dist.all_reduce(g, op=dist.ReduceOp.SUM)
            scale = 1.0 / w
        F.sgd_step(self.master, self.mom, g, self.shadow if self.dtype == torch.bfloat16 else None, cfg.lr,
                   cfg.momentum, cfg.weight_decay, scale)

    # ---- evaluation ----------------------------------------------------------------------------------
    def before_eval(self) -> None:
        """Part 3: every rank evaluates with rank 0's BN running statistics (DDP's buffer broadcast,
        SURVEY 2.4.2 X5).  Broadcasting once before evaluation is observably identical to DDP's
        per-forward broadcast because training forwards never read the running statistics."""
        if self.cfg.sync == "bucketed" and self.world > 1:
            self.sync_buffers()

    def eval_step(self, data: torch.Tensor, target: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        with self.on_stream():
            n = self._stage(data, target)
            s = stream_handle()
            check(self._lib.ddpb_engine_eval_reset(self._handle, s), "eval_reset")
            check(self._lib.ddpb_engine_eval_step(self._handle, n, s), "engine_eval_step")
            return self.eval_loss[0].clone(), self.correct[0].clone()

    # ---- collectives -----------------------------------------------------------------------------------
    def sync_buffers(self) -> None:
        if self.world == 1:
            return
        if self._comm in ("p2p", "nvls"):
            check(self._lib.ddpb_engine_collective(self._handle, 1, stream_handle()), "bcast buffers")
        else:
            dist.broadcast(self.running_mean, src=0)
            dist.broadcast(self.running_var, src=0)
        dist.broadcast(self.nbt, src=0)

    def plan_signature(self) -> List[int]:
        """Everything the cross-rank kernels assume to be identical on all ranks: the parameter shapes
        (in arena order), the arena size and the bucket ranges / modes they will be called with."""
        sig: List[int] = [int(self.info.arena_elems), int(self.info.total_channels), int(self.info.num_buckets),
                          1 if self.dtype == torch.bfloat16 else 0, STRATEGY_CODE[self.cfg.sync]]
        for _, (off, shape, _) in sorted(self._slots().items(), key=lambda kv: kv[1][0]):
            sig.append(off)
            sig.extend(int(d) for d in shape)
        for b in self.buckets():
            sig.extend(b)
        return sig

    def verify_plan_across_ranks(self) -> None:
        """``DDP.__init__`` verifies parameter shapes across processes before its first broadcast
        (torch ``nn/parallel/distributed.py:881``, which ref ``Part 3/main.py:61`` gets implicitly).
        Here a mismatch would be worse than a wrong answer: the peer-to-peer kernels index every
        rank's arena with THIS rank's offsets.  All ranks compare their layout with rank 0's and fail
        together."""
        if self.world == 1:
            return
        mine = torch.tensor(self.plan_signature(), dtype=torch.int64, device=self.device)
        n = torch.tensor([mine.numel()], dtype=torch.int64, device=self.device)
        n0 = n.clone()
        dist.broadcast(n0, src=0)
        same = int(n0) == int(n)
        if same:
            ref = mine.clone()
            dist.broadcast(ref, src=0)
            same = bool(torch.equal(ref, mine))
        else:
            dist.broadcast(torch.empty(int(n0), dtype=torch.int64, device=self.device), src=0)
        ok = torch.tensor([1 if same else 0], dtype=torch.int32, device=self.device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok) == 0:
            raise RuntimeError(f"rank {self.rank}: model / arena / bucket plan differs across ranks "
                               f"({'this rank differs from rank 0' if not same else 'another rank differs'}); "
                               "all ranks must be started with the same --model, --dtype, --sync and bucket options")

    def broadcast_from_rank0(self) -> None:
        if self.world == 1:
            return
        if self._comm in ("p2p", "nvls"):
            torch.cuda.synchronize(self.device)
            dist.barrier()
            self.gather_master()                    # rank 0 must hold every slice before it is the source
            check(self._lib.ddpb_engine_collective(self._handle, 0, stream_handle()), "bcast weights")
        else:
            dist.broadcast(self.master, src=0)
            check(self._lib.ddpb_engine_refresh_shadow(self._handle, stream_handle()), "refresh_shadow")
        self.sync_buffers()

    def gather_master(self) -> None:
        """The fp32 master of a conv weight lives, current, only on the rank that owns its slice (peers get
        the bf16 copy they compute with); bring all slices everywhere before the weights are read in fp32
        (``state_dict``, a broadcast, a consistency check).  A collective: every rank must call it."""
        if self.world > 1 and self._comm in ("p2p", "nvls"):
            check(self._lib.ddpb_engine_collective(self._handle, 3, stream_handle()), "gather master")

    def gather_momentum(self) -> None:
        """The two-shot reduce keeps each momentum slice only on the rank that owns it; bring all
        shards everywhere before the optimizer state is read (checkpoint)."""
        if self.world > 1 and self._comm in ("p2p", "nvls"):
            check(self._lib.ddpb_engine_collective(self._handle, 2, stream_handle()), "gather momentum")

    # ---- parameter access ---------------------------------------------------------------------------------
    def _slots(self) -> Dict[str, Tuple[int, Tuple[int, ...], str]]:
        """name -> (arena offset, torch shape, kind)"""
        info = self.info
        out: Dict[str, Tuple[int, Tuple[int, ...], str]] = {}
        for i, b in enumerate(self.blocks):
            out[f"layers.{b.seq_conv}.weight"] = (int(info.conv_w_off[i]), (b.cout, b.cin, 3, 3), "conv_w")
            out[f"layers.{b.seq_conv}.bias"] = (int(info.conv_b_off[i]), (b.cout,), "vec")
            out[f"layers.{b.seq_bn}.weight"] = (int(info.bn_w_off[i]), (b.cout,), "vec")
            out[f"layers.{b.seq_bn}.bias"] = (int(info.bn_b_off[i]), (b.cout,), "vec")
        out["fc1.weight"] = (int(info.fc_w_off), (10, 512), "mat")
        out["fc1.bias"] = (int(info.fc_b_off), (10,), "vec")
        return out

    def _export(self, flat: torch.Tensor) -> Dict[str, torch.Tensor]:
        out = {}
        for name, (off, shape, kind) in self._slots().items():
            n = 1
            for s_ in shape:
                n *= s_
            t = flat[off: off + n]
            if kind == "conv_w":
                t = krsc_to_oihw(t.view(shape[0], 3, 3, shape[1]))
            else:
                t = t.view(shape).clone()
            out[name] = t
        return out

    def _import(self, flat: torch.Tensor, tensors: Dict[str, torch.Tensor]) -> None:
        for name, (off, shape, kind) in self._slots().items():
            t = tensors[name].to(self.device, torch.float32)
            if tuple(t.shape) != tuple(shape):
                raise ValueError(f"{name}: shape {tuple(t.shape)} != expected {shape}")
            if kind == "conv_w":
                t = oihw_to_krsc(t)
            flat[off: off + t.numel()].copy_(t.reshape(-1))

    def state_dict(self) -> Dict[str, torch.Tensor]:
        """The reference's 58 keys.  On a multi-GPU native run this is a collective (see ``gather_master``)."""
        self.gather_master()
        params = self._export(self.master)
        sd: Dict[str, torch.Tensor] = {}
        for key in state_dict_keys(self.cfg.model):
            if key in params:
                sd[key] = params[key]
                continue
            mod, leaf = key.rsplit(".", 1)
            i = next(j for j, b in enumerate(self.blocks) if f"layers.{b.seq_bn}" == mod)
            c0 = int(self.info.ch_off[i])
            c = self.blocks[i].cout
            if leaf == "running_mean":
                sd[key] = self.running_mean[c0: c0 + c].clone()
            elif leaf == "running_var":
                sd[key] = self.running_var[c0: c0 + c].clone()
            else:
                sd[key] = self.nbt[i].clone()
        return sd

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        sd = strip_module_prefix(sd)
        missing = [k for k in state_dict_keys(self.cfg.model) if k not in sd]
        if missing:
            raise KeyError(f"state_dict is missing keys: {missing[:4]}...")
        self.master.zero_()
        self._import(self.master, sd)
        for i, b in enumerate(self.blocks):
            c0, c = int(self.info.ch_off[i]), b.cout
            self.running_mean[c0: c0 + c].copy_(sd[f"layers.{b.seq_bn}.running_mean"])
            self.running_var[c0: c0 + c].copy_(sd[f"layers.{b.seq_bn}.running_var"])
            self.nbt[i].copy_(sd[f"layers.{b.seq_bn}.num_batches_tracked"])
        check(self._lib.ddpb_engine_refresh_shadow(self._handle, stream_handle()), "refresh_shadow")

    def optimizer_state_dict(self) -> Dict[str, Any]:
        """Same layout as ``torch.optim.SGD.state_dict()`` for the reference's parameter order."""
        self.gather_momentum()
        bufs = self._export(self.mom)
        names = [s.name for s in param_specs(self.cfg.model)]
        return {
            "state": {i: {"momentum_buffer": bufs[n]} for i, n in enumerate(names)},
            "param_groups": [{"lr": self.cfg.lr, "momentum": self.cfg.momentum, "dampening": 0,
                              "weight_decay": self.cfg.weight_decay, "nesterov": False,
                              "params": list(range(len(names)))}],
        }

    def load_optimizer_state_dict(self, sd: Optional[Dict[str, Any]]) -> None:
        if not sd or not sd.get("state"):
            return
        names = [s.name for s in param_specs(self.cfg.model)]
        bufs = {names[int(i)]: st["momentum_buffer"] for i, st in sd["state"].items()
                if st.get("momentum_buffer") is not None}
        if len(bufs) == len(names):
            self.mom.zero_()
            self._import(self.mom, bufs)

    def grads(self) -> Dict[str, torch.Tensor]:
        return self._export(self.grad)

    # ---- misc ----------------------------------------------------------------------------------------
    def buckets(self) -> List[Tuple[int, int, int, int]]:
        i = self.info
        return [(int(i.bucket_off[b]), int(i.bucket_numel[b]), int(i.bucket_mode[b]), int(i.bucket_last_block[b]))
                for b in range(int(i.num_buckets))]

    def profile_step(self, data: torch.Tensor, target: torch.Tensor) -> Dict[str, Any]:
        """One instrumented (eager) training step: device timeline of forward, backward and every
        bucket's fused reduce.  ``exposed_comm_ms`` is what the gradient sync adds after the last
        backward kernel finished; the per-bucket rows show how far ahead of that each reduce ended."""
        nb = int(self.info.num_buckets)
        buf = (ctypes.c_float * (3 + 2 * nb))()
        n = ctypes.c_int(0)
        with self.on_stream():
            b = self._stage(data, target)
            check(self._lib.ddpb_engine_profile_step(self._handle, b, stream_handle(), buf, len(buf),
                                                     ctypes.byref(n)), "engine_profile_step")
        fwd, bwd, end = buf[0], buf[1], buf[2]
        rows = [{"bucket": i, "bytes": int(self.info.bucket_numel[i]) * 4, "ready_ms": buf[3 + 2 * i],
                 "done_ms": buf[4 + 2 * i]} for i in range(nb) if buf[3 + 2 * i] >= 0]
        return {"forward_ms": fwd, "backward_compute_end_ms": bwd, "step_ms": end,
                "exposed_comm_ms": end - bwd, "exposed_comm_frac": (end - bwd) / end if end > 0 else 0.0,
                "buckets": rows}

    def trace_step(self, data: torch.Tensor, target: torch.Tensor, mode: str = "stamps") -> List[Dict[str, Any]]:
        """One training step with per-kernel timing on every stream.  ``mode``:

        * ``"stamps"`` - the production CUDA graph, unmodified; every kernel stamps ``%globaltimer`` at its
          first CTA's start and its last warp's end (the real timeline, programmatic launches included);
        * ``"graph"``  - the graph with an event-record node after every kernel (no overlap across nodes);
        * ``"eager"``  - eager launches with an event after each (host-bound for the small kernels).

        Rows in launch order: ``{kernel, block, stream, begin_ms, end_ms, ms}``; for the event modes
        ``begin_ms`` is -1 and ``ms`` is the time since the previous completion on the same stream.
        A real step (weights are updated); on several GPUs every rank must call it."""
        code = {"eager": 0, "graph": 1, "stamps": 2}[mode]
        cap = 256
        names = ctypes.create_string_buffer(48 * cap)
        block = (ctypes.c_int * cap)()
        sid = (ctypes.c_int * cap)()
        begin = (ctypes.c_float * cap)()
        end = (ctypes.c_float * cap)()
        n = ctypes.c_int(0)
        with self.on_stream():
            b = self._stage(data, target)
            check(self._lib.ddpb_engine_trace_step(self._handle, b, stream_handle(), code, names, block, sid, begin, end,
                                                   cap, ctypes.byref(n)), "engine_trace_step")
        rows, last = [], {}
        label = {0: "main", 1: "wgrad", 2: "comm", 3: "comm2", 4: "other"}
        for i in range(n.value):
            name = names.raw[48 * i: 48 * (i + 1)].split(b"\0", 1)[0].decode()
            prev = last.get(sid[i], 0.0)
            dur = float(end[i]) - (float(begin[i]) if code == 2 else prev)
            rows.append({"kernel": name, "block": int(block[i]), "stream": label[int(sid[i])],
                         "begin_ms": float(begin[i]), "end_ms": float(end[i]), "ms": dur})
            last[sid[i]] = float(end[i])
        return rows

    def launch_count(self) -> int:
        self._refresh_info()
        return int(self.info.launches)

    def launch_counts(self) -> Dict[str, int]:
        """Kernel launches per C entry point since the engine was created (graph replays included)."""
        buf = ctypes.create_string_buffer(4096)
        self._lib.ddpb_engine_launch_counts(self._handle, buf, len(buf))
        out: Dict[str, int] = {}
        for item in buf.value.decode().split(";"):
            if "=" in item:
                k, v = item.split("=")
                out[k] = int(v)
        return out

    def comm_name(self) -> str:
        return self._comm

    def uses_graph(self) -> bool:
        self._refresh_info()
        return bool(self.info.graph_active)

    def synchronize(self) -> None:
        torch.cuda.synchronize(self.device)
