"""Peer-to-peer collectives over symmetric memory (the GPU gradient engine, outside the trainer).

``SymmetricArena`` is the Python face of ``csrc/runtime/symm_mem.cpp``: every rank allocates the
same layout (gradient / master / momentum / bf16 copy / signal pad), the CUDA IPC handles travel
over the torch.distributed control plane once, and afterwards ``csrc/kernels/sync_p2p.cu`` reads
and writes the peers' buffers directly over NVLink.

Used by the bandwidth sweep (``tools/allreduce_sweep.py``, BASELINE.json config 5) and by the
multi-GPU tests; the training step reaches the same kernels through the C++ engine.

Replaces the reference's ``dist.gather`` / ``dist.scatter`` / ``dist.all_reduce`` calls
(ref ``Part 2a/main.py:117-127``, ``Part 2b/main.py:116-119``).
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import functional as F
from ..ops import native as nv
from ..ops.native import check, lib


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class SymmetricArena:
    """``numel`` fp32 elements of gradient, master weights and momentum plus a bf16 copy and the
    flag pad, allocated identically on every rank and mapped into every peer."""

    FLAG_BYTES = nv.P2P_MAX_SLOTS * 2 * nv.P2P_MAX_BLOCKS * nv.MAX_RANKS * 4

    def __init__(self, numel: int, device: torch.device, with_shadow: bool = True, nvls: bool = False) -> None:
        if not dist.is_initialized():
            raise RuntimeError("SymmetricArena needs an initialised process group (control plane)")
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        if self.world > nv.MAX_RANKS:
            raise RuntimeError(f"at most {nv.MAX_RANKS} ranks per node are supported")
        self.device = device
        self.numel = _round_up(int(numel), 64)
        n4 = self.numel * 4
        self.off_grad, self.off_master, self.off_mom = 0, _round_up(n4, 1024), 2 * _round_up(n4, 1024)
        self.off_shadow = 3 * _round_up(n4, 1024)
        self.off_flags = self.off_shadow + _round_up(self.numel * 2, 1024)
        self.nbytes = self.off_flags + self.FLAG_BYTES
        self._lib = lib()
        torch.cuda.set_device(device)
        self.mc_base = 0                     # multicast mapping (NVLS) or 0
        self._vmm = None
        if nvls:
            from .vmm import VmmArena
            self._vmm = VmmArena.create(self.nbytes, device, tag=f"symm{numel}")
        if self._vmm is not None:
            self.base, self.peer_base, self.mc_base = self._vmm.local, list(self._vmm.peers), self._vmm.mc
        else:
            base = ctypes.c_void_p()
            handle = (ctypes.c_ubyte * 64)()
            check(self._lib.ddpb_symm_alloc(ctypes.c_size_t(self.nbytes), ctypes.byref(base), handle), "symm_alloc")
            self.base = int(base.value)
            handles: List[Optional[bytes]] = [None] * self.world
            dist.all_gather_object(handles, bytes(handle))
            self.peer_base: List[int] = []
            for r, h in enumerate(handles):
                if r == self.rank:
                    self.peer_base.append(self.base)
                    continue
                p = ctypes.c_void_p()
                buf = (ctypes.c_ubyte * 64).from_buffer_copy(h)
                check(self._lib.ddpb_symm_open(buf, ctypes.byref(p)), f"symm_open(rank {r})")
                self.peer_base.append(int(p.value))
        self.with_shadow = with_shadow
        view = lambda off, n, dt: nv.DeviceBuffer(self.base + off, n * torch.empty((), dtype=dt).element_size()) \
            .as_tensor(device).view(dt)
        self.grad = view(self.off_grad, self.numel, torch.float32)
        self.master = view(self.off_master, self.numel, torch.float32)
        self.mom = view(self.off_mom, self.numel, torch.float32)
        self.shadow = view(self.off_shadow, self.numel, torch.bfloat16)
        self.epochs = torch.zeros(nv.P2P_MAX_SLOTS * 2 * nv.P2P_MAX_BLOCKS, dtype=torch.int32, device=device)
        torch.cuda.synchronize(device)
        dist.barrier()

    def _table(self, off: int) -> List[int]:
        return [b + off for b in self.peer_base]

    def sync(self, mode: int, off: int = 0, n: Optional[int] = None, lr: float = 0.0, momentum: float = 0.0,
             wd: float = 0.0, slot: int = 0, blocks: int = 0, stream=None, shard=None) -> None:
        """One fused reduce(+SGD) / broadcast / all-gather over ``[off, off+n)`` (see sync_p2p.cu).
        ``shard``: flat [lo, hi, ...] ranges whose fp32 master is not published (bf16 copy only)."""
        n = self.numel - off if n is None else n
        F.p2p_sync(mode, self._table(self.off_grad), self._table(self.off_master),
                   self._table(self.off_shadow) if self.with_shadow else None, self._table(self.off_flags),
                   self.mom, self.epochs, self.rank, self.world, slot, off, n, lr, momentum, wd, blocks, stream, shard)

    def allgather_master(self, off: int = 0, n: Optional[int] = None, slot: int = 2) -> None:
        """Publish this rank's two-shot slice of the fp32 master (after sharded syncs)."""
        n = self.numel - off if n is None else n
        F.p2p_sync(nv.P2P_ALLGATHER, self._table(self.off_master), None, None, self._table(self.off_flags), None,
                   self.epochs, self.rank, self.world, slot, off, n)

    def sync_nvls(self, off: int = 0, n: Optional[int] = None, lr: float = 0.0, momentum: float = 0.0,
                  wd: float = 0.0, slot: int = 1, blocks: int = 0, stream=None, shard=None,
                  one_shot: bool = False) -> None:
        """Reduce + SGD with the reduction done inside the NVSwitch (needs ``nvls=True``): two-shot (the
        owner of a slice updates it and multicasts the result) or one-shot (every rank pulls the reduced
        sum of the whole range and updates its own replica)."""
        if not self.mc_base:
            raise RuntimeError("this arena has no multicast mapping")
        n = self.numel - off if n is None else n
        tbl = (ctypes.c_void_p * self.world)(*[ctypes.c_void_p(b + self.off_flags) for b in self.peer_base])
        n_sh = len(shard) // 2 if shard else 0
        tab = (ctypes.c_longlong * (2 * n_sh))(*shard) if n_sh else None
        check(self._lib.ddpb_p2p_sync_nvls_sharded(
            nv.P2P_ONESHOT if one_shot else nv.P2P_TWOSHOT,
            ctypes.c_void_p(self.mc_base + self.off_grad), ctypes.c_void_p(self.mc_base + self.off_master),
            ctypes.c_void_p(self.mc_base + self.off_shadow if self.with_shadow else 0), nv.ptr(self.master),
            nv.ptr(self.shadow) if self.with_shadow else None, tbl,
            nv.ptr(self.mom), nv.ptr(self.epochs), self.rank, self.world, slot, nv.cll(off), nv.cll(n), nv.cf(lr),
            nv.cf(momentum), nv.cf(wd), blocks, 0, tab, n_sh, nv.stream_handle(stream)), "p2p_sync_nvls")

    def allreduce_mean_(self, off: int = 0, n: Optional[int] = None, two_shot: Optional[bool] = None,
                        blocks: int = 0) -> torch.Tensor:
        """Average ``grad[off:off+n]`` over ranks.  With lr = wd = momentum = 0 the fused epilogue
        degenerates to ``mom <- mean(grad)``, so the result is returned as a view of ``mom``
        (two-shot: each rank holds its own slice only; call ``allgather_mom`` for all of it)."""
        n = self.numel - off if n is None else n
        if two_shot is None:
            two_shot = n * 4 >= 256 * 1024
        self.sync(nv.P2P_TWOSHOT if two_shot else nv.P2P_ONESHOT, off, n, blocks=blocks)
        return self.mom[off: off + n]

    def close(self) -> None:
        if self._vmm is not None:        # VMM mappings live until process exit
            self.peer_base = []
            return
        for r, p in enumerate(self.peer_base):
            if r != self.rank:
                self._lib.ddpb_symm_close(ctypes.c_void_p(p))
        self._lib.ddpb_symm_free(ctypes.c_void_p(self.base))
        self.peer_base = []
