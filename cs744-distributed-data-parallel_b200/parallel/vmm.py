"""NVLS set-up: a cuMemCreate-backed symmetric arena bound to an NVSwitch multicast object.

Control-plane half of ``csrc/runtime/vmm_arena.cpp``.  POSIX file
descriptors of the allocations have to reach the other per-GPU processes; they travel as
``SCM_RIGHTS`` messages over abstract unix sockets, with torch.distributed (the existing TCPStore
control plane, ref ``Part 2a/main.py:148-153``) used only to agree on socket names and for barriers.

``VmmArena.create`` returns None (after telling every rank so) whenever anything in the chain is
unavailable - no multicast support, a driver error, fd passing blocked - and the caller falls back to
the cudaMalloc + CUDA-IPC arena and the unicast kernels.
"""
from __future__ import annotations

import ctypes
import os
import socket
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import native as nv
from ..ops.native import check, lib


def _sock_name(tag: str, rank: int) -> str:
    return f"\0ddpb-{tag}-{rank}"


def exchange_fds(my_fds: List[int], tag: str) -> List[List[int]]:
    """Every rank publishes ``my_fds``; returns ``result[r]`` = this process's copies of rank r's
    descriptors (``result[own rank]`` is ``my_fds`` itself)."""
    world, rank = dist.get_world_size(), dist.get_rank()
    server = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    server.bind(_sock_name(tag, rank))
    server.listen(world)
    dist.barrier()                                  # every listener is up
    out: List[List[int]] = [[] for _ in range(world)]
    out[rank] = list(my_fds)
    # deterministic order avoids connect/accept deadlocks: lower ranks are served first
    for peer in range(world):
        if peer == rank:
            for _ in range(world - 1):
                conn, _ = server.accept()
                socket.send_fds(conn, [b"f"], list(my_fds))
                conn.close()
        else:
            c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            c.connect(_sock_name(tag, peer))
            _, fds, _, _ = socket.recv_fds(c, 16, len(my_fds))
            out[peer] = list(fds)
            c.close()
        dist.barrier()
    server.close()
    return out


def _all_ok(ok: bool, device: torch.device) -> bool:
    t = torch.tensor([1 if ok else 0], device=device, dtype=torch.int32)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(int(t))


class VmmArena:
    """One rank's view: local mapping, peer mappings and the multicast mapping of the same bytes."""

    def __init__(self) -> None:
        self.bytes = 0
        self.local = 0
        self.peers: List[int] = []
        self.mc = 0
        self.mem_handle = 0
        self.mc_handle = 0

    @staticmethod
    def supported(device: torch.device) -> bool:
        torch.cuda.set_device(device)
        flag = ctypes.c_int(0)
        check(lib().ddpb_vmm_supported(ctypes.byref(flag)), "vmm_supported")
        return bool(flag.value)

    @classmethod
    def create(cls, nbytes: int, device: torch.device, tag: str = "arena") -> Optional["VmmArena"]:
        L = lib()
        world, rank = dist.get_world_size(), dist.get_rank()
        torch.cuda.set_device(device)
        ok = True
        try:
            ok = cls.supported(device)
        except nv.NativeError:
            ok = False
        if not _all_ok(ok, device):
            return None
        self = cls()
        tag = f"{tag}-{os.environ.get('MASTER_PORT', '0')}"
        try:
            padded = ctypes.c_size_t(0)
            check(L.ddpb_vmm_padded_size(ctypes.c_size_t(nbytes), world, ctypes.byref(padded)), "vmm_padded_size")
            self.bytes = int(padded.value)
            ptr, fd, h = ctypes.c_void_p(), ctypes.c_int(-1), ctypes.c_ulonglong(0)
            check(L.ddpb_vmm_alloc(ctypes.c_size_t(self.bytes), ctypes.byref(ptr), ctypes.byref(fd), ctypes.byref(h)),
                  "vmm_alloc")
            self.local, self.mem_handle = int(ptr.value), int(h.value)
            my_fds = [int(fd.value)]
            mc_fd = ctypes.c_int(-1)
            mch = ctypes.c_ulonglong(0)
            if rank == 0:
                check(L.ddpb_mc_create(ctypes.c_size_t(self.bytes), world, ctypes.byref(mc_fd), ctypes.byref(mch)),
                      "mc_create")
                my_fds.append(int(mc_fd.value))
            else:
                my_fds.append(my_fds[0])            # placeholder so every rank sends two descriptors
        except nv.NativeError as e:
            print(f"[ddpb] rank {rank}: VMM arena unavailable: {e}", flush=True)
            ok = False
        if not _all_ok(ok, device):
            return None
        try:
            table = exchange_fds(my_fds, tag)
            self.peers = []
            for r in range(world):
                if r == rank:
                    self.peers.append(self.local)
                    continue
                p = ctypes.c_void_p()
                check(L.ddpb_vmm_import(table[r][0], ctypes.c_size_t(self.bytes), ctypes.byref(p)), f"vmm_import({r})")
                self.peers.append(int(p.value))
            if rank != 0:
                check(L.ddpb_mc_import(table[0][1], ctypes.byref(mch)), "mc_import")
            self.mc_handle = int(mch.value)
            check(L.ddpb_mc_add_device(ctypes.c_ulonglong(self.mc_handle)), "mc_add_device")
        except (nv.NativeError, OSError) as e:
            print(f"[ddpb] rank {rank}: multicast set-up failed: {e}", flush=True)
            ok = False
        if not _all_ok(ok, device):                 # also the "everyone has joined" barrier
            return None
        try:
            mcp = ctypes.c_void_p()
            check(L.ddpb_mc_bind_and_map(ctypes.c_ulonglong(self.mc_handle), ctypes.c_ulonglong(self.mem_handle),
                                         ctypes.c_size_t(self.bytes), ctypes.byref(mcp)), "mc_bind_and_map")
            self.mc = int(mcp.value)
        except nv.NativeError as e:
            print(f"[ddpb] rank {rank}: multicast bind failed: {e}", flush=True)
            ok = False
        if not _all_ok(ok, device):
            return None
        for fds in table:
            for f in set(fds):
                try:
                    os.close(f)
                except OSError:
                    pass
        return self
