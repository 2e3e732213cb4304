"""Process-group bootstrap (ref ``init_process``, ``Part 2a/main.py:148-153``).

Same contract as the reference: set ``MASTER_ADDR`` / ``MASTER_PORT``, ``env://`` rendezvous on a
TCPStore hosted by rank 0 (port 6585), block until every rank joined, then call
``fn(rank, size, ...)``.  Differences, all additive:

* the backend is chosen from the device (gloo on CPU, NCCL on CUDA) instead of being hard-wired
  to gloo; NCCL is used for bootstrap/metadata and as the baseline path, while the product
  gradient path on GPUs is the peer-to-peer engine (``parallel/p2p.py``);
* a rendezvous timeout and an orderly ``shutdown`` (the reference never tears down, A.3 #10);
* one process per GPU with ``rank == local device index`` (the CLI has no local-rank flag), or
  the ``LOCAL_RANK`` that ``torchrun`` exports when present.
"""
from __future__ import annotations

import datetime
import os
from typing import Callable, Optional

import torch
import torch.distributed as dist


def pick_backend(device: str, backend: str = "auto") -> str:
    if backend not in ("auto", None):
        return backend
    return "nccl" if device.startswith("cuda") else "gloo"


def local_device_index(rank: int) -> int:
    if "LOCAL_RANK" in os.environ:
        return int(os.environ["LOCAL_RANK"])
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    return rank % n if n else 0


def init_process(master: str, port: str, rank: int, size: int, fn: Optional[Callable] = None,
                 *fn_args, backend: str = "gloo", timeout_s: float = 600.0, device: str = "cpu",
                 **fn_kwargs):
    """Initialise the distributed environment, then run ``fn(rank, size, *args)``."""
    os.environ["MASTER_ADDR"] = str(master)
    os.environ["MASTER_PORT"] = str(port)
    kwargs = dict(rank=rank, world_size=size, timeout=datetime.timedelta(seconds=timeout_s))
    if backend == "nccl" and device.startswith("cuda"):
        idx = local_device_index(rank)
        torch.cuda.set_device(idx)
        kwargs["device_id"] = torch.device("cuda", idx)
    if not dist.is_initialized():
        dist.init_process_group(backend, **kwargs)
    if fn is not None:
        return fn(rank, size, *fn_args, **fn_kwargs)
    return None


def shutdown() -> None:
    if dist.is_available() and dist.is_initialized():
        try:
            dist.barrier()
        except Exception:
            pass
        dist.destroy_process_group()
