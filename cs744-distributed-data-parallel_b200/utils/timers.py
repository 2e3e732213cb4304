"""Iteration timers.

The reference brackets regions with ``time.time()`` (ref ``Part 1/main.py:33-43``): forward =
zero_grad + forward; backward = loss + backward + sync + step; average = the whole iteration.
Host clocks around asynchronous CUDA launches measure nothing, so on a GPU the same three
regions are bracketed with CUDA events on the launching stream and read back once per log
window; on CPU the host clock is used as in the reference.  ``max_over_ranks`` reduces a window's
numbers so that what is reported is the slowest rank (SURVEY K-T1).
"""
from __future__ import annotations

import time
from typing import List, Sequence, Tuple

import torch


class HostRegionTimer:
    """time.time() brackets; returns seconds immediately."""
    is_async = False

    def __init__(self) -> None:
        self._t = [0.0, 0.0, 0.0]

    def mark(self, i: int) -> None:
        self._t[i] = time.time()

    def collect(self) -> Tuple[float, float, float]:
        t0, t1, t2 = self._t
        return t1 - t0, t2 - t1, t2 - t0

    def flush(self) -> List[Tuple[float, float, float]]:
        return []


class CudaRegionTimer:
    """Three events per iteration, resolved lazily (no host sync inside the hot loop)."""
    is_async = True

    def __init__(self, stream=None, depth: int = 64) -> None:
        self.stream = stream
        self._pool: List[Tuple[torch.cuda.Event, ...]] = []
        self._pending: List[Tuple[torch.cuda.Event, ...]] = []
        self._cur: List[torch.cuda.Event] = []

    def _events(self):
        if self._pool:
            return list(self._pool.pop())
        return [torch.cuda.Event(enable_timing=True) for _ in range(3)]

    def mark(self, i: int) -> None:
        if i == 0:
            self._cur = self._events()
        self._cur[i].record(self.stream) if self.stream is not None else self._cur[i].record()
        if i == 2:
            self._pending.append(tuple(self._cur))

    def flush(self) -> List[Tuple[float, float, float]]:
        """Synchronise on the last event and return (fwd, bwd, total) seconds per iteration."""
        out = []
        if not self._pending:
            return out
        self._pending[-1][2].synchronize()
        for e0, e1, e2 in self._pending:
            out.append((e0.elapsed_time(e1) * 1e-3, e1.elapsed_time(e2) * 1e-3,
                        e0.elapsed_time(e2) * 1e-3))
            self._pool.append((e0, e1, e2))
        self._pending = []
        return out


def max_over_ranks(values: Sequence[float], device="cpu") -> List[float]:
    """Element-wise max across ranks (identity when no process group is up)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return list(values)
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()
