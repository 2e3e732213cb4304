"""Gradient bucket planning.

The reference gets its buckets from torch DDP (ref ``Part 3/main.py:61``): 25 MiB cap, 1 MiB
first bucket, tensors visited in reverse declaration order and, after the first iteration, in
the observed gradient-ready order (SURVEY 2.5.4, C23 b/d).  Because this framework owns the
backward pass, the ready order is static and known before the first step, so the plan is
computed once, up front, and never rebuilt.

Plans available (``--bucket-plan``):

* ``ddp``       – the size-capped greedy assignment torch uses (same result as
  ``dist._compute_bucket_assignment_by_size`` on the same order; verified in the tests);
* ``per_layer`` – one bucket per conv block (BN pair + conv pair), classifier folded into the
  first: launches a reduction as soon as each layer's wgrad lands; best overlap on NVSwitch
  where per-collective cost is latency, not link count;
* ``single``    – everything in one bucket (what DDP does in its very first iteration).

The C++ scheduler (``csrc/runtime/engine.cpp``: ``plan_buckets`` / ``buckets_ready`` / ``launch_bucket``) computes
the same plan natively and consumes the plan as arrays of
``(offset, numel)`` ranges over the flat gradient arena, which is laid out in ready order so
that every bucket is one contiguous range.
"""
from __future__ import annotations

import dataclasses
from typing import Dict, List, Sequence, Tuple

MIB = 1024 * 1024
ARENA_ALIGN_ELEMS = 64      # every tensor starts on a 256-byte boundary in the fp32 arena


@dataclasses.dataclass(frozen=True)
class Bucket:
    index: int
    names: Tuple[str, ...]
    nbytes: int


def assign_by_size(sizes_bytes: Sequence[int], first_cap: int, cap: int) -> List[List[int]]:
    """Greedy size-capped grouping: a bucket closes once its size reaches its cap; the first
    bucket uses ``first_cap`` and all later ones ``cap``."""
    groups: List[List[int]] = []
    cur: List[int] = []
    cur_bytes = 0
    limit = first_cap
    for i, nb in enumerate(sizes_bytes):
        cur.append(i)
        cur_bytes += nb
        if cur_bytes >= limit:
            groups.append(cur)
            cur, cur_bytes = [], 0
            limit = cap
    if cur:
        groups.append(cur)
    return groups


def plan_buckets(names: Sequence[str], numels: Sequence[int], plan: str = "ddp",
                 first_cap_mb: float = 1.0, cap_mb: float = 25.0, elem_size: int = 4,
                 layer_of: Dict[str, int] | None = None) -> List[Bucket]:
    """``names`` must already be in gradient-ready order."""
    sizes = [n * elem_size for n in numels]
    if plan == "ddp":
        groups = assign_by_size(sizes, int(first_cap_mb * MIB), int(cap_mb * MIB))
    elif plan == "single":
        groups = [list(range(len(names)))]
    elif plan == "per_layer":
        if layer_of is None:
            raise ValueError("per_layer plan needs layer_of")
        groups, seen = [], {}
        for i, nm in enumerate(names):
            key = layer_of[nm]
            if key not in seen:
                seen[key] = len(groups)
                groups.append([])
            groups[seen[key]].append(i)
    else:
        raise ValueError(f"unknown bucket plan {plan!r}")
    return [Bucket(b, tuple(names[i] for i in g), sum(sizes[i] for i in g))
            for b, g in enumerate(groups)]


@dataclasses.dataclass(frozen=True)
class ArenaSlot:
    name: str
    offset: int      # in elements
    numel: int


@dataclasses.dataclass(frozen=True)
class ArenaLayout:
    """Flat layout of all parameters in gradient-ready order, each slot 256-byte aligned."""
    slots: Tuple[ArenaSlot, ...]
    total: int       # padded element count

    def slot(self, name: str) -> ArenaSlot:
        for s in self.slots:
            if s.name == name:
                return s
        raise KeyError(name)

    def bucket_ranges(self, buckets: Sequence[Bucket]) -> List[Tuple[int, int]]:
        """(offset, numel) of each bucket, padding between member tensors included."""
        out = []
        for b in buckets:
            first, last = self.slot(b.names[0]), self.slot(b.names[-1])
            end = _round_up(last.offset + last.numel, ARENA_ALIGN_ELEMS)
            out.append((first.offset, end - first.offset))
        return out


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def arena_layout(names: Sequence[str], numels: Sequence[int]) -> ArenaLayout:
    slots, off = [], 0
    for nm, n in zip(names, numels):
        slots.append(ArenaSlot(nm, off, n))
        off = _round_up(off + n, ARENA_ALIGN_ELEMS)
    return ArenaLayout(tuple(slots), off)
