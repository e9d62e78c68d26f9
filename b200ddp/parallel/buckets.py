"""Gradient bucket planner and flat-arena layout.

Replaces the planning half of torch DDP's C++ ``Reducer`` (the reference gets it
implicitly from ``DDP(model)``, reference ``train.py:233``).  Policy reproduced
(SURVEY §2.4): parameters are visited in reverse registration order — a proxy
for gradient-ready order — with a 1 MiB cap on the first bucket and 25 MiB on
the rest; a bucket closes once its byte size reaches the cap, so buckets
overshoot.  ``plan_buckets`` is tested index-for-index against
``torch.distributed._compute_bucket_assignment_by_size``.

What is different from the Reducer: a bucket here is not a staging buffer that
gradients are copied into and out of.  ``ArenaLayout`` places every parameter at
a fixed, 256-byte-aligned offset of one flat arena in bucket order, so a bucket
is a contiguous slice, ``param.grad`` is a view of it, and the fused
all-reduce+optimizer kernel consumes it in place.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Sequence, Tuple

import torch

MiB = 1024 * 1024
DEFAULT_FIRST_BUCKET_BYTES = 1 * MiB
DEFAULT_BUCKET_CAP_BYTES = 25 * MiB
# Elements.  64 fp32 = 256 B, 64 bf16 = 128 B: keeps every tensor start legal for
# TMA (16 B) and for 16-byte vector loads in every arena dtype.
ARENA_ALIGN_ELEMS = 64


def plan_buckets(
    sizes_bytes: Sequence[int],
    keys: Sequence[Tuple] | None = None,
    caps: Sequence[int] = (DEFAULT_FIRST_BUCKET_BYTES, DEFAULT_BUCKET_CAP_BYTES),
) -> List[List[int]]:
    """Greedy size-capped bucketing over tensors *in the given order*.

    ``sizes_bytes[i]`` is tensor i's byte size, ``keys[i]`` its (dtype, device)
    class (tensors of different classes never share a bucket).  The first bucket
    of a class uses ``caps[0]``, later ones ``caps[1]`` (and so on, sticking at
    the last cap).  Result buckets are ordered by their smallest index.
    """
    if keys is None:
        keys = [()] * len(sizes_bytes)
    open_buckets: Dict[Tuple, Tuple[List[int], int]] = {}
    cap_pos: Dict[Tuple, int] = {}
    done: List[List[int]] = []
    for i, (nbytes, key) in enumerate(zip(sizes_bytes, keys)):
        idxs, size = open_buckets.get(key, ([], 0))
        idxs = idxs + [i]
        size += int(nbytes)
        pos = cap_pos.setdefault(key, 0)
        if size >= caps[pos]:
            done.append(idxs)
            open_buckets[key] = ([], 0)
            if pos + 1 < len(caps):
                cap_pos[key] = pos + 1
        else:
            open_buckets[key] = (idxs, size)
    for idxs, _ in open_buckets.values():
        if idxs:
            done.append(idxs)
    done.sort(key=lambda b: min(b))
    return done


@dataclass
class ParamSlot:
    name: str
    index: int  # position in module.parameters() order (optimizer index)
    shape: torch.Size
    numel: int
    offset: int  # element offset in the flat arena
    bucket: int


@dataclass
class Bucket:
    index: int
    start: int  # element offset (aligned)
    end: int  # element offset one past the padded end
    params: List[int] = field(default_factory=list)  # optimizer indices, ready order

    @property
    def numel(self) -> int:
        return self.end - self.start


@dataclass
class ArenaLayout:
    slots: List[ParamSlot]  # indexed by optimizer index
    buckets: List[Bucket]
    total: int  # padded element count

    def slot_by_name(self, name: str) -> ParamSlot:
        for s in self.slots:
            if s.name == name:
                return s
        raise KeyError(name)


def _align(n: int, a: int = ARENA_ALIGN_ELEMS) -> int:
    return (n + a - 1) // a * a


def build_layout(
    named_params: Sequence[Tuple[str, torch.Tensor]],
    elem_bytes: int = 4,
    caps: Sequence[int] = (DEFAULT_FIRST_BUCKET_BYTES, DEFAULT_BUCKET_CAP_BYTES),
    ready_order: Sequence[int] | None = None,
) -> ArenaLayout:
    """Lay parameters out in bucket order.

    ``ready_order`` lists optimizer indices in the order their gradients become
    ready; default is reverse registration order, which is also what DDP assumes
    before it has observed a backward pass.  ``elem_bytes`` is the size of one
    *gradient* element (4 for fp32 wire, 2 for bf16 wire) and only affects where
    bucket boundaries fall.
    """
    n = len(named_params)
    order = list(ready_order) if ready_order is not None else list(range(n - 1, -1, -1))
    assert sorted(order) == list(range(n)), "ready_order must be a permutation"
    sizes = [named_params[i][1].numel() * elem_bytes for i in order]
    groups = plan_buckets(sizes, None, caps)
    slots: List[ParamSlot | None] = [None] * n
    buckets: List[Bucket] = []
    cursor = 0
    for b, grp in enumerate(groups):
        start = cursor
        members = []
        for pos in grp:
            i = order[pos]
            name, p = named_params[i]
            slots[i] = ParamSlot(name, i, p.shape, p.numel(), cursor, b)
            cursor = _align(cursor + p.numel())
            members.append(i)
        buckets.append(Bucket(b, start, cursor, members))
    return ArenaLayout([s for s in slots if s is not None], buckets, cursor)
