"""Bucket planner vs torch's Reducer planner (SURVEY §2.4, §4 unit tier)."""
import torch
import torch.distributed as dist
from hypothesis import given, settings, strategies as st

from b200ddp.models.simplenet import SimpleNet
from b200ddp.parallel.buckets import ARENA_ALIGN_ELEMS, MiB, build_layout, plan_buckets


def _torch_plan(tensors, caps):
    idx, _ = dist._compute_bucket_assignment_by_size(tensors, list(caps))
    return [list(b) for b in idx]


def test_simplenet_single_bucket():
    params = list(SimpleNet().parameters())[::-1]
    sizes = [p.numel() * 4 for p in params]
    ours = plan_buckets(sizes)
    assert ours == _torch_plan(params, (MiB, 25 * MiB))
    assert len(ours) == 1 and sum(sizes) == 1_077_288


@settings(max_examples=80, deadline=None)
@given(numels=st.lists(st.integers(1, 3000), min_size=1, max_size=40),
       cap0=st.integers(64, 8000), cap1=st.integers(64, 20000))
def test_matches_torch_planner(numels, cap0, cap1):
    tensors = [torch.empty(n) for n in numels]
    ours = plan_buckets([n * 4 for n in numels], None, (cap0, cap1))
    assert ours == _torch_plan(tensors, (cap0, cap1))


def test_gpt2_small_bucket_table():
    # 13 fp32 buckets for GPT-2 small with tied wte, last one holds wte.
    shapes = [(50257, 768), (1024, 768)]
    for _ in range(12):
        shapes += [(768,), (768,), (2304, 768), (2304,), (768, 768), (768,), (768,), (768,),
                   (3072, 768), (3072,), (768, 3072), (768,)]
    shapes += [(768,), (768,)]
    numels = [int(torch.Size(s).numel()) for s in shapes][::-1]
    plan = plan_buckets([n * 4 for n in numels])
    assert len(plan) == 13
    sizes_mib = [sum(numels[i] for i in b) * 4 / MiB for b in plan]
    assert abs(sizes_mib[0] - 9.0) < 0.1 and abs(sizes_mib[-1] - 168.3) < 0.5
    assert sum(numels) == 124_439_808


def test_layout_is_bucket_ordered_and_aligned():
    named = list(SimpleNet().named_parameters())
    lay = build_layout(named, caps=(100_000, 300_000))
    assert len(lay.buckets) >= 2
    seen = 0
    for b in lay.buckets:
        assert b.start == seen and b.start % ARENA_ALIGN_ELEMS == 0
        for i in b.params:
            s = lay.slots[i]
            assert b.start <= s.offset and s.offset + s.numel <= b.end
            assert s.offset % ARENA_ALIGN_ELEMS == 0
        seen = b.end
    assert seen == lay.total
    # reverse registration order: the last layer's bias is first in the arena
    assert lay.slots[len(named) - 1].offset == 0


@settings(max_examples=200, deadline=None)
@given(world=st.sampled_from([2, 4, 8]), vec=st.sampled_from([4, 8]), start_v=st.integers(0, 4096),
       numel_v=st.integers(1, 200_000))
def test_two_shot_shard_ranges_partition_the_bucket(world, vec, start_v, numel_v):
    """Host mirror of comm.cu's shard ownership (who reduces, updates and later re-broadcasts which
    slice; `consolidate_state` relies on it): contiguous, ordered, disjoint, covering, and every
    shard but the last a multiple of 16 vectors."""
    from types import SimpleNamespace

    from b200ddp.parallel.symm import SymmetricArena
    arena = object.__new__(SymmetricArena)
    arena.world, arena.vec_elems = world, vec
    b = SimpleNamespace(start=start_v * vec, numel=numel_v * vec)
    ranges = arena.shard_ranges(b)
    assert len(ranges) == world
    assert ranges[0][0] == b.start
    cursor = b.start
    for lo, hi in ranges:
        assert lo <= hi
        if hi > lo:
            assert lo == cursor
            cursor = hi
    assert cursor == b.start + b.numel
    sizes = [hi - lo for lo, hi in ranges if hi > lo]
    for sz in sizes[:-1]:
        assert sz % (16 * vec) == 0
