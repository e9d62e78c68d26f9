"""ShardSampler vs torch DistributedSampler (SURVEY §2.3 D6, §4 unit tier)."""
import torch
from hypothesis import given, settings, strategies as st
from torch.utils.data.distributed import DistributedSampler

from b200ddp.data import BatchLoader, ShardSampler, SyntheticDataset, create_data_loader


def test_golden_vectors_n10_r3():
    # survey run, N=10 R=3 epoch 0 seed 0
    want = {0: [4, 5, 0, 2], 1: [1, 3, 8, 4], 2: [7, 9, 6, 1]}
    for r, idx in want.items():
        assert ShardSampler(10, 3, r).indices().tolist() == idx


@settings(max_examples=60, deadline=None)
@given(n=st.integers(1, 300), world=st.integers(1, 9), epoch=st.integers(0, 5),
       seed=st.integers(0, 3), shuffle=st.booleans(), drop_last=st.booleans())
def test_matches_distributed_sampler(n, world, epoch, seed, shuffle, drop_last):
    ds = list(range(n))
    for rank in range(world):
        ref = DistributedSampler(ds, num_replicas=world, rank=rank, shuffle=shuffle, seed=seed,
                                 drop_last=drop_last)
        ref.set_epoch(epoch)
        ours = ShardSampler(n, world, rank, shuffle=shuffle, seed=seed, drop_last=drop_last)
        ours.set_epoch(epoch)
        assert list(ours) == list(ref)
        assert len(ours) == len(ref)


def test_loader_batches_and_short_tail():
    # N=10000 R=8 B=64 -> 1250 per rank -> 20 batches, last 34
    ds = SyntheticDataset(10000, seed=0)
    loader, sampler = create_data_loader(ds, 64, rank=3, world_size=8)
    assert len(loader) == 20
    sizes = [x.shape[0] for x, _ in loader]
    assert sizes == [64] * 19 + [34]
    idx = sampler.indices()
    x0, y0 = next(iter(loader))
    assert torch.equal(x0, ds.data[idx[:64]]) and torch.equal(y0, ds.labels[idx[:64]])


def test_dataset_shapes_and_dtypes():
    ds = SyntheticDataset(17)
    x, y = ds[3]
    assert x.shape == (784,) and x.dtype == torch.float32
    assert y.dtype == torch.int64 and 0 <= int(y) < 10
    assert len(ds) == 17


def test_loader_handles_sequence_labels():
    """Token datasets have [N, T] labels; batches keep that shape (CPU path)."""
    from b200ddp.models.datasets import SyntheticTokens
    ds = SyntheticTokens(20, seq_len=16, vocab_size=50, seed=0)
    loader, _ = create_data_loader(ds, 8, rank=0, world_size=1)
    shapes = [(x.shape, y.shape) for x, y in loader]
    assert shapes[0] == (torch.Size([8, 16]), torch.Size([8, 16])) and shapes[-1][0][0] == 4
    assert torch.equal(ds.data[:, 1:], ds.labels[:, :-1])  # labels are the inputs shifted by one


def test_host_batches_cover_the_shard_in_order_and_reuse_the_ring():
    """BatchLoader.host_batches (pinned batches for the native runner), with pinning stubbed out."""
    from unittest import mock

    import pytest

    from b200ddp.data import BatchLoader, ShardSampler, SyntheticDataset
    ds = SyntheticDataset(200, seed=1)
    sm = ShardSampler(len(ds), 2, 0)
    ld = BatchLoader(ds, 64, sm)
    with pytest.raises(RuntimeError, match="CUDA"):
        next(ld.host_batches())
    ld._cuda, ld._resident = True, False          # pretend the target is a GPU
    with mock.patch.object(torch.Tensor, "pin_memory", lambda self: self):
        got = [(x.clone(), y.clone()) for x, y in ld.host_batches()]
        ring = ld._host_ring
        assert len(ring) >= 3
        list(ld.host_batches())
        assert ld._host_ring is ring             # pinned once, reused across epochs
    idx = sm.indices()
    assert [x.shape[0] for x, _ in got] == [64, 36]
    assert torch.equal(torch.cat([x for x, _ in got]), ds.data[idx])
    assert torch.equal(torch.cat([y for _, y in got]), ds.labels[idx])


def test_cpu_engine_does_not_take_the_host_stepper_path():
    from b200ddp import DataParallel, SimpleNet
    from b200ddp.data import BatchLoader, ShardSampler, SyntheticDataset
    from b200ddp.train import _use_host_stepper
    ds = SyntheticDataset(64, seed=2)
    ld = BatchLoader(ds, 32, ShardSampler(len(ds), 1, 0))
    eng = DataParallel(SimpleNet(), backend="gloo", init_sync=False)
    assert not _use_host_stepper(eng, ld)
