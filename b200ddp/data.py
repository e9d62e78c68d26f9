"""Synthetic dataset, shard sampler and the batch pipeline.

Capabilities mirrored from the reference:

* ``SyntheticDataset`` — reference ``train.py:53-67``: ``randn(N, 784)`` fp32
  features and ``randint(0, 10, (N,))`` int64 labels, generated per process and
  unseeded unless ``seed`` is given.
* ``ShardSampler`` — index-for-index parity with
  ``torch.utils.data.DistributedSampler`` as configured at reference
  ``train.py:104-106`` (shuffle, seed 0, pad-by-wrapping, strided rank slice).
* ``BatchLoader`` — replaces ``DataLoader(num_workers=2, pin_memory=...)``
  (reference ``train.py:108-114``).  The data set is one resident tensor, so a
  batch is a single gather, not ``batch_size`` ``__getitem__`` calls pickled
  through worker pipes.  On CUDA the gather lands in a pinned ring and is
  copied on a dedicated copy stream one batch ahead of compute.
"""
from __future__ import annotations

import math
from typing import Iterator, Optional, Tuple

import torch
from torch.utils.data import Dataset


class SyntheticDataset(Dataset):
    def __init__(
        self,
        num_samples: int = 10000,
        input_size: int = 784,
        num_classes: int = 10,
        seed: Optional[int] = None,
        device: Optional[torch.device] = None,
    ):
        self.num_samples = num_samples
        self.input_size = input_size
        self.num_classes = num_classes
        gen = None
        if seed is not None:
            gen = torch.Generator().manual_seed(seed)
        self.data = torch.randn(num_samples, input_size, generator=gen)
        self.labels = torch.randint(0, num_classes, (num_samples,), generator=gen)
        if device is not None and torch.device(device).type != "cpu":
            # --synthetic-on-device: the whole set lives in HBM (31 MB at the
            # default size); batches become on-device gathers.
            self.data = self.data.to(device)
            self.labels = self.labels.to(device)

    def __len__(self) -> int:
        return self.num_samples

    def __getitem__(self, idx):
        return self.data[idx], self.labels[idx]


class ShardSampler:
    """Deterministic rank shard of a shuffled index space.

    Same arithmetic as the sampler the reference instantiates
    (``train.py:104-106``): ``num_samples = ceil(N / R)``, a ``randperm`` seeded
    with ``seed + epoch``, padding by wrapping the head of the permutation, and
    ``indices[rank::R]``.
    """

    def __init__(
        self,
        dataset_len: int,
        num_replicas: int,
        rank: int,
        shuffle: bool = True,
        seed: int = 0,
        drop_last: bool = False,
    ):
        if not (0 <= rank < num_replicas):
            raise ValueError(f"rank {rank} outside [0, {num_replicas})")
        self.n = int(dataset_len)
        self.num_replicas = num_replicas
        self.rank = rank
        self.shuffle = shuffle
        self.seed = seed
        self.drop_last = drop_last
        self.epoch = 0
        if drop_last and self.n % num_replicas != 0:
            self.num_samples = math.ceil((self.n - num_replicas) / num_replicas)
        else:
            self.num_samples = math.ceil(self.n / num_replicas)
        self.total_size = self.num_samples * num_replicas

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch

    def indices(self) -> torch.Tensor:
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            idx = torch.randperm(self.n, generator=g)
        else:
            idx = torch.arange(self.n)
        if not self.drop_last:
            pad = self.total_size - idx.numel()
            if pad > 0:
                if pad <= idx.numel():
                    idx = torch.cat([idx, idx[:pad]])
                else:
                    reps = math.ceil(pad / idx.numel())
                    idx = torch.cat([idx, idx.repeat(reps)[:pad]])
        else:
            idx = idx[: self.total_size]
        return idx[self.rank : self.total_size : self.num_replicas].contiguous()

    def __iter__(self) -> Iterator[int]:
        return iter(self.indices().tolist())

    def __len__(self) -> int:
        return self.num_samples


class BatchLoader:
    """Batches a tensor-backed dataset through a sampler.

    ``len(loader)`` and the short last batch follow ``DataLoader`` with
    ``drop_last=False``.  Iteration yields ``(data, target)`` already on
    ``device``.
    """

    def __init__(
        self,
        dataset: SyntheticDataset,
        batch_size: int,
        sampler: ShardSampler,
        device: torch.device = torch.device("cpu"),
        prefetch: int = 2,
    ):
        self.dataset = dataset
        self.batch_size = batch_size
        self.sampler = sampler
        self.device = torch.device(device)
        self.prefetch = max(1, prefetch)
        self._cuda = self.device.type == "cuda"
        self._resident = self._cuda and dataset.data.device == self.device
        self._ring = None
        self._copy_stream = None

    def __len__(self) -> int:
        return math.ceil(len(self.sampler) / self.batch_size)

    # -- host staging ------------------------------------------------------
    def _make_ring(self):
        d = self.dataset
        slots = self.prefetch + 1
        self._ring = [
            (
                torch.empty(self.batch_size, *d.data.shape[1:], dtype=d.data.dtype).pin_memory(),
                torch.empty(self.batch_size, *d.labels.shape[1:], dtype=d.labels.dtype).pin_memory(),
                torch.empty(self.batch_size, *d.data.shape[1:], dtype=d.data.dtype, device=self.device),
                torch.empty(self.batch_size, *d.labels.shape[1:], dtype=d.labels.dtype, device=self.device),
                torch.cuda.Event(),  # copy done
                torch.cuda.Event(),  # consumer done with the device buffers
            )
            for _ in range(slots)
        ]
        self._copy_stream = torch.cuda.Stream(device=self.device)

    def _stage(self, slot: int, idx: torch.Tensor):
        hx, hy, dx, dy, copied, consumed = self._ring[slot]
        n = idx.numel()
        # The previous H2D out of this pinned slot must have drained before the
        # host overwrites it (the consumer only waited on it in stream order).
        copied.synchronize()
        torch.index_select(self.dataset.data, 0, idx, out=hx[:n])
        torch.index_select(self.dataset.labels, 0, idx, out=hy[:n])
        with torch.cuda.stream(self._copy_stream):
            self._copy_stream.wait_event(consumed)
            dx[:n].copy_(hx[:n], non_blocking=True)
            dy[:n].copy_(hy[:n], non_blocking=True)
            copied.record(self._copy_stream)
        return n

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        idx_all = self.sampler.indices()
        chunks = list(torch.split(idx_all, self.batch_size))
        if not self._cuda:
            for idx in chunks:
                yield self.dataset.data[idx], self.dataset.labels[idx]
            return
        if self._resident:
            idx_dev = idx_all.to(self.device, non_blocking=True)
            for idx in torch.split(idx_dev, self.batch_size):
                yield self.dataset.data.index_select(0, idx), self.dataset.labels.index_select(0, idx)
            return
        if self._ring is None:
            self._make_ring()
        slots = len(self._ring)
        cur = torch.cuda.current_stream(self.device)
        for s in range(slots):
            self._ring[s][5].record(cur)
        sizes = {}
        nxt = 0
        for _ in range(min(self.prefetch, len(chunks))):
            sizes[nxt] = self._stage(nxt % slots, chunks[nxt])
            nxt += 1
        for i in range(len(chunks)):
            slot = i % slots
            _, _, dx, dy, copied, consumed = self._ring[slot]
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(copied)
            n = sizes.pop(i)
            yield dx[:n], dy[:n]
            consumed.record(torch.cuda.current_stream(self.device))
            if nxt < len(chunks):
                sizes[nxt] = self._stage(nxt % slots, chunks[nxt])
                nxt += 1


    # -- host batches for the native end-to-end runner ------------------------
    def host_batches(self, slots: int = 4) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        """Yields each batch as pinned host tensors (views of a small ring) for
        ``HostStepper.step``, which does the H2D itself.  A slot is reused `slots` batches
        later; the stepper returns from step i+1 only after step i has finished, so with
        slots >= 3 the asynchronous H2D out of a slot has always drained by then."""
        if not self._cuda or self._resident:
            raise RuntimeError("host_batches needs a CUDA target device and a host-resident data set")
        d = self.dataset
        # pinned once and kept with the loader: the runner's asynchronous H2D of the last batches
        # of an epoch may still be reading a slot when the generator is exhausted
        ring = getattr(self, "_host_ring", None)
        if ring is None or len(ring) < max(3, slots):
            ring = self._host_ring = [
                (torch.empty(self.batch_size, *d.data.shape[1:], dtype=d.data.dtype).pin_memory(),
                 torch.empty(self.batch_size, *d.labels.shape[1:], dtype=d.labels.dtype).pin_memory())
                for _ in range(max(3, slots))]
        for i, idx in enumerate(torch.split(self.sampler.indices(), self.batch_size)):
            hx, hy = ring[i % len(ring)]
            n = idx.numel()
            torch.index_select(d.data, 0, idx, out=hx[:n])
            torch.index_select(d.labels, 0, idx, out=hy[:n])
            yield hx[:n], hy[:n]


    def host_epoch(self, max_bytes: int = 1 << 30):
        """This epoch's shard as ONE pinned pool for ``HostStepper.run_pool``: ``(x [nb, B, ...],
        y [nb, B], tail)`` with the nb full batches in sampler order and ``tail`` the short last
        batch ``(hx, hy)`` or None.  The pinned buffers belong to the loader and are
        refilled every epoch; returns None when the shard is larger than `max_bytes`."""
        if not self._cuda or self._resident:
            raise RuntimeError("host_epoch needs a CUDA target device and a host-resident data set")
        d = self.dataset
        idx = self.sampler.indices()
        B = self.batch_size
        nb, rem = idx.numel() // B, idx.numel() % B
        per = d.data[0].numel() * d.data.element_size() + d.labels[0].numel() * d.labels.element_size()
        if nb == 0 or idx.numel() * per > max_bytes:
            return None
        pool = getattr(self, "_host_pool", None)
        if pool is None or pool[0].shape[0] < nb + 1:
            pool = self._host_pool = (
                torch.empty(nb + 1, B, *d.data.shape[1:], dtype=d.data.dtype).pin_memory(),
                torch.empty(nb + 1, B, *d.labels.shape[1:], dtype=d.labels.dtype).pin_memory())
        xp, yp = pool
        torch.index_select(d.data, 0, idx[:nb * B], out=xp[:nb].view(nb * B, *d.data.shape[1:]))
        torch.index_select(d.labels, 0, idx[:nb * B], out=yp[:nb].view(nb * B, *d.labels.shape[1:]))
        tail = None
        if rem:
            torch.index_select(d.data, 0, idx[nb * B:], out=xp[nb, :rem])
            torch.index_select(d.labels, 0, idx[nb * B:], out=yp[nb, :rem])
            tail = (xp[nb, :rem], yp[nb, :rem])
        return xp[:nb], yp[:nb], tail


def create_data_loader(
    dataset: SyntheticDataset,
    batch_size: int,
    rank: int,
    world_size: int,
    device: torch.device = torch.device("cpu"),
    seed: int = 0,
) -> Tuple[BatchLoader, ShardSampler]:
    """Same return contract as reference ``train.py:101-116``: ``(loader, sampler)``."""
    sampler = ShardSampler(len(dataset), num_replicas=world_size, rank=rank, shuffle=True, seed=seed)
    loader = BatchLoader(dataset, batch_size, sampler, device=device)
    return loader, sampler
