"""Process-group runtime: env:// rendezvous, device selection, teardown.

Mirrors reference ``train.py:70-98`` (``setup_distributed``, ``cleanup_distributed``,
``get_device``) including the log lines, with one correction the reference
needs on a GPU box: the host-side plumbing backend is chosen per device instead
of hard-coding ``gloo`` (reference ``train.py:71``).

Backends (``--backend``):

* ``b200``  — product path on a B200 box.  torch.distributed (NCCL) is used only
  for bootstrap (exchanging IPC handles, barriers at epoch boundaries); the
  gradient all-reduce is our fused peer-memory kernel (``b200ddp/parallel/symm.py``).
* ``nccl``  — baseline: one NCCL all-reduce per bucket, then our fused optimizer
  kernel.  Kept selectable as the comparator and as the multi-node fallback.
* ``gloo``  — CPU plumbing configuration (BASELINE.json config 1).
* ``auto``  — ``b200`` when CUDA is present, else ``gloo``.
"""
from __future__ import annotations

import datetime
import os
from typing import Tuple

import torch
import torch.distributed as dist

from ..utils.logging import get_logger

logger = get_logger("b200ddp")

REQUIRED_ENV = ("RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT")


def resolve_backend(backend: str) -> str:
    if backend == "auto":
        return "b200" if torch.cuda.is_available() else "gloo"
    if backend not in ("b200", "nccl", "gloo"):
        raise ValueError(f"unknown backend {backend!r}")
    if backend in ("b200", "nccl") and not torch.cuda.is_available():
        raise RuntimeError(f"--backend {backend} needs a CUDA device; none is visible")
    return backend


def pick_engine_backend(backend: str, device) -> str:
    """Gradient-path backend of the engine for ``--backend auto`` (collective: call on every rank
    after ``setup_distributed``).  The ``b200`` backend maps every peer's arenas through CUDA IPC /
    NVSwitch multicast, which exists inside ONE NVSwitch domain of at most 8 GPUs; the reference's
    own deployment (``entrypoint.sh``: REPLICAS > 1 nodes, ref ``entrypoint.sh:33-39``) spans hosts.
    There, or beyond 8 ranks, the engine falls back to NCCL all-reduce + the fused optimizer kernel."""
    if backend != "auto":
        return backend
    if torch.device(device).type != "cuda":
        return "gloo"
    world_size = dist.get_world_size() if dist.is_initialized() else 1
    if world_size == 1:
        return "b200"
    import socket
    from .._native.bindings_comm import MAX_RANKS
    hosts = [None] * world_size
    dist.all_gather_object(hosts, socket.gethostname())
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", world_size))
    one_box = len(set(hosts)) == 1 and local_world == world_size
    if one_box and world_size <= MAX_RANKS:
        return "b200"
    why = (f"{len(set(hosts))} hosts" if not one_box else f"{world_size} ranks > {MAX_RANKS}")
    get_logger("b200ddp").info(f"backend auto -> nccl ({why}: peer-mapped arenas need one NVSwitch domain)")
    return "nccl"


def setup_distributed(backend: str = "auto", timeout_s: int = 600) -> Tuple[int, int, int]:
    """Joins the env:// rendezvous.  Returns ``(rank, world_size, local_rank)``.

    A single process started without torchrun (no ``RANK`` in the environment)
    becomes a world of one on 127.0.0.1, so ``python -m b200ddp.train`` works
    without a launcher.
    """
    backend = resolve_backend(backend)
    if "RANK" not in os.environ:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        os.environ["RANK"] = "0"
        os.environ["WORLD_SIZE"] = "1"
        os.environ.setdefault("LOCAL_RANK", "0")
    missing = [k for k in REQUIRED_ENV if k not in os.environ]
    if missing:
        raise RuntimeError(f"env:// rendezvous needs {', '.join(missing)} (set by the launcher)")
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    c10d_backend = "gloo" if backend == "gloo" else "nccl"
    kwargs = {}
    if c10d_backend == "nccl":
        torch.cuda.set_device(local_rank)
        kwargs["device_id"] = torch.device("cuda", local_rank)
    dist.init_process_group(backend=c10d_backend, timeout=datetime.timedelta(seconds=timeout_s), **kwargs)
    rank = dist.get_rank()
    world_size = dist.get_world_size()
    logger.info(
        f"Initialized process group: rank={rank}, world_size={world_size}, "
        f"local_rank={local_rank}"
    )
    return rank, world_size, local_rank


def cleanup_distributed() -> None:
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()


def get_device(local_rank: int) -> torch.device:
    if torch.cuda.is_available():
        device = torch.device(f"cuda:{local_rank}")
        torch.cuda.set_device(device)
        logger.info(f"Using GPU: {torch.cuda.get_device_name(device)}")
    else:
        device = torch.device("cpu")
        logger.info("Using CPU for training")
    return device


def world() -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1
