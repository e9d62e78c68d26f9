"""Symmetric peer memory and the fused all-reduce + optimizer launch.

The process-group half of the engine on the ``b200`` backend: every
rank owns one symmetric segment — flag block + gradient arena (+ parameter
arena for the two-shot push) — allocated by the native library, and maps every
peer's segment through CUDA IPC handles exchanged once over the bootstrap
process group.  After that no host-side collective runs on the gradient path:
each bucket is one kernel (``csrc/comm.cu``) that synchronises with its peers
through epoch flags in this memory.

ref: ``train.py:233`` ``DDP(model)`` (torch Reducer + ProcessGroup) and
``train.py:139`` ``optimizer.step()``.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from .. import _native as nat
from .._native import bindings
from .._native.bindings_comm import FLAG_WORDS, MAX_RANKS, Comm

_DT = {torch.float32: ("<f4", 4), torch.bfloat16: ("<u2", 2), torch.int32: ("<i4", 4),
       torch.uint8: ("|u1", 1)}

# One-shot makes every rank pull the whole bucket from its W-1 peers; it is used while that
# remote traffic, bucket_bytes * (W - 1), stays at or below this many bytes (latency path).
# Larger buckets use two-shot (bandwidth path).  Overridable for the sweep.
ONE_SHOT_MAX_BYTES = int(os.environ.get("B200DDP_ONE_SHOT_MAX_BYTES", 2 * 1024 * 1024))


class _RawCuda:
    """Minimal __cuda_array_interface__ carrier so torch can wrap a raw device pointer."""

    def __init__(self, ptr: int, n: int, typestr: str):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False),
                                         "version": 3, "strides": None}


def _wrap(ptr: int, n: int, dtype: torch.dtype, device: torch.device) -> torch.Tensor:
    typestr, _ = _DT[dtype]
    t = torch.as_tensor(_RawCuda(ptr, n, typestr), device=device)
    return t.view(dtype) if t.dtype != dtype else t


class SymmetricArena:
    """Collective allocator of peer-mapped buffers.  ``alloc`` must be called by every rank
    with the same arguments in the same order."""

    def __init__(self, device: torch.device, rank: int, world: int, group=None):
        if world > MAX_RANKS:
            raise RuntimeError(f"b200 backend supports up to {MAX_RANKS} ranks per NVSwitch domain")
        self.device = torch.device(device)
        self.rank, self.world, self.group = rank, world, group
        self.L = nat.lib()
        self._segments: List[Dict] = []
        self.flags_local, self.flags_peers = self._alloc_raw(FLAG_WORDS * 4)
        self.timeout_ms = int(os.environ.get("B200DDP_PEER_TIMEOUT_MS", "30000"))
        self.two_shot_buckets = set()
        self.vec_elems = 4
        self._comm_cache: Dict = {}
        # NVLS: allocate from multicast-capable memory when the fabric offers it.
        # B200DDP_NVLS=0 disables, =1 requires, default (auto) falls back to CUDA IPC.
        self.nvls_mode = os.environ.get("B200DDP_NVLS", "auto")
        self.nvls = False
        if self.nvls_mode != "0" and world > 1:
            self.nvls = self._probe_multicast()
            if self.nvls_mode == "1" and not self.nvls:
                raise RuntimeError("B200DDP_NVLS=1 but NVSwitch multicast memory is unavailable")

    # ------------------------------------------------------------------ multicast arena
    def _probe_multicast(self) -> bool:
        try:
            t = self._alloc_mc(1024, torch.float32)
            ok = bool(getattr(t, "_b200_mc", 0))
        except Exception as e:  # no fabric / old driver / container without the capability
            ok, self._nvls_error = False, repr(e)
        votes = [None] * self.world
        dist.all_gather_object(votes, ok, group=self.group)
        return all(votes)

    def _alloc_mc(self, n: int, dtype: torch.dtype) -> torch.Tensor:
        """Multicast-capable symmetric allocation.  The VMM allocation, handle exchange and
        multicast binding are torch's CUDASymmetricMemory (bootstrap plumbing); what runs on
        the addresses it returns — multimem.ld_reduce / multimem.st — is ours (comm.cu)."""
        import torch.distributed._symmetric_memory as sm
        t = sm.empty(n, dtype=dtype, device=self.device)
        hdl = sm.rendezvous(t, self.group if self.group is not None else dist.group.WORLD)
        t.zero_()
        t._b200_peers = [int(p) for p in hdl.buffer_ptrs]  # type: ignore[attr-defined]
        t._b200_mc = int(hdl.multicast_ptr or 0)           # type: ignore[attr-defined]
        t._b200_hdl = hdl                                   # keep the mapping alive
        return t

    def _alloc_raw(self, nbytes: int):
        nbytes = (nbytes + 255) // 256 * 256
        ptr = C.c_void_p()
        handle = (C.c_char * 64)()
        nat.check(self.L.b200_symm_alloc(nbytes, C.byref(ptr), handle), "b200_symm_alloc")
        handles: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(handles, bytes(handle), group=self.group)
        peers = []
        for r, h in enumerate(handles):
            if r == self.rank:
                peers.append(ptr.value)
            else:
                p = C.c_void_p()
                buf = (C.c_char * 64).from_buffer_copy(h)
                nat.check(self.L.b200_symm_open(buf, C.byref(p)), f"b200_symm_open(rank {r})")
                peers.append(p.value)
        self._segments.append({"local": ptr.value, "peers": peers, "bytes": nbytes})
        return ptr.value, peers

    def alloc(self, n: int, dtype: torch.dtype) -> torch.Tensor:
        if self.nvls:
            torch.cuda.synchronize(self.device)
            return self._alloc_mc(n, dtype)
        _, esize = _DT[dtype]
        local, peers = self._alloc_raw(n * esize)
        t = _wrap(local, n, dtype, self.device)
        t._b200_peers = peers  # type: ignore[attr-defined]
        return t

    def comm(self, grad: torch.Tensor, push: Optional[torch.Tensor] = None, *,
             master: Optional[torch.Tensor] = None, f32_mask: Optional[torch.Tensor] = None,
             parity_stride: int = 0, no_exit_wait: bool = False) -> Comm:
        """Kernel-side view of the symmetric arenas.  `master` + `f32_mask`: the sharded kernels
        also push the fp32 master of the marked parameters (biases, norm scales) to every rank.
        `parity_stride` / `no_exit_wait`: double-buffered gradients (include/b200ddp.h)."""
        c = Comm()
        c.rank, c.world = self.rank, self.world
        for j in range(self.world):
            c.grad[j] = grad._b200_peers[j]
            c.push[j] = push._b200_peers[j] if push is not None else None
            c.flags[j] = self.flags_peers[j]
        c.grad_mc = getattr(grad, "_b200_mc", 0) or None
        c.push_mc = (getattr(push, "_b200_mc", 0) or None) if push is not None else None
        c.timeout_ms = self.timeout_ms
        c.max_blocks = 0
        if f32_mask is not None and master is not None and hasattr(master, "_b200_peers"):
            for j in range(self.world):
                c.master_peer[j] = master._b200_peers[j]
            c.f32_mask = f32_mask.data_ptr()
        c.parity_stride = parity_stride
        c.no_exit_wait = int(no_exit_wait)
        return c

    def pick_algo(self, engine, b) -> int:
        """0 one-shot, 1 two-shot, 2 NVLS for bucket `b` (identical on every rank)."""
        push = engine.shadow if engine.shadow is not None else engine.master
        can_push = hasattr(push, "_b200_peers")
        esize = engine.grad.element_size()
        algo = 1 if (can_push and engine.allow_two_shot
                     and b.numel * esize * (self.world - 1) > ONE_SHOT_MAX_BYTES) else 0
        if (algo == 1 and self.world > 2 and getattr(engine.grad, "_b200_mc", 0)
                and getattr(push, "_b200_mc", 0)):
            # in-switch reduction + broadcast; measured faster than P2P two-shot at 8 ranks for
            # every size, slower at 2 ranks (no traffic saved, multimem overhead)
            algo = 2
        return algo

    # ------------------------------------------------------------------ engine hooks
    def shard_ranges(self, b) -> List[tuple]:
        """Element ranges [lo, hi) each rank owns in a two-shot bucket (mirrors comm.cu)."""
        w = self.vec_elems
        lo, hi = b.start // w, (b.start + b.numel) // w
        per = ((hi - lo + self.world - 1) // self.world + 15) // 16 * 16
        out = []
        for r in range(self.world):
            s = lo + per * r
            e = min(s + per, hi)
            out.append((w * s, w * max(s, e)))
        return out

    def launch_bucket(self, engine, b, scale: float, *, step_host: int = 0,
                      device_step: bool = False, advance: bool = False, max_blocks: int = 0) -> None:
        """One fused kernel: all-reduce bucket `b` over peer memory, scale, optimizer update."""
        opt = engine.optimizer
        push = engine.shadow if engine.shadow is not None else engine.master
        can_push = hasattr(push, "_b200_peers")
        wire_bf16 = engine.grad.dtype == torch.bfloat16
        algo = self.pick_algo(engine, b)
        self.vec_elems = 8 if wire_bf16 else 4  # one 16-byte vector of the wire dtype
        if algo >= 1:
            self.two_shot_buckets.add(b.index)
        key = (id(engine), can_push)
        c = self._comm_cache.get(key)
        if c is None:
            c = self._comm_cache[key] = self.comm(
                engine.grad, push if can_push else None, master=engine.master,
                f32_mask=engine.f32_mask if engine.shadow is not None else None)
        c.max_blocks = max_blocks
        hp = bindings.make_hparams(opt, scale)
        nat.check(self.L.b200_allreduce_optimizer(
            C.byref(c), algo, C.byref(hp), b.start, b.numel, int(wire_bf16),
            opt.master.data_ptr(), opt.exp_avg.data_ptr(),
            opt.exp_avg_sq.data_ptr() if opt.kind != "sgd" else None,
            opt.shadow.data_ptr() if opt.shadow is not None else None,
            int(push.dtype == torch.bfloat16), opt.step_dev.data_ptr() if device_step else None,
            step_host, int(advance), nat.stream_ptr(self.device)), "b200_allreduce_optimizer")

    def allreduce_optimizer_step(self, engine, scale: float, *, device_step: bool = False) -> None:
        """All buckets back to back, in bucket order (identical on every rank)."""
        nb = len(engine.layout.buckets)
        for k, b in enumerate(engine.layout.buckets):
            self.launch_bucket(engine, b, scale, step_host=engine.optimizer.step_count,
                               device_step=device_step, advance=device_step and k == nb - 1)

    def allreduce_(self, t: torch.Tensor, start: int, n: int, *, algo: int, scale: float = 1.0,
                   out: Optional[torch.Tensor] = None) -> None:
        c = self.comm(t)
        nat.check(self.L.b200_allreduce(
            C.byref(c), algo, start, n, int(t.dtype == torch.bfloat16),
            out.data_ptr() if out is not None else None, scale, nat.stream_ptr(self.device)),
            "b200_allreduce")
