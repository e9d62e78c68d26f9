"""DataParallel engine — the replacement for ``DDP(model)`` + ``optim.Adam``.

The reference wraps its model in ``torch.nn.parallel.DistributedDataParallel``
(reference ``train.py:233``) and steps ``optim.Adam`` (``train.py:249,139``).  Per
step that costs: 6 grad->bucket copies, one all-reduce, a divide, 6 bucket->grad
copies and ~50 small optimizer launches.  This engine owns
both jobs and lays memory out so none of those copies exist:

* every parameter is a view of one flat fp32 *master* arena, every ``.grad`` a
  view of one flat *gradient* arena, both in bucket (gradient-ready) order
  (``buckets.py``);
* a bucket is a contiguous slice; when its last gradient lands, ONE operation
  reduces it across ranks, scales by ``1/world``, and applies the optimizer
  update in place.  On the ``b200`` backend that operation is a single sm_100a
  kernel over NVSwitch peer memory (``symm.py`` / ``csrc/comm``); on ``nccl`` /
  ``gloo`` it is an async all-reduce followed by the flat optimizer pass;
* buckets are launched from autograd hooks while backward is still running
  (side stream on CUDA), so communication overlaps compute like the Reducer.

DDP constructor behaviour that is preserved: parameter shapes
are verified across ranks, then parameters and buffers are broadcast from rank
0; buffers are re-broadcast before every training forward when they exist (K4).
The wrapper exposes ``.module`` like DDP, so the reference's
``save_checkpoint`` / ``load_checkpoint`` work on it unchanged.
"""
from __future__ import annotations

import hashlib
import os
import weakref
from typing import Callable, Dict, List, Optional

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F

from ..optim import FlatOptimizer
from ..utils.logging import get_logger
from .buckets import ArenaLayout, Bucket, MiB, build_layout
from .process_group import world

logger = get_logger("b200ddp")
_SYMM_DTYPES = (torch.float32, torch.bfloat16, torch.int32, torch.uint8, torch.int64)


def _native():
    from .. import _native as nat
    return nat


class _StepGraph:
    """A captured training step: static input placeholders, the graph, the loss it writes."""
    __slots__ = ("x", "y", "graph", "loss")


class DataParallel(nn.Module):
    def __init__(
        self,
        module: nn.Module,
        *,
        optimizer: str = "adam",
        lr: float = 1e-3,
        betas=(0.9, 0.999),
        eps: float = 1e-8,
        weight_decay: Optional[float] = None,
        momentum: float = 0.9,
        backend: str = "auto",
        compute_dtype: torch.dtype = torch.float32,
        wire_dtype: Optional[torch.dtype] = None,
        bucket_cap_mb: float = 25.0,
        first_bucket_mb: float = 1.0,
        broadcast_buffers: bool = True,
        init_sync: bool = True,
        loss_fn: Optional[Callable] = None,
        use_program: bool = True,
        overlap: bool = True,
        seed: int = 0,
        step_graph: Optional[bool] = None,
    ):
        super().__init__()
        self.module = module
        self.rank, self.world_size = world()
        params = [(n, p) for n, p in module.named_parameters() if p.requires_grad]
        if not params:
            raise ValueError("module has no trainable parameters")
        frozen = [n for n, p in module.named_parameters() if not p.requires_grad]
        if frozen:
            raise ValueError(f"frozen parameters are not supported: {frozen[:3]}")
        self.device = params[0][1].device
        self.is_cuda = self.device.type == "cuda"
        if backend == "auto":
            backend = "b200" if self.is_cuda else "gloo"
        if backend in ("b200", "nccl") and not self.is_cuda:
            raise RuntimeError(f"backend {backend!r} requires the model on a CUDA device")
        self.backend = backend
        self.compute_dtype = compute_dtype
        # SimpleNet-class MLPs on a B200 run as a fused step program: bf16 compute, fp32
        # gradients written straight into the arena by the wgrad epilogues.
        want_program = (use_program and self.is_cuda and hasattr(module, "mlp_spec")
                        and compute_dtype == torch.bfloat16 and backend in ("b200",)
                        and _native().available())
        if wire_dtype is None:
            wire_dtype = torch.float32 if want_program else compute_dtype
        self.wire_dtype = wire_dtype
        want_program = want_program and wire_dtype == torch.float32
        self.broadcast_buffers = broadcast_buffers
        if loss_fn is None:
            from ..ops.functional import cross_entropy as loss_fn  # fused kernel on CUDA, torch on CPU
        self.loss_fn = loss_fn
        self.overlap = overlap
        self.seed = seed
        self._step_index = 0

        if init_sync and self.world_size > 1:
            self._verify_shapes(params)

        # ---- flat arenas ---------------------------------------------------------
        grad_elem = 2 if self.wire_dtype == torch.bfloat16 else 4
        self.layout: ArenaLayout = build_layout(
            params, elem_bytes=grad_elem,
            caps=(int(first_bucket_mb * MiB), int(bucket_cap_mb * MiB)))
        total = self.layout.total
        self.symm = None
        if backend == "b200" and self.world_size > 1:
            from .symm import SymmetricArena
            self.symm = SymmetricArena(self.device, self.rank, self.world_size)
        # The master arena is peer-mapped: with fp32 compute the sharded kernels push the updated
        # fp32 parameters straight into every rank's master; with a bf16 shadow they push bf16 and,
        # for the parameters forward reads in fp32 (biases, norm scales / shifts: `f32_mask`), the
        # fp32 value as well — otherwise those would be current only on the shard owner.
        self.master = self._alloc(total, torch.float32, peer=True)
        self.allow_two_shot = True
        self.shadow = None
        if compute_dtype != torch.float32:
            self.shadow = self._alloc(total, compute_dtype, peer=True)
        # Gradients are what peers read: they live in peer-mapped memory on b200.  The fused MLP
        # step double-buffers them on step parity (two arenas back to back), which lets the
        # one-shot all-reduce leave without an exit barrier (csrc/comm.cu).
        self.grad_parity_stride = 0
        if want_program and self.symm is not None:
            both = self._alloc(2 * total, self.wire_dtype, peer=True)
            self.grad = both[:total]
            for attr in ("_b200_peers", "_b200_mc", "_b200_hdl"):
                if hasattr(both, attr):
                    setattr(self.grad, attr, getattr(both, attr))
            self._grad_both = both
            self.grad_parity_stride = total * both.element_size()
        else:
            self.grad = self._alloc(total, self.wire_dtype, peer=True)
        self.f32_mask = None
        if self.symm is not None and self.shadow is not None:
            mask = torch.zeros((total + 7) // 8, dtype=torch.uint8)
            for (name, p), slot in zip(params, self.layout.slots):
                if p.dim() <= 1:
                    mask[slot.offset // 8:(slot.offset + slot.numel + 7) // 8] = 1
            self.f32_mask = mask.to(self.device)
        for (name, p), slot in zip(params, self.layout.slots):
            sl = slice(slot.offset, slot.offset + slot.numel)
            with torch.no_grad():
                self.master[sl].copy_(p.detach().reshape(-1).float())
                if self.shadow is not None:
                    self.shadow[sl].copy_(self.master[sl])
                    p.data = self.shadow[sl].view(slot.shape)
                else:
                    p.data = self.master[sl].view(slot.shape)
            if self.wire_dtype == p.dtype:
                p.grad = self.grad[sl].view(slot.shape)
        self._params = [p for _, p in params]
        # Module buffers (BatchNorm running statistics) move into one flat arena per dtype and
        # become views of it, so the per-step rank-0 broadcast is one collective
        # per dtype with no gather/scatter copies.  They keep their dtype (fp32 statistics).
        self._buffer_arenas: List[torch.Tensor] = []
        by_dtype: Dict[torch.dtype, List[torch.Tensor]] = {}
        for b in module.buffers():
            by_dtype.setdefault(b.dtype, []).append(b)
        for dtype, group in by_dtype.items():
            # peer-mapped on the b200 backend: the per-forward broadcast (K4) is then one kernel of
            # ours inside the step graph (multimem.st through the switch) instead of an NCCL call
            flat = self._alloc(sum(b.numel() for b in group), dtype, peer=dtype in _SYMM_DTYPES)
            off = 0
            for b in group:
                flat[off:off + b.numel()].copy_(b.detach().reshape(-1))
                b.data = flat[off:off + b.numel()].view(b.shape)
                off += b.numel()
            self._buffer_arenas.append(flat)

        if init_sync and self.world_size > 1:
            self._broadcast_state()

        self.optimizer = FlatOptimizer(
            self.layout, self.master, self.grad, kind=optimizer, lr=lr, betas=betas, eps=eps,
            weight_decay=weight_decay, momentum=momentum, shadow=self.shadow)

        # ---- bucket-ready tracking -----------------------------------------------
        self._pending: List[int] = [0] * len(self.layout.buckets)
        self._launched: List[bool] = [False] * len(self.layout.buckets)
        self._works: Dict[int, object] = {}
        self._in_backward = False
        self._next_bucket = 0
        self._step_t = 1
        self.comm_max_blocks = 64
        self._comm_stream = torch.cuda.Stream(self.device) if self.is_cuda else None
        for p, slot in zip(self._params, self.layout.slots):
            hook = self._make_hook(slot.bucket)
            p.register_post_accumulate_grad_hook(hook)
            sl = slice(slot.offset, slot.offset + slot.numel)
            p._b200_arena = self.grad[sl].view(slot.shape)
            if self.is_cuda:
                # lets the native autograd ops write gradients straight into the arena
                # (ops/functional.py) and signal the bucket themselves
                p._b200_grad = p._b200_arena
                p._b200_master = self.master[sl].view(slot.shape)
                p._b200_ready = (lambda h=hook, q=p: h(q))

        # ---- fused step program (SimpleNet-class MLPs on a B200) ---------------------
        self.program = None
        if want_program:
            from .mlp_program import MLPProgram
            self.program = MLPProgram(self)

        # ---- whole-step CUDA graph for every other model on the b200 backend ----------
        # zero-grad, forward, loss, backward (with the per-bucket fused all-reduce+optimizer
        # kernels on the side stream) and the update are captured once per input shape and
        # replayed: a ResNet-50 step is ~1000 launches and otherwise bound by the host.
        if step_graph is None:
            step_graph = os.environ.get("B200DDP_NO_STEP_GRAPH") is None
        self._graph_ok = bool(step_graph and self.program is None and self.is_cuda and backend == "b200"
                              and _native().available())
        self._step_graphs: Dict[tuple, "_StepGraph"] = {}
        self._graph_warmup: Dict[tuple, int] = {}
        self._capturing = False
        self._dev_step_synced = False
        self._graphs_hp_version = self.optimizer.hp_version
        self._steppers = weakref.WeakSet()

    # ------------------------------------------------------------------ allocation
    def _alloc(self, n: int, dtype: torch.dtype, peer: bool = False) -> torch.Tensor:
        if peer and self.symm is not None:
            return self.symm.alloc(n, dtype)
        return torch.zeros(n, dtype=dtype, device=self.device)

    # ------------------------------------------------------------------ ctor sync
    def _verify_shapes(self, params) -> None:
        desc = ";".join(f"{n}:{tuple(p.shape)}" for n, p in params)
        digest = hashlib.sha1(desc.encode()).hexdigest()
        got: List[Optional[str]] = [None] * self.world_size
        dist.all_gather_object(got, digest)
        if any(g != got[0] for g in got):
            bad = [r for r, g in enumerate(got) if g != got[0]]
            raise RuntimeError(f"parameter shapes differ from rank 0 on ranks {bad}")

    def _broadcast_state(self) -> None:
        dist.broadcast(self.master, src=0)
        if self.shadow is not None:
            self.shadow.copy_(self.master)
        self._sync_buffers()

    @property
    def _buffers_in_graph(self) -> bool:
        """True when every buffer arena is peer-mapped: the K4 broadcast is then our kernel and is
        captured with the step."""
        return self.symm is not None and all(hasattr(f, "_b200_peers") for f in self._buffer_arenas)

    def _sync_buffers(self) -> None:
        if self.world_size == 1:
            return
        for flat in self._buffer_arenas:  # buffers are views of these (see __init__)
            if self.symm is not None and hasattr(flat, "_b200_peers"):
                self.symm.broadcast_(flat, root=0)
            else:
                dist.broadcast(flat, src=0)

    # ------------------------------------------------------------------ nn.Module
    def forward(self, *args, **kwargs):
        if self.training and self.broadcast_buffers and torch.is_grad_enabled() and (
                not self._capturing or self._buffers_in_graph):
            self._sync_buffers()
        if self.compute_dtype != torch.float32:
            args = tuple(a.to(self.compute_dtype) if torch.is_tensor(a) and a.dtype == torch.float32
                         else a for a in args)
        return self.module(*args, **kwargs)

    def parameters_count(self) -> int:
        return sum(s.numel for s in self.layout.slots)

    # ------------------------------------------------------------------ step API
    def zero_grad(self) -> None:
        self.grad.zero_()

    def train_step(self, data: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        """One full optimisation step; returns the (detached, on-device) loss."""
        self.check_graphs_current()
        if self.program is not None:
            loss = self.program.train_step(data, target)
            self.optimizer.step_count += 1
            self._step_index += 1
            return loss
        if self._graph_ok:
            return self._graphed_step(data, target)
        return self._eager_step(data, target)

    def train_steps(self, x_pool: torch.Tensor, y_pool: torch.Tensor, first: int, count: int) -> None:
        """`count` optimisation steps over an on-device pool of batches (x_pool [P, B, ...],
        y_pool [P, B]); step i uses batch (first + i) % P.  With the fused MLP session this is ONE
        native call that issues one launch per up to 256 steps (no Python, no graph replay in the loop);
        other models loop over ``train_step``.  The loss of the last step is ``last_loss()``."""
        self.check_graphs_current()
        if self.program is not None and count > 0:
            loss = self.program.train_many(x_pool, y_pool, first, count)
            if loss is not None:
                self._last_loss = loss
            self.optimizer.step_count += count
            self._step_index += count
            return
        P = x_pool.shape[0]
        for i in range(count):
            self._last_loss = self.train_step(x_pool[(first + i) % P], y_pool[(first + i) % P])

    def last_loss(self) -> torch.Tensor:
        if self.program is not None and self.program.fused is not None and self.program._last_fused:
            return self.program.fused.loss[0].clone()
        return self._last_loss

    def check_graphs_current(self) -> None:
        """Captured steps bake lr / betas / eps / weight decay into kernel arguments: when the
        optimizer's hyper-parameters changed since capture, drop the graphs and re-capture."""
        if self.optimizer.hp_version != self._graphs_hp_version:
            self.invalidate_graphs()

    def invalidate_graphs(self) -> None:
        for st in list(self._steppers):
            st.reset_runners()        # native runners hold raw graph handles
        self._step_graphs.clear()
        if self.program is not None:
            self.program.invalidate()
        self._graphs_hp_version = self.optimizer.hp_version

    def _eager_step(self, data: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        if not self._capturing:
            self._dev_step_synced = False   # eager steps advance the host counter only
        self.module.train()
        self.zero_grad()
        self.begin_backward()
        if getattr(self.module, "fused_loss", False):
            loss = self(data, target)
        else:
            out = self(data)
            loss = self.loss_fn(out, target)
        loss.backward()
        self.finish_step()
        return loss.detach()

    # -- whole-step CUDA graph ---------------------------------------------------------
    GRAPH_WARMUP_STEPS = 2     # eager steps per input shape before capture (lazy initialisation)
    GRAPH_MAX_SHAPES = 2       # e.g. the full batch and the short last batch of an epoch

    def _graphed_step(self, data: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        key = (tuple(data.shape), data.dtype, tuple(target.shape), target.dtype)
        st = self._step_graphs.get(key)
        if st is None:
            seen = self._graph_warmup.get(key, 0)
            if seen < self.GRAPH_WARMUP_STEPS or len(self._step_graphs) >= self.GRAPH_MAX_SHAPES:
                self._graph_warmup[key] = seen + 1
                return self._eager_step(data, target)
            st = self._capture_step(key, data, target)
            if st is None:   # capture failed: stay eager from now on
                return self._eager_step(data, target)
        opt = self.optimizer
        if not self._dev_step_synced:   # eager steps advance the host counter only
            opt.step_dev[0] = opt.step_count
            self._dev_step_synced = True
        if self.training and self.broadcast_buffers and not self._buffers_in_graph:
            self._sync_buffers()   # SURVEY K4 over NCCL stays outside the graph; our kernel is captured in it
        st.x.copy_(data, non_blocking=True)
        st.y.copy_(target, non_blocking=True)
        st.graph.replay()
        opt.step_count += 1
        self._step_index += 1
        return st.loss.clone()

    def _capture_step(self, key, data: torch.Tensor, target: torch.Tensor):
        opt = self.optimizer
        saved = (opt.step_count, self._step_index)
        st = _StepGraph()
        st.x, st.y = torch.empty_like(data), torch.empty_like(target)
        st.x.copy_(data)
        st.y.copy_(target)
        torch.cuda.synchronize(self.device)
        st.graph = torch.cuda.CUDAGraph()
        self._capturing = True
        try:
            # thread_local: NCCL's watchdog and the allocator may make CUDA calls on other threads
            with torch.cuda.graph(st.graph, capture_error_mode="thread_local"):
                st.loss = self._eager_step(st.x, st.y)
        except Exception as e:  # noqa: BLE001 - any capture-illegal op in a user model
            self._graph_ok = False
            import warnings
            warnings.warn(f"b200ddp: step graph capture failed, running eagerly ({type(e).__name__}: {e})")
            st = None
        finally:
            self._capturing = False
            opt.step_count, self._step_index = saved   # nothing ran: capture only records
        if st is not None:
            self._step_graphs[key] = st
        return st

    def host_stepper(self, slots: int = 3):
        """End-to-end stepping from pinned host batches with one native call per step
        (``parallel/host_stepper.py``): H2D, staging, the captured step and the loss read-back."""
        from .host_stepper import HostStepper
        st = HostStepper(self, slots=slots)
        self._steppers.add(st)
        return st

    def eval_step(self, data: torch.Tensor, target: torch.Tensor):
        """Validation batch: returns ``(mean loss, correct count)`` as device tensors."""
        if self.program is not None:
            return self.program.eval_step(data, target)
        with torch.no_grad():
            out = self(data).float()
            if out.dim() > 2:  # token models: [B, T, V] vs [B, T]
                out, target = out.reshape(-1, out.shape[-1]), target.reshape(-1)
            loss = self.loss_fn(out, target)
            correct = out.argmax(dim=1).eq(target).sum()
        return loss, correct

    # -- granular API (zero_grad -> forward -> loss.backward -> step), DDP-style -------
    def begin_backward(self) -> None:
        for b in self.layout.buckets:
            self._pending[b.index] = len(b.params)
            self._launched[b.index] = False
        self._works.clear()
        self._in_backward = True
        self._next_bucket = 0
        self._step_t = self.optimizer.step_count + 1

    def _make_hook(self, bucket_index: int):
        def hook(param):
            if not self._in_backward:
                return
            g = param.grad
            if g is not None and g.data_ptr() != param._b200_arena.data_ptr():
                # autograd produced a fresh tensor (the wire dtype differs from the parameter's):
                # cast it into the arena slice the reduction reads
                param._b200_arena.copy_(g)
                param.grad = None
            self._pending[bucket_index] -= 1
            if self._pending[bucket_index] == 0 and self.overlap:
                self._launch_ready()
        return hook

    def _launch_bucket(self, b: Bucket) -> None:
        """Starts the reduction of one bucket.  Buckets go out strictly in index order on every
        rank (the in-kernel collectives pair up by launch order)."""
        if self._launched[b.index]:
            return
        self._launched[b.index] = True
        if self.world_size == 1:
            return
        g = self.grad[b.start:b.end]
        if self.backend == "b200":
            # fused all-reduce + optimizer on the side stream; a capped grid leaves SMs for
            # the backward kernels that are still running
            self._comm_stream.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self._comm_stream):
                last = b.index == len(self.layout.buckets) - 1
                self.symm.launch_bucket(self, b, 1.0 / self.world_size, step_host=self._step_t,
                                        device_step=self._capturing, advance=self._capturing and last,
                                        max_blocks=self.comm_max_blocks if self._in_backward else 0)
            return
        if self.is_cuda:
            self._comm_stream.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self._comm_stream):
                self._works[b.index] = dist.all_reduce(g, op=dist.ReduceOp.SUM, async_op=True)
        else:
            self._works[b.index] = dist.all_reduce(g, op=dist.ReduceOp.SUM, async_op=True)

    def _launch_ready(self) -> None:
        nb = len(self.layout.buckets)
        while self._next_bucket < nb and self._pending[self._next_bucket] == 0:
            self._launch_bucket(self.layout.buckets[self._next_bucket])
            self._next_bucket += 1

    def finish_step(self) -> None:
        """Waits for every bucket's reduction and applies the optimizer."""
        self._in_backward = False
        opt = self.optimizer
        scale = 1.0 / self.world_size
        for b in self.layout.buckets:  # parameters that received no gradient: reduce zeros
            self._pending[b.index] = 0
        self._launch_ready()
        opt.step_count = self._step_t
        nat = _native() if self.is_cuda else None
        use_native = nat is not None and nat.available()
        if self.backend == "b200" and self.world_size > 1:
            torch.cuda.current_stream(self.device).wait_stream(self._comm_stream)
        else:
            for b in self.layout.buckets:
                w = self._works.get(b.index)
                if w is not None:
                    w.wait()
                if use_native:
                    nat.optimizer_step(opt, b.start, b.end, scale, device_step=self._capturing,
                                       advance=self._capturing and b.index == len(self.layout.buckets) - 1)
                else:
                    opt.step(b.start, b.end, grad_scale=scale, advance=False)
        self._step_index += 1

    def consolidate_state(self) -> None:
        """Makes master/exp_avg/exp_avg_sq complete on every rank (collective; call on all
        ranks before a checkpoint).  Buckets updated by the two-shot kernel keep optimizer
        state — and in bf16 mode the fp32 master — only on the shard owner."""
        if self.symm is None or not self.symm.two_shot_buckets:
            return
        opt = self.optimizer
        tensors = [opt.exp_avg] + ([opt.exp_avg_sq] if opt.kind != "sgd" else [])
        if self.shadow is not None:
            tensors.append(self.master)
        for k in sorted(self.symm.two_shot_buckets):
            b = self.layout.buckets[k]
            for r, (lo, hi) in enumerate(self.symm.shard_ranges(b)):
                if hi > lo:
                    for t in tensors:
                        dist.broadcast(t[lo:hi], src=r)

    # reference-style spelling: ``loss.backward(); optimizer.step()``
    def step(self) -> None:
        self.finish_step()

    # ------------------------------------------------------------------ checkpoint surface
    def model_state_dict(self) -> Dict[str, torch.Tensor]:
        """fp32, un-prefixed, one storage per tensor."""
        out: Dict[str, torch.Tensor] = {}
        by_name = {s.name: s for s in self.layout.slots}
        for k, v in self.module.state_dict().items():
            s = by_name.get(k)
            if s is not None:
                out[k] = self.master[s.offset:s.offset + s.numel].view(s.shape).clone()
            else:
                out[k] = v.detach().clone()
        return out

    def load_model_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        self.module.load_state_dict(sd)
        with torch.no_grad():
            for s in self.layout.slots:
                self.master[s.offset:s.offset + s.numel].copy_(sd[s.name].reshape(-1).float())
            if self.shadow is not None:
                self.shadow.copy_(self.master)
        if self.program is not None:
            self.program.on_params_changed()
