"""Fused step program for MLP classifiers (the reference's SimpleNet).

The reference's hot loop (``train.py:132-148``) issues ~100 kernels, 12 bucket
copies, one all-reduce and 1-2 host syncs per step for ~77 MFLOP of math
(SURVEY §2.5).  Here the whole step is a fixed schedule of sm_100a kernels that
is captured once per batch size into a CUDA graph and replayed:

    stage      x fp32 -> bf16, labels -> fixed buffer, clear the atomically
               accumulated words (bias grads, loss)               (1 launch, outside the graph)
    forward    3 tcgen05 GEMMs; epilogues fuse bias+ReLU+dropout; the last one
               fuses softmax cross-entropy forward AND backward (dlogits), the
               bias-3 gradient and the loss/accuracy reduction
    backward   2 dgrad GEMMs (epilogue = ReLU/dropout gate + next bias gradient)
               and 3 wgrad GEMMs writing fp32 straight into the gradient arena
               (wgrad runs on a forked stream, off the critical path)
    update     ONE kernel: peer-memory all-reduce + 1/W + Adam + bf16 shadow
               (plain fused Adam when world size is 1)

No activation is transposed or copied: dgrad/wgrad read the forward tensors
through MN-major UMMA descriptors.  Dropout masks are never stored: they are a
pure function of (seed, step, layer, element) and the backward gate is derived
from the stored activation (H > 0 <=> kept and positive).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional, Tuple

import torch

from .. import _native as nat
from .._native import bindings as B


class FusedSession:
    """Native session of the fused step (``csrc/mlp_fused.cu``): ONE launch per step (or per up to
    256 steps) — an 8-CTA cluster for forward + cross-entropy + backward and ``update_clusters``
    more clusters of the same grid for the all-reduce + optimizer; two launches per step when the
    update cannot ride along.  Issued from C without a CUDA graph.  Built for the reference's model
    class (ref ``train.py:32-50``): two hidden layers of 256, at most 16 classes, at most 1024
    inputs, batches of at most 64; everything else runs the GEMM-per-layer schedule below."""

    MAX_BATCH = 64

    @staticmethod
    def supports(prog: "MLPProgram") -> bool:
        L = prog.layers
        if os.environ.get("B200DDP_NO_FUSED_MLP") or len(L) != 3:
            return False
        return (L[0]["n_out"] == 256 and L[1]["n_in"] == 256 and L[1]["n_out"] == 256
                and L[2]["n_in"] == 256 and L[2]["n_out"] <= 16
                and 64 <= L[0]["n_in"] <= 1024 and L[0]["n_in"] % 8 == 0)

    def __init__(self, prog: "MLPProgram"):
        eng = prog.engine
        self.prog, self.engine, self.device = prog, eng, prog.device
        self.L = nat.lib()
        self.in_features = prog.layers[0]["n_in"]
        self.scratch = torch.empty(int(self.L.b200_mlp_scratch_bytes()), dtype=torch.uint8, device=self.device)
        self.loss = torch.zeros(2, dtype=torch.float32, device=self.device)
        self.handle: Optional[int] = None
        self.host_ready = False
        self._build()

    def _build(self) -> None:
        eng, prog, opt = self.engine, self.prog, self.engine.optimizer
        d = B.MlpDesc()
        d.in_features, d.hidden, d.classes = self.in_features, 256, prog.layers[2]["n_out"]
        (d.W1, d.W2, d.W3) = (l["w"].data_ptr() for l in prog.layers)
        (d.b1, d.b2, d.b3) = (l["b"].data_ptr() for l in prog.layers)
        (d.gW1, d.gW2, d.gW3) = (l["gw"].data_ptr() for l in prog.layers)
        (d.gb1, d.gb2, d.gb3) = (l["gb"].data_ptr() for l in prog.layers)
        d.grad_parity_stride = eng.grad_parity_stride
        d.scratch, d.loss_out = self.scratch.data_ptr(), self.loss.data_ptr()
        d.step_dev = opt.step_dev.data_ptr()
        d.dropout_p1, d.dropout_p2 = prog.layers[0]["p"], prog.layers[1]["p"]
        d.seed = prog.seed
        d.hp = B.make_hparams(opt, 1.0 / eng.world_size)
        d.master, d.grad = opt.master.data_ptr(), eng.grad.data_ptr()
        d.exp_avg = opt.exp_avg.data_ptr()
        d.exp_avg_sq = opt.exp_avg_sq.data_ptr() if opt.kind != "sgd" else None
        d.shadow = opt.shadow.data_ptr()
        buckets = eng.layout.buckets
        if len(buckets) > len(d.bucket_start):
            raise RuntimeError("too many buckets for the fused MLP session")
        d.n_buckets = len(buckets)
        comm = None
        if eng.world_size > 1:
            symm = eng.symm
            algos = [symm.pick_algo(eng, b) for b in buckets]
            symm.vec_elems = 4
            symm.two_shot_buckets.update(b.index for b, al in zip(buckets, algos) if al >= 1)
            comm = symm.comm(eng.grad, eng.shadow, master=eng.master, f32_mask=eng.f32_mask,
                             parity_stride=eng.grad_parity_stride,
                             no_exit_wait=eng.grad_parity_stride > 0)
            d.comm = C.pointer(comm)
        else:
            algos = [0] * len(buckets)
        for k, b in enumerate(buckets):
            d.bucket_start[k], d.bucket_n[k], d.bucket_algo[k] = b.start, b.numel, algos[k]
        self.algos = algos
        h = C.c_void_p()
        nat.check(self.L.b200_mlp_create(C.byref(h), C.byref(d)), "b200_mlp_create")
        self.handle = h.value
        self._keep = (d, comm)
        # > 0: the update runs on extra clusters of the step launch itself (one launch per step)
        self.update_clusters = int(self.L.b200_mlp_update_clusters(self.handle))

    def rebuild(self) -> None:
        """Hyper-parameters are baked into the session: drop it and build a new one."""
        self.close()
        self.host_ready = False
        self._build()

    def close(self) -> None:
        if self.handle is not None:
            torch.cuda.synchronize(self.device)
            self.L.b200_mlp_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001 - interpreter shutdown
            pass

    # ------------------------------------------------------------------ stepping
    def forward_backward(self, x: torch.Tensor, y: torch.Tensor, train: bool,
                         logits_out: Optional[torch.Tensor] = None) -> None:
        nat.check(self.L.b200_mlp_forward_backward(
            self.handle, x.data_ptr(), y.data_ptr(), x.shape[0], int(train),
            logits_out.data_ptr() if logits_out is not None else None, nat.stream_ptr(self.device)),
            "b200_mlp_forward_backward")

    def train(self, x_pool: torch.Tensor, y_pool: torch.Tensor, first: int, count: int) -> None:
        """`count` steps over the device-resident pool x_pool [P, bsz, in] / y_pool [P, bsz]."""
        nat.check(self.L.b200_mlp_train(
            self.handle, x_pool.data_ptr(), y_pool.data_ptr(), x_pool.shape[-2],
            x_pool.shape[0] if x_pool.dim() == 3 else 1, first, count, nat.stream_ptr(self.device)),
            "b200_mlp_train")


class _Buffers:
    def __init__(self, prog: "MLPProgram", bsz: int):
        dev, bf = prog.device, torch.bfloat16
        dims = prog.dims
        self.bsz = bsz
        self.x = torch.zeros(bsz, dims[0], dtype=bf, device=dev)
        self.y = torch.zeros(bsz, dtype=torch.int64, device=dev)
        self.h = [torch.zeros(bsz, d, dtype=bf, device=dev) for d in dims[1:-1]]
        self.ncls_ld = (dims[-1] + 15) // 16 * 16
        self.dlogits = torch.zeros(bsz, self.ncls_ld, dtype=bf, device=dev)
        self.dz = [torch.zeros(bsz, d, dtype=bf, device=dev) for d in dims[1:-1]]
        self.loss = torch.zeros(2, dtype=torch.float32, device=dev)
        # fp32 words the step adds into with atomics; the staging kernel clears them
        self.zero_train = B.make_zero_list([l["gb"] for l in prog.layers] + [self.loss])
        self.zero_eval = B.make_zero_list([self.loss])
        self.train_graph: torch.cuda.CUDAGraph | None = None
        self.eval_graph: torch.cuda.CUDAGraph | None = None


class MLPProgram:
    def __init__(self, engine):
        self.engine = engine
        self.device = engine.device
        spec = engine.module.mlp_spec()
        if engine.shadow is None or engine.shadow.dtype != torch.bfloat16:
            raise RuntimeError("the fused MLP program needs compute_dtype=bfloat16")
        if engine.grad.dtype != torch.float32:
            raise RuntimeError("the fused MLP program writes fp32 gradients (wire_dtype=float32)")
        self.layers = []
        by_param = {id(p): s for p, s in zip(engine._params, engine.layout.slots)}
        for lin, relu, p in spec:
            ws, bs = by_param[id(lin.weight)], by_param[id(lin.bias)]
            self.layers.append(dict(
                n_out=lin.out_features, n_in=lin.in_features, relu=relu, p=float(p),
                w=engine.shadow[ws.offset:ws.offset + ws.numel].view(ws.shape),          # bf16 [out,in]
                b=engine.master[bs.offset:bs.offset + bs.numel],                           # fp32 [out]
                gw=engine.grad[ws.offset:ws.offset + ws.numel].view(ws.shape),            # fp32 [out,in]
                gb=engine.grad[bs.offset:bs.offset + bs.numel]))
        for lay in self.layers[:-1]:
            if not lay["relu"]:
                raise RuntimeError("hidden layers must be Linear+ReLU(+Dropout)")
        self.dims = [self.layers[0]["n_in"]] + [l["n_out"] for l in self.layers]
        if self.dims[-1] > 32:
            raise RuntimeError("fused cross-entropy epilogue supports up to 32 classes")
        for d in self.dims[:-1]:
            if d % 8:
                raise RuntimeError("layer widths must be multiples of 8 (TMA 16-byte pitch)")
        self.seed = (int(engine.seed) * 1000003 + engine.rank * 7919 + 12345) & 0xFFFFFFFFFFFFFFFF
        self._bufs: Dict[int, _Buffers] = {}
        self._side = torch.cuda.Stream(self.device)
        self.launches_per_step = 0
        # the fused single-launch step when the model is the reference's shape class
        self.fused: Optional[FusedSession] = FusedSession(self) if FusedSession.supports(self) else None
        self._last_fused = False
        if self.fused is not None:
            self.launches_per_step = 1 if self.fused.update_clusters else 1 + len(engine.layout.buckets)

    def on_params_changed(self) -> None:
        pass  # weights are read in place from the arenas; nothing is cached

    def invalidate(self) -> None:
        """Drops the captured graphs and the native session (both bake the optimizer
        hyper-parameters into kernel arguments); the next step rebuilds."""
        for b in self._bufs.values():
            b.train_graph = None
            b.eval_graph = None
        if self.fused is not None:
            self.fused.rebuild()

    def _fused_ok(self, data: torch.Tensor, target: torch.Tensor) -> bool:
        return (self.fused is not None and data.shape[0] <= FusedSession.MAX_BATCH
                and data.dtype == torch.float32 and target.dtype == torch.int64)

    @staticmethod
    def _flat(data: torch.Tensor, target: torch.Tensor):
        x = data.reshape(data.shape[0], -1)
        if not x.is_contiguous():
            x = x.contiguous()
        return x, target.contiguous()

    def train_many(self, x_pool: torch.Tensor, y_pool: torch.Tensor, first: int, count: int):
        """`count` optimisation steps over an on-device pool of batches ([P, bsz, in], [P, bsz]):
        step i uses batch (first + i) % P.  One native call; one launch per up to 256 steps when the
        update rides in the step launch (FusedSession.update_clusters), else two launches per step."""
        if not self._fused_ok(x_pool[0], y_pool[0]) or not x_pool.is_contiguous() or not y_pool.is_contiguous():
            loss = None
            for i in range(count):
                k = (first + i) % x_pool.shape[0]
                loss = self.train_step(x_pool[k], y_pool[k])
            return loss
        self._leave_gemm_path()
        self.fused.train(x_pool.reshape(x_pool.shape[0], x_pool.shape[1], -1), y_pool, first, count)
        return None

    def _leave_gemm_path(self) -> None:
        self._last_fused = True

    def _enter_gemm_path(self) -> None:
        """The fused session's one-shot all-reduce leaves without an exit barrier (its gradients
        are double-buffered); the GEMM schedule writes buffer 0 every step, so a switch between
        the two needs the peers to have finished reading first."""
        if self._last_fused and self.engine.world_size > 1 and self.engine.grad_parity_stride:
            import torch.distributed as dist
            torch.cuda.synchronize(self.device)
            dist.barrier()
        self._last_fused = False

    # ------------------------------------------------------------------ schedule
    def _forward(self, bufs: _Buffers, train: bool) -> None:
        step_dev = self.engine.optimizer.step_dev
        a = bufs.x
        L = len(self.layers)
        for i, lay in enumerate(self.layers[:-1]):
            B.gemm(a, lay["w"], bufs.h[i], M=bufs.bsz, N=lay["n_out"], K=lay["n_in"], bias=lay["b"],
                   epilogue=B.EPI_RELU_DROPOUT, dropout_p=lay["p"] if train else 0.0,
                   seed=self.seed, step_dev=step_dev, layer_id=i)
            a = bufs.h[i]
        last = self.layers[-1]
        B.gemm(a, last["w"], bufs.dlogits, M=bufs.bsz, N=last["n_out"], K=last["n_in"],
               bias=last["b"], epilogue=B.EPI_SOFTMAX_CE, targets=bufs.y,
               ce_grad_scale=1.0 / bufs.bsz, loss_out=bufs.loss, loss_scale=1.0 / bufs.bsz,
               ce_write_grad=train, colsum=last["gb"] if train else None)

    def _backward(self, bufs: _Buffers) -> None:
        main = torch.cuda.current_stream(self.device)
        side = self._side
        acts = [bufs.x] + bufs.h  # input of layer i
        dout = bufs.dlogits        # gradient w.r.t. layer i's pre-activation, [B, n_out]
        for i in range(len(self.layers) - 1, -1, -1):
            lay = self.layers[i]
            n_out, n_in = lay["n_out"], lay["n_in"]
            # wgrad on the side stream: gW[n_out, n_in] = dout^T @ act   (both operands MN-major)
            side.wait_stream(main)
            with torch.cuda.stream(side):
                B.gemm(dout, acts[i], lay["gw"], M=n_out, N=n_in, K=bufs.bsz,
                       a_mn_major=True, b_mn_major=True)
            if i > 0:
                prev = self.layers[i - 1]
                keep = 1.0 / (1.0 - prev["p"]) if prev["p"] > 0 else 1.0
                # dgrad: dz_{i-1} = (dout @ W_i) * gate(h_{i-1});  W_i stored [n_out, n_in] = [K, N]
                B.gemm(dout, lay["w"], bufs.dz[i - 1], M=bufs.bsz, N=n_in, K=n_out, b_mn_major=True,
                       epilogue=B.EPI_GATE_POSITIVE, aux=bufs.h[i - 1], gate_scale=keep,
                       colsum=prev["gb"])
                dout = bufs.dz[i - 1]
        main.wait_stream(side)

    def _update(self) -> None:
        eng = self.engine
        scale = 1.0 / eng.world_size
        if eng.world_size > 1 and eng.symm is not None:
            eng.symm.allreduce_optimizer_step(eng, scale, device_step=True)
        elif eng.world_size > 1:
            raise RuntimeError("fused MLP program needs the b200 backend when world size > 1")
        else:
            nb = len(eng.layout.buckets)
            for k, b in enumerate(eng.layout.buckets):
                B.optimizer_step(eng.optimizer, b.start, b.end, scale, device_step=True,
                                 advance=(k == nb - 1))

    def _train_body(self, bufs: _Buffers) -> None:
        # no memsets: weight gradients are overwritten by the wgrad epilogues, the bias
        # gradients and the loss accumulator were cleared by the staging kernel
        self._forward(bufs, True)
        self._backward(bufs)
        self._update()

    def _eval_body(self, bufs: _Buffers) -> None:
        self._forward(bufs, False)

    def _capture(self, body, bufs: _Buffers) -> torch.cuda.CUDAGraph:
        # Warm up eagerly once (sets kernel attributes, allocates nothing afterwards), restoring
        # the optimizer state the warm-up step would otherwise consume.
        g = torch.cuda.CUDAGraph()
        before = B.launch_count()
        with torch.cuda.graph(g, stream=torch.cuda.Stream(self.device)):
            body(bufs)
        self.launches_per_step = B.launch_count() - before
        return g

    def _get(self, bsz: int) -> _Buffers:
        b = self._bufs.get(bsz)
        if b is None:
            b = self._bufs[bsz] = _Buffers(self, bsz)
        return b

    # ------------------------------------------------------------------ public
    def _stage(self, bufs: _Buffers, data: torch.Tensor, target: torch.Tensor,
               train: bool = True) -> None:
        x = data.reshape(data.shape[0], -1)
        if x.dtype != torch.float32 or not x.is_contiguous():
            x = x.float().contiguous()
        if x.shape[1] != self.dims[0]:
            raise ValueError(f"expected {self.dims[0]} features, got {x.shape[1]}")
        B.stage_batch(x, target.contiguous(), bufs.x, bufs.y,
                      bufs.zero_train if train else bufs.zero_eval)

    def train_step(self, data: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        if self._fused_ok(data, target):
            x, y = self._flat(data, target)
            self._leave_gemm_path()
            self.fused.train(x, y, 0, 1)
            return self.fused.loss[0].clone()
        self._enter_gemm_path()
        bufs = self._get(int(data.shape[0]))
        self._stage(bufs, data, target)
        if bufs.train_graph is None:
            bufs.train_graph = self._capture(self._train_body, bufs)
        bufs.train_graph.replay()
        return bufs.loss[0].clone()

    def eval_step(self, data: torch.Tensor, target: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        if self._fused_ok(data, target):
            x, y = self._flat(data, target)
            self.fused.forward_backward(x, y, train=False)
            out = self.fused.loss.clone()
            return out[0], out[1]
        bufs = self._get(int(data.shape[0]))
        self._stage(bufs, data, target, train=False)
        if bufs.eval_graph is None:
            bufs.eval_graph = self._capture(self._eval_body, bufs)
        bufs.eval_graph.replay()
        out = bufs.loss.clone()
        return out[0], out[1]

    def logits(self, data: torch.Tensor) -> torch.Tensor:
        """Eval-mode logits [B, classes] fp32 (uncaptured; for tests and ad-hoc inference)."""
        if self._fused_ok(data, torch.empty(0, dtype=torch.int64)):
            x, y = self._flat(data, torch.zeros(data.shape[0], dtype=torch.int64, device=self.device))
            out = torch.empty(x.shape[0], self.dims[-1], dtype=torch.float32, device=self.device)
            self.fused.forward_backward(x, y, train=False, logits_out=out)
            return out
        bufs = self._get(int(data.shape[0]))
        self._stage(bufs, data, torch.zeros(data.shape[0], dtype=torch.int64, device=self.device),
                    train=False)
        self._eval_body(bufs)
        return bufs.dlogits[:, : self.dims[-1]].float()
