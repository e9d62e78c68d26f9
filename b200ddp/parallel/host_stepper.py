"""End-to-end stepping from pinned host batches through the native runner (csrc/runner.cu).

The reference's loop moves every batch with ``.to(device)`` and reads ``loss.item()`` after every
step (ref ``train.py:133,141``).  ``HostStepper`` is the same contract — host tensors in, a Python
float out — with the per-step host work done by ONE C call: H2D on a copy stream, staging, launch
of the captured step, D2H of the loss, and a wait for the *previous* step's loss, so one step is
always queued behind the one the GPU is running.

    stepper = engine.host_stepper()
    for hx, hy in pinned_batches:
        loss_of_previous_step = stepper.step(hx, hy)      # None on the first call
    last = stepper.finish()

Until the engine has a captured step for the batch shape (the MLP program captures on first use;
other models after ``DataParallel.GRAPH_WARMUP_STEPS`` eager steps) ``step`` takes the ordinary
``train_step`` path, so it is always correct to use.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional

import torch

from .. import _native as nat
from .._native import bindings as B


class HostStepper:
    def __init__(self, engine, slots: int = 3):
        if not engine.is_cuda or not nat.available():
            raise RuntimeError("HostStepper needs the model on a CUDA device and the native library")
        self.engine = engine
        self.device = engine.device
        self.slots = slots
        self._runners: Dict[tuple, int] = {}        # shape key -> native runner handle
        self._live: Optional[int] = None            # runner with a step in flight
        self._slow_pending: Optional[torch.Tensor] = None
        self._carry: Optional[float] = None         # loss collected by reset_runners(), not yet returned
        self._prev = C.c_float(0.0)
        self._valid = C.c_int(0)

    # ------------------------------------------------------------------ graph lookup
    def _captured(self, hx: torch.Tensor, hy: torch.Tensor):
        """(graph, mode, dst0, dst1, loss_ptr, zero_list) of the captured step for this shape,
        or None while the engine is still warming up."""
        eng = self.engine
        if eng.program is not None:
            bsz = int(hx.shape[0])
            prog = eng.program
            bufs = prog._get(bsz)
            if bufs.train_graph is None:
                bufs.train_graph = prog._capture(prog._train_body, bufs)
            return bufs.train_graph, 1, bufs.x, bufs.y, bufs.loss.data_ptr(), bufs.zero_train
        key = (tuple(hx.shape), hx.dtype, tuple(hy.shape), hy.dtype)
        st = eng._step_graphs.get(key)
        if st is None:
            return None
        return st.graph, 0, st.x, st.y, st.loss.data_ptr(), None

    def _runner_for(self, hx: torch.Tensor, hy: torch.Tensor) -> Optional[int]:
        key = (tuple(hx.shape), hx.dtype, tuple(hy.shape), hy.dtype)
        h = self._runners.get(key)
        if h is not None:
            return h
        cap = self._captured(hx, hy)
        if cap is None:
            return None
        graph, mode, dst0, dst1, loss_ptr, zero_list = cap
        d = B.RunnerDesc()
        d.graph_exec = graph.raw_cuda_graph_exec()
        d.stream = torch.cuda.current_stream(self.device).cuda_stream
        d.mode, d.slots = mode, self.slots
        d.in_bytes[0], d.in_bytes[1] = hx.numel() * hx.element_size(), hy.numel() * hy.element_size()
        d.dst[0], d.dst[1] = dst0.data_ptr(), dst1.data_ptr()
        d.loss_dev = loss_ptr
        d.nzero = 0
        if zero_list is not None:
            ptrs, counts, n = zero_list
            d.nzero = n
            for k in range(n):
                d.zero_ptrs[k], d.zero_counts[k] = ptrs[k], counts[k]
        if mode == 1 and (hx.dtype != torch.float32 or hy.dtype != torch.int64):
            raise TypeError("the MLP program stages fp32 features and int64 labels")
        if mode == 0 and (hx.dtype != dst0.dtype or hy.dtype != dst1.dtype):
            raise TypeError("host batch dtypes differ from the captured step's inputs")
        handle = C.c_void_p()
        nat.check(nat.lib().b200_runner_create(C.byref(handle), C.byref(d)), "b200_runner_create")
        self._runners[key] = handle.value
        return handle.value

    # ------------------------------------------------------------------ stepping
    def step(self, hx: torch.Tensor, hy: torch.Tensor) -> Optional[float]:
        """Submits one training step on pinned host tensors; returns the loss of the step
        submitted before it (``None`` on the very first call)."""
        if not (hx.is_pinned() and hy.is_pinned() and hx.is_contiguous() and hy.is_contiguous()):
            raise ValueError("HostStepper.step wants contiguous pinned host tensors")
        eng = self.engine
        eng.check_graphs_current()       # hyper-parameters changed: graphs (and runners) are rebuilt
        if eng.program is not None and eng.program._fused_ok(hx, hy):
            return self._fused_step(hx, hy)
        runner = self._runner_for(hx, hy)
        if runner is None:   # engine still warming up for this shape: ordinary path
            prev = self._take_previous()
            self._slow_pending = eng.train_step(hx.to(self.device, non_blocking=True),
                                                hy.to(self.device, non_blocking=True))
            return prev
        # the previous step's loss comes back from the native call itself when it ran on the
        # same runner; otherwise collect it first
        prev = None if self._live == runner else self._take_previous()
        opt = eng.optimizer
        if eng.program is None:
            if not eng._dev_step_synced:     # eager steps advance the host counter only
                opt.step_dev[0] = opt.step_count
                eng._dev_step_synced = True
            if eng.training and eng.broadcast_buffers and not eng._buffers_in_graph:
                eng._sync_buffers()          # SURVEY K4 over NCCL (our broadcast kernel is inside the graph)
        same = self._live == runner
        nat.check(nat.lib().b200_runner_step(runner, hx.data_ptr(), hy.data_ptr(), C.byref(self._prev),
                                             C.byref(self._valid)), "b200_runner_step")
        self._live = runner
        opt.step_count += 1
        eng._step_index += 1
        if same and self._valid.value:
            return float(self._prev.value)
        return prev

    FUSED = -1   # `_live` value while the fused MLP session has a step in flight

    def _fused_step(self, hx: torch.Tensor, hy: torch.Tensor) -> Optional[float]:
        """Fused single-launch MLP step (csrc/mlp_fused.cu): the session copies the pinned batch into
        its device ring on a copy stream, launches the step kernel (which stores the loss into
        mapped pinned memory itself) and the update, and hands back the previous step's loss."""
        eng = self.engine
        fs = eng.program.fused
        L = nat.lib()
        if not fs.host_ready:
            nat.check(L.b200_mlp_host_init(fs.handle, eng.program.fused.MAX_BATCH, self.slots), "b200_mlp_host_init")
            fs.host_ready = True
        prev = None if self._live == self.FUSED else self._take_previous()
        same = self._live == self.FUSED
        eng.program._leave_gemm_path()
        x = hx.reshape(hx.shape[0], -1)
        nat.check(L.b200_mlp_host_step(fs.handle, x.data_ptr(), hy.data_ptr(), x.shape[0], C.byref(self._prev),
                                       C.byref(self._valid), nat.stream_ptr(self.device)), "b200_mlp_host_step")
        self._live = self.FUSED
        eng.optimizer.step_count += 1
        eng._step_index += 1
        if same and self._valid.value:
            return float(self._prev.value)
        return prev

    # ------------------------------------------------------------------ many steps per call
    def run(self, batches) -> list:
        """Trains on every (pinned_x, pinned_y) of `batches` and returns their losses.  For the fused MLP
        session this is ONE native call: each step still copies its own inputs host -> device and
        hands its own loss back to the host (ref train.py:133, :141), but the kernel launches run
        up to 256 steps each and wait for every step's input copy themselves.  Other models: a
        loop over ``step``."""
        batches = list(batches)
        if not batches:
            return []
        ptrs = self._fused_run_ok(batches[0][0], batches[0][1])
        if not ptrs or any(b[0].shape != batches[0][0].shape or not b[0].is_pinned() for b in batches):
            out = []
            for hx, hy in batches:
                prev = self.step(hx, hy)
                if prev is not None:
                    out.append(prev)
            out.append(self.finish())
            return out[-len(batches):]
        import numpy as np
        xp = np.fromiter((b[0].data_ptr() for b in batches), dtype=np.uint64, count=len(batches))
        yp = np.fromiter((b[1].data_ptr() for b in batches), dtype=np.uint64, count=len(batches))
        return self._fused_run(xp, yp, batches[0][0].shape[0])

    def run_pool(self, x_pool: torch.Tensor, y_pool: torch.Tensor, first: int, count: int) -> list:
        """``run`` over a pinned host pool x_pool [P, B, ...] / y_pool [P, B]: step i trains on batch
        (first + i) % P.  The host-memory counterpart of ``engine.train_steps``."""
        P = x_pool.shape[0]
        if (count > 0 and x_pool.is_pinned() and y_pool.is_pinned() and x_pool.is_contiguous() and y_pool.is_contiguous()
                and self._fused_run_ok(x_pool[0], y_pool[0])):
            import numpy as np
            idx = (first + np.arange(count, dtype=np.uint64)) % np.uint64(P)
            xp = np.uint64(x_pool.data_ptr()) + idx * np.uint64(x_pool.stride(0) * x_pool.element_size())
            yp = np.uint64(y_pool.data_ptr()) + idx * np.uint64(y_pool.stride(0) * y_pool.element_size())
            return self._fused_run(xp, yp, x_pool.shape[1])
        return self.run([(x_pool[(first + i) % P], y_pool[(first + i) % P]) for i in range(count)])

    def _fused_run_ok(self, hx: torch.Tensor, hy: torch.Tensor) -> bool:
        eng = self.engine
        return bool(eng.program is not None and eng.program._fused_ok(hx, hy) and eng.program.fused.update_clusters > 0
                    and hx.is_contiguous() and hy.is_contiguous() and os.environ.get("B200DDP_MLP_EVENT_WAIT") is None)

    def _fused_run(self, xp, yp, bsz: int) -> list:
        import numpy as np
        eng = self.engine
        fs = eng.program.fused
        L = nat.lib()
        if not fs.host_ready:
            nat.check(L.b200_mlp_host_init(fs.handle, fs.MAX_BATCH, self.slots), "b200_mlp_host_init")
            fs.host_ready = True
        carried = self._take_previous()       # a step submitted through step() is finished first
        if carried is not None:
            self._carry = carried
        eng.program._leave_gemm_path()
        n = int(xp.shape[0])
        losses = np.empty(n, dtype=np.float32)
        nat.check(L.b200_mlp_host_run(fs.handle, xp.ctypes.data, yp.ctypes.data, n, int(bsz), losses.ctypes.data,
                                      nat.stream_ptr(self.device)), "b200_mlp_host_run")
        eng.optimizer.step_count += n
        eng._step_index += n
        return losses.tolist()

    def _take_previous(self) -> Optional[float]:
        """Loss of the last submitted step, whichever path it took; clears the pending state."""
        if self._carry is not None:
            prev, self._carry = self._carry, None
            return prev
        if self._live == self.FUSED:
            fs = self.engine.program.fused
            nat.check(nat.lib().b200_mlp_host_finish(fs.handle, C.byref(self._prev), C.byref(self._valid),
                                                     nat.stream_ptr(self.device)), "b200_mlp_host_finish")
            self._live = None
            return float(self._prev.value) if self._valid.value else None
        if self._live is not None:
            nat.check(nat.lib().b200_runner_finish(self._live, C.byref(self._prev), C.byref(self._valid)),
                      "b200_runner_finish")
            self._live = None
            return float(self._prev.value) if self._valid.value else None
        if self._slow_pending is not None:
            prev, self._slow_pending = float(self._slow_pending), None
            return prev
        return None

    def finish(self) -> Optional[float]:
        """Waits for the last submitted step and returns its loss."""
        return self._take_previous()

    def reset_runners(self) -> None:
        """Destroys the native runners (the engine is about to drop the graphs they launch); the
        loss of the step in flight is kept and returned by the next ``step`` / ``finish``."""
        prev = self._take_previous()
        if prev is not None:
            self._carry = prev
        for h in self._runners.values():
            nat.lib().b200_runner_destroy(h)
        self._runners.clear()

    def close(self) -> None:
        self.reset_runners()
        self._carry = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001 - interpreter shutdown
            pass
