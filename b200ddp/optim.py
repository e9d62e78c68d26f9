"""Flat-arena optimizers (Adam / AdamW / SGD-momentum).

The reference builds ``optim.Adam(model.parameters(), lr)`` (reference
``train.py:249``) and relies on torch's multi-tensor path: ~8 foreach launches
per step plus one host-side step counter per parameter (SURVEY §2.3 D8, §2.5
G13).  Here every parameter, gradient and moment lives at a fixed offset of one
flat arena, so the update is ONE pass over a contiguous range, and on a B200 that
pass is fused into the gradient all-reduce kernel (``csrc/comm``): the reduced
gradient never goes back to HBM before it is consumed.

Parity that is kept:

* the update arithmetic is op-for-op the one in ``torch/optim/adam.py`` /
  ``sgd.py`` (bit-exact in fp32 at world size 1, tested);
* ``state_dict()`` / ``load_state_dict()`` use torch's layout — per-parameter
  ``step`` (0-dim fp32 CPU tensor), ``exp_avg``, ``exp_avg_sq`` (each its own
  storage) and the exact ``param_groups`` keys of the installed torch — so the
  reference's ``load_checkpoint`` (``train.py:196-209``) accepts our files and we
  accept its files.
"""
from __future__ import annotations

import math
from typing import Any, Dict, Optional

import torch

from .parallel.buckets import ArenaLayout

OPT_ADAM, OPT_ADAMW, OPT_SGD = 0, 1, 2
_KINDS = {"adam": OPT_ADAM, "adamw": OPT_ADAMW, "sgd": OPT_SGD}


def _torch_param_group_template(kind: str, hp: Dict[str, Any]) -> Dict[str, Any]:
    """param_groups[0] exactly as the installed torch would emit it."""
    dummy = [torch.nn.Parameter(torch.zeros(1))]
    if kind == "adam":
        opt = torch.optim.Adam(dummy, lr=hp["lr"], betas=hp["betas"], eps=hp["eps"],
                               weight_decay=hp["weight_decay"])
    elif kind == "adamw":
        opt = torch.optim.AdamW(dummy, lr=hp["lr"], betas=hp["betas"], eps=hp["eps"],
                                weight_decay=hp["weight_decay"])
    else:
        opt = torch.optim.SGD(dummy, lr=hp["lr"], momentum=hp["momentum"],
                              weight_decay=hp["weight_decay"])
    return opt.state_dict()["param_groups"][0]


class FlatOptimizer:
    """Optimizer over a flat fp32 master arena.

    ``master`` — fp32 parameters (the values that are checkpointed);
    ``grad``   — gradient arena, fp32 or bf16, same element offsets;
    ``shadow`` — optional low-precision copy of the parameters that the compute
                 kernels read (bf16 mode); rewritten by every update.
    """

    def __init__(
        self,
        layout: ArenaLayout,
        master: torch.Tensor,
        grad: torch.Tensor,
        kind: str = "adam",
        lr: float = 1e-3,
        betas=(0.9, 0.999),
        eps: float = 1e-8,
        weight_decay: Optional[float] = None,
        momentum: float = 0.9,
        shadow: Optional[torch.Tensor] = None,
    ):
        if kind not in _KINDS:
            raise ValueError(f"unknown optimizer {kind!r}")
        self.kind = kind
        self.kind_id = _KINDS[kind]
        self.layout = layout
        self.master = master
        self.grad = grad
        self.shadow = shadow
        if weight_decay is None:
            weight_decay = 0.01 if kind == "adamw" else 0.0
        self.hp = dict(lr=float(lr), betas=(float(betas[0]), float(betas[1])), eps=float(eps),
                       weight_decay=float(weight_decay), momentum=float(momentum))
        self.step_count = 0
        # Both moment arenas exist for every kind so one kernel signature serves
        # all; SGD uses exp_avg as its momentum buffer and never touches the other.
        self.exp_avg = torch.zeros_like(master)
        self.exp_avg_sq = torch.zeros_like(master) if kind != "sgd" else master.new_zeros(1)
        # Device-resident step counter: the captured CUDA graph cannot take a new
        # host scalar every replay, so bias correction is computed on device.
        self.step_dev = torch.zeros(2, dtype=torch.int32, device=master.device)
        self.hp_version = 0   # bumped whenever hp changes (see set_hparams)
        self._group_template = _torch_param_group_template(kind, self.hp)

    # -- torch.optim surface ---------------------------------------------------
    @property
    def param_groups(self):
        g = dict(self._group_template)
        g["lr"] = self.hp["lr"]
        g["params"] = list(range(len(self.layout.slots)))
        return [g]

    def zero_grad(self, set_to_none: bool = False) -> None:
        self.grad.zero_()

    def set_lr(self, lr: float) -> None:
        self.set_hparams(lr=lr)

    def set_hparams(self, **kw) -> None:
        """Changes lr / betas / eps / weight_decay / momentum.  Captured steps bake these into
        their kernel arguments; bumping ``hp_version`` makes the engine re-capture."""
        for k, v in kw.items():
            if k not in self.hp:
                raise KeyError(f"unknown optimizer hyper-parameter {k!r}")
            self.hp[k] = (float(v[0]), float(v[1])) if k == "betas" else float(v)
        self.hp_version += 1

    # -- the update -------------------------------------------------------------
    def step(self, start: int = 0, end: Optional[int] = None, grad_scale: float = 1.0,
             advance: bool = True) -> None:
        """Reference (torch-op) implementation of the update on ``[start, end)``.

        Used on CPU and as the numerics oracle for the CUDA kernels.  ``advance``
        is False when several bucket-slices belong to the same optimizer step.
        """
        end = self.layout.total if end is None else end
        if advance:
            self.step_count += 1
        t = self.step_count
        hp = self.hp
        p = self.master[start:end]
        g = self.grad[start:end]
        if g.dtype != torch.float32:
            g = g.float()
        if grad_scale != 1.0:
            g = g * grad_scale
        lr, wd = hp["lr"], hp["weight_decay"]
        if self.kind == "sgd":
            if wd != 0.0:
                g = g.add(p, alpha=wd)
            mu = hp["momentum"]
            if mu != 0.0:
                buf = self.exp_avg[start:end]
                if t == 1:
                    buf.copy_(g)
                else:
                    buf.mul_(mu).add_(g)
                g = buf
            p.add_(g, alpha=-lr)
        else:
            b1, b2 = hp["betas"]
            m = self.exp_avg[start:end]
            v = self.exp_avg_sq[start:end]
            if self.kind == "adamw":
                p.mul_(1.0 - lr * wd)
            elif wd != 0.0:
                g = g.add(p, alpha=wd)
            m.lerp_(g, 1.0 - b1)
            v.mul_(b2).addcmul_(g, g, value=1.0 - b2)
            bc1 = 1.0 - b1 ** t
            bc2 = 1.0 - b2 ** t
            denom = (v.sqrt() / math.sqrt(bc2)).add_(hp["eps"])
            p.addcdiv_(m, denom, value=-(lr / bc1))
        if self.shadow is not None:
            self.shadow[start:end].copy_(p)

    # -- checkpoint surface --------------------------------------------------------
    def state_dict(self) -> Dict[str, Any]:
        state: Dict[int, Dict[str, torch.Tensor]] = {}
        if self.step_count > 0:
            for s in self.layout.slots:
                sl = slice(s.offset, s.offset + s.numel)
                if self.kind == "sgd":
                    if self.hp["momentum"] != 0.0:
                        state[s.index] = {"momentum_buffer": self.exp_avg[sl].view(s.shape).clone()}
                else:
                    state[s.index] = {
                        "step": torch.tensor(float(self.step_count), dtype=torch.float32),
                        "exp_avg": self.exp_avg[sl].view(s.shape).clone(),
                        "exp_avg_sq": self.exp_avg_sq[sl].view(s.shape).clone(),
                    }
        return {"state": state, "param_groups": self.param_groups}

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        groups = sd["param_groups"]
        if len(groups) != 1:
            raise ValueError("flat optimizer supports exactly one param group")
        g = groups[0]
        if len(g["params"]) != len(self.layout.slots):
            raise ValueError(
                f"checkpoint has {len(g['params'])} parameters, model has {len(self.layout.slots)}")
        self.hp["lr"] = float(g["lr"])
        if "betas" in g:
            self.hp["betas"] = (float(g["betas"][0]), float(g["betas"][1]))
        if "eps" in g:
            self.hp["eps"] = float(g["eps"])
        if "weight_decay" in g:
            self.hp["weight_decay"] = float(g["weight_decay"])
        if "momentum" in g:
            self.hp["momentum"] = float(g["momentum"])
        self._group_template = {k: v for k, v in g.items()}
        self.hp_version += 1
        steps = set()
        index_of = {pid: i for i, pid in enumerate(g["params"])}
        self.exp_avg.zero_()
        if self.kind != "sgd":
            self.exp_avg_sq.zero_()
        for pid, st in sd["state"].items():
            s = self.layout.slots[index_of[pid]]
            sl = slice(s.offset, s.offset + s.numel)
            if self.kind == "sgd":
                if st.get("momentum_buffer") is not None:
                    self.exp_avg[sl].copy_(st["momentum_buffer"].reshape(-1))
                    steps.add(1)
            else:
                if tuple(st["exp_avg"].shape) != tuple(s.shape):
                    raise ValueError(f"shape mismatch for parameter {s.name}")
                self.exp_avg[sl].copy_(st["exp_avg"].reshape(-1))
                self.exp_avg_sq[sl].copy_(st["exp_avg_sq"].reshape(-1))
                steps.add(int(float(st["step"])))
        if len(steps) > 1:
            raise ValueError(f"per-parameter step counts disagree: {sorted(steps)}")
        self.step_count = steps.pop() if steps else 0
        self.step_dev[0] = self.step_count
