"""GPT-2 (BASELINE.json config 4: "GPT-2 small DDP bf16 seq 1024").

Not part of the reference repository — it is one of the benchmark extensions the
baseline definition names; the comparator is HF ``GPT2LMHeadModel`` under the
reference's DDP loop.  Architecture and parameter count match HF
GPT-2 small exactly (124,439,808 with the LM head tied to ``wte``); weights use
the ``nn.Linear`` [out, in] layout, which is the K-major B operand of our GEMM.

Every matmul (qkv, attention projection, MLP up/down, LM head) runs on the
tcgen05 GEMM with fused epilogues (bias, GELU + pre-activation save, residual
add, GELU-backward gate); LayerNorm and the LM-head cross-entropy are our
kernels; gradients are written into the engine's arena by the wgrad epilogues.
The softmax(QK^T)V core is the fused flash-style tcgen05 kernel pair of
``csrc/flash_attention.cu`` (``B200DDP_ATTENTION=gemm|sdpa`` select the batched-GEMM path or torch's
library kernel for comparison, see ops/functional.py).
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch
import torch.nn as nn

from ..ops import functional as Fn


@dataclass
class GPT2Config:
    vocab_size: int = 50257
    n_positions: int = 1024
    n_embd: int = 768
    n_layer: int = 12
    n_head: int = 12
    layer_norm_eps: float = 1e-5


class Block(nn.Module):
    def __init__(self, cfg: GPT2Config):
        super().__init__()
        C = cfg.n_embd
        self.n_head = cfg.n_head
        self.eps = cfg.layer_norm_eps
        self.ln_1 = nn.LayerNorm(C, eps=cfg.layer_norm_eps)
        self.attn_qkv = nn.Linear(C, 3 * C)
        self.attn_proj = nn.Linear(C, C)
        self.ln_2 = nn.LayerNorm(C, eps=cfg.layer_norm_eps)
        self.mlp_fc = nn.Linear(C, 4 * C)
        self.mlp_proj = nn.Linear(4 * C, C)

    def forward(self, x):
        h = Fn.layer_norm(x, self.ln_1.weight, self.ln_1.bias, self.eps)
        qkv = Fn.linear(h, self.attn_qkv.weight, self.attn_qkv.bias)
        a = Fn.causal_attention(qkv, self.n_head)
        x = Fn.linear(a, self.attn_proj.weight, self.attn_proj.bias, residual=x)
        h = Fn.layer_norm(x, self.ln_2.weight, self.ln_2.bias, self.eps)
        return Fn.mlp(h, self.mlp_fc.weight, self.mlp_fc.bias, self.mlp_proj.weight,
                      self.mlp_proj.bias, residual=x)


class GPT2(nn.Module):
    fused_loss = True  # forward(idx, targets) returns the loss (LM head + CE are one op)

    def __init__(self, cfg: GPT2Config = GPT2Config()):
        super().__init__()
        self.cfg = cfg
        self.wte = nn.Embedding(cfg.vocab_size, cfg.n_embd)
        self.wpe = nn.Embedding(cfg.n_positions, cfg.n_embd)
        self.h = nn.ModuleList(Block(cfg) for _ in range(cfg.n_layer))
        self.ln_f = nn.LayerNorm(cfg.n_embd, eps=cfg.layer_norm_eps)
        self.apply(self._init)
        for blk in self.h:  # GPT-2 residual-projection scaling
            for lin in (blk.attn_proj, blk.mlp_proj):
                nn.init.normal_(lin.weight, mean=0.0, std=0.02 / math.sqrt(2 * cfg.n_layer))

    @staticmethod
    def _init(m):
        if isinstance(m, nn.Linear):
            nn.init.normal_(m.weight, mean=0.0, std=0.02)
            nn.init.zeros_(m.bias)
        elif isinstance(m, nn.Embedding):
            nn.init.normal_(m.weight, mean=0.0, std=0.02)

    def hidden(self, idx: torch.Tensor) -> torch.Tensor:
        x = Fn.embedding(idx, self.wte.weight, self.wpe.weight)
        for blk in self.h:
            x = blk(x)
        return Fn.layer_norm(x, self.ln_f.weight, self.ln_f.bias, self.cfg.layer_norm_eps)

    def forward(self, idx: torch.Tensor, targets: torch.Tensor | None = None):
        x = self.hidden(idx)
        if targets is None:
            return torch.nn.functional.linear(x, self.wte.weight)
        # wte also feeds the embedding: its bucket is signalled by the embedding backward
        return Fn.lm_head_cross_entropy(x, self.wte.weight, targets, signal_weight=False)
