"""ctypes binding of the host numerics oracle (``oracle/oracle.cpp``).

Plain C++ restatements of what the sm_100a kernels compute, built with g++ (no CUDA), so the
kernel arithmetic has a reference that runs on any box.  Arguments are CPU tensors; bf16 tensors
are passed as their raw 16-bit storage.  The optimizer update is the very code the device kernels
compile (``csrc/optim_math.h``).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from .bindings import OptHParams

_lib: Optional[C.CDLL] = None

EPI_NONE, EPI_RELU, EPI_GELU, EPI_GELU_BWD = 0, 1, 3, 4


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        from .build import build_oracle
        _lib = C.CDLL(build_oracle())
        for name in ("oracle_optimizer_step", "oracle_allreduce_sum", "oracle_gemm_bf16", "oracle_cross_entropy",
                     "oracle_layernorm_fwd", "oracle_layernorm_bwd", "oracle_batchnorm_fwd", "oracle_im2col",
                     "oracle_col2im", "oracle_causal_attention", "oracle_maxpool_fwd", "oracle_maxpool_bwd",
                     "oracle_avgpool_fwd"):
            getattr(_lib, name).restype = None
    return _lib


def _p(t: Optional[torch.Tensor]):
    if t is None:
        return None
    assert t.device.type == "cpu" and t.is_contiguous()
    return C.c_void_p(t.data_ptr())


def optimizer_step(kind: str, p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: Optional[torch.Tensor], t: int, *,
                   lr: float, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0,
                   momentum: float = 0.0, grad_scale: float = 1.0, shadow: Optional[torch.Tensor] = None) -> None:
    """In-place update of fp32 p/m/v with the kernels' own update code."""
    hp = OptHParams({"adam": 0, "adamw": 1, "sgd": 2}[kind], lr, betas[0], betas[1], eps, weight_decay, momentum,
                    grad_scale)
    lib().oracle_optimizer_step(C.byref(hp), _p(p), _p(g), _p(m), _p(v), _p(shadow), C.c_int64(p.numel()), C.c_int(t))


def allreduce_sum(parts, scale: float = 1.0) -> torch.Tensor:
    out = torch.empty_like(parts[0])
    arr = (C.c_void_p * len(parts))(*[p.data_ptr() for p in parts])
    lib().oracle_allreduce_sum(arr, C.c_int(len(parts)), C.c_int64(out.numel()), C.c_float(scale), _p(out))
    return out


def gemm(a: torch.Tensor, b: torch.Tensor, bias: Optional[torch.Tensor] = None, epilogue: int = EPI_NONE,
         aux: Optional[torch.Tensor] = None, want_aux_out: bool = False) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """epi(a[M,K] @ b[N,K]^T) on bf16 tensors; returns (D bf16, pre-activation bf16 or None)."""
    assert a.dtype == b.dtype == torch.bfloat16
    M, K = a.shape
    N = b.shape[0]
    d = torch.empty(M, N, dtype=torch.bfloat16)
    z = torch.empty(M, N, dtype=torch.bfloat16) if want_aux_out else None
    lib().oracle_gemm_bf16(_p(a), _p(b), _p(d), M, N, K, _p(bias), epilogue, _p(aux), _p(z))
    return d, z


def cross_entropy(logits: torch.Tensor, targets: torch.Tensor, grad_scale: float):
    """(mean loss, correct count, dlogits) of fp32 logits."""
    rows, cols = logits.shape
    dl = torch.empty_like(logits)
    out = torch.zeros(2)
    lib().oracle_cross_entropy(_p(logits), _p(targets), C.c_int64(rows), cols, _p(dl), C.c_float(grad_scale), _p(out))
    return float(out[0]), float(out[1]), dl


def layernorm_fwd(x, gamma, beta, eps: float = 1e-5):
    rows, Cc = x.shape
    y = torch.empty_like(x)
    mean, rstd = torch.empty(rows), torch.empty(rows)
    lib().oracle_layernorm_fwd(_p(x), _p(gamma), _p(beta), _p(y), _p(mean), _p(rstd), C.c_int64(rows), Cc, C.c_float(eps))
    return y, mean, rstd


def layernorm_bwd(dy, x, gamma, mean, rstd):
    rows, Cc = x.shape
    dx = torch.empty_like(x)
    dg, db = torch.empty(Cc), torch.empty(Cc)
    lib().oracle_layernorm_bwd(_p(dy), _p(x), _p(gamma), _p(mean), _p(rstd), _p(dx), _p(dg), _p(db), C.c_int64(rows), Cc)
    return dx, dg, db


def batchnorm_fwd(x, gamma, beta, running_mean, running_var, eps=1e-5, momentum=0.1, relu=False, residual=None):
    """x [rows, C] bf16 (channels-last, flattened); running statistics updated in place."""
    rows, Cc = x.shape
    y = torch.empty_like(x)
    lib().oracle_batchnorm_fwd(_p(x), _p(residual), _p(y), _p(gamma), _p(beta), _p(running_mean), _p(running_var),
                               C.c_int64(rows), Cc, C.c_float(eps), C.c_float(momentum), int(relu))
    return y


def im2col(x_nhwc, R, S, stride, pad):
    N, H, W, Cc = x_nhwc.shape
    P, Q = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1
    col = torch.empty(N * P * Q, R * S * Cc, dtype=torch.bfloat16)
    lib().oracle_im2col(_p(x_nhwc), _p(col), N, H, W, Cc, R, S, stride, pad)
    return col


def col2im(dcol, shape_nhwc, R, S, stride, pad):
    N, H, W, Cc = shape_nhwc
    dx = torch.empty(N, H, W, Cc, dtype=torch.bfloat16)
    lib().oracle_col2im(_p(dcol), _p(dx), N, H, W, Cc, R, S, stride, pad)
    return dx


def maxpool_fwd(x_nhwc, k, stride, pad):
    N, H, W, Cc = x_nhwc.shape
    P, Q = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    y = torch.empty(N, P, Q, Cc, dtype=torch.bfloat16)
    idx = torch.empty(N, P, Q, Cc, dtype=torch.uint8)
    lib().oracle_maxpool_fwd(_p(x_nhwc), _p(y), _p(idx), N, H, W, Cc, k, k, stride, pad)
    return y, idx


def maxpool_bwd(dy_nhwc, idx, shape_nhwc, k, stride, pad):
    N, H, W, Cc = shape_nhwc
    dx = torch.empty(N, H, W, Cc, dtype=torch.bfloat16)
    lib().oracle_maxpool_bwd(_p(dy_nhwc), _p(idx), _p(dx), N, H, W, Cc, k, k, stride, pad)
    return dx


def avgpool_fwd(x_nhwc):
    N, H, W, Cc = x_nhwc.shape
    y = torch.empty(N, Cc, dtype=torch.bfloat16)
    lib().oracle_avgpool_fwd(_p(x_nhwc), _p(y), N, H * W, Cc)
    return y


def causal_attention(qkv: torch.Tensor, n_head: int):
    """qkv [B, T, 3*H*D] bf16 -> (out [B, T, H*D] bf16, lse [B, H, T] fp32)."""
    B, T, C3 = qkv.shape
    D = C3 // 3 // n_head
    out = torch.empty(B, T, n_head * D, dtype=torch.bfloat16)
    lse = torch.empty(B, n_head, T)
    lib().oracle_causal_attention(_p(qkv), _p(out), _p(lse), B, T, n_head, D)
    return out, lse
