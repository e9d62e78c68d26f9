"""ctypes signatures and thin tensor-level wrappers for ``include/b200ddp.h``."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import check, lib, stream_ptr

EPI_NONE, EPI_RELU_DROPOUT, EPI_GATE_POSITIVE, EPI_GELU, EPI_GELU_BWD, EPI_SOFTMAX_CE = range(6)


class OptHParams(C.Structure):
    _fields_ = [("kind", C.c_int), ("lr", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float),
                ("eps", C.c_float), ("weight_decay", C.c_float), ("momentum", C.c_float),
                ("grad_scale", C.c_float)]


class GemmDesc(C.Structure):
    _fields_ = [
        ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
        ("A", C.c_void_p), ("lda", C.c_int64), ("a_mn_major", C.c_int),
        ("B", C.c_void_p), ("ldb", C.c_int64), ("b_mn_major", C.c_int),
        ("D", C.c_void_p), ("ldd", C.c_int64), ("d_fp32", C.c_int),
        ("bias", C.c_void_p),
        ("epilogue", C.c_int),
        ("aux", C.c_void_p), ("ldaux", C.c_int64),
        ("aux_out", C.c_void_p),
        ("colsum", C.c_void_p),
        ("dropout_p", C.c_float), ("gate_scale", C.c_float),
        ("seed", C.c_uint64),
        ("step_dev", C.c_void_p),
        ("layer_id", C.c_uint32),
        ("targets", C.c_void_p),
        ("ce_grad_scale", C.c_float),
        ("loss_out", C.c_void_p),
        ("loss_scale", C.c_float),
        ("ce_write_grad", C.c_int),
        ("tile_n", C.c_int),
        ("max_ctas", C.c_int),
        ("batch_inner", C.c_int), ("batch_outer", C.c_int),
        ("a_stride_inner", C.c_int64), ("a_stride_outer", C.c_int64),
        ("b_stride_inner", C.c_int64), ("b_stride_outer", C.c_int64),
        ("d_stride_inner", C.c_int64), ("d_stride_outer", C.c_int64),
        ("causal", C.c_int),
        ("tile_m", C.c_int),
        ("split_k", C.c_int),
        ("cta_pair", C.c_int),
        ("a_conv", C.c_int), ("b_conv", C.c_int),
        ("conv_n", C.c_int), ("conv_h", C.c_int), ("conv_w", C.c_int), ("conv_c", C.c_int),
        ("conv_r", C.c_int), ("conv_s", C.c_int), ("conv_stride", C.c_int), ("conv_pad", C.c_int),
    ]


class RunnerDesc(C.Structure):
    _fields_ = [
        ("graph_exec", C.c_void_p), ("stream", C.c_void_p),
        ("mode", C.c_int), ("slots", C.c_int),
        ("in_bytes", C.c_int64 * 2),
        ("dst", C.c_void_p * 2),
        ("loss_dev", C.c_void_p),
        ("zero_ptrs", C.c_void_p * 8),
        ("zero_counts", C.c_int * 8),
        ("nzero", C.c_int),
    ]


class MlpDesc(C.Structure):
    from .bindings_comm import Comm as _Comm
    _fields_ = [
        ("in_features", C.c_int), ("hidden", C.c_int), ("classes", C.c_int),
        ("W1", C.c_void_p), ("W2", C.c_void_p), ("W3", C.c_void_p),
        ("b1", C.c_void_p), ("b2", C.c_void_p), ("b3", C.c_void_p),
        ("gW1", C.c_void_p), ("gb1", C.c_void_p), ("gW2", C.c_void_p), ("gb2", C.c_void_p),
        ("gW3", C.c_void_p), ("gb3", C.c_void_p),
        ("grad_parity_stride", C.c_int64),
        ("scratch", C.c_void_p), ("loss_out", C.c_void_p), ("step_dev", C.c_void_p),
        ("dropout_p1", C.c_float), ("dropout_p2", C.c_float), ("seed", C.c_uint64),
        ("hp", OptHParams),
        ("master", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
        ("shadow", C.c_void_p),
        ("n_buckets", C.c_int),
        ("bucket_start", C.c_int64 * 4), ("bucket_n", C.c_int64 * 4), ("bucket_algo", C.c_int * 4),
        ("comm", C.POINTER(_Comm)), ("comm_copy", _Comm),
    ]


def declare(L: C.CDLL) -> None:
    L.b200_embedding_fwd.argtypes = [C.c_void_p] * 4 + [C.c_int64, C.c_int, C.c_int, C.c_int, C.c_void_p]
    L.b200_embedding_bwd.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int,
                                     C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    L.b200_embedding_scratch_bytes.argtypes = [C.c_int64]
    L.b200_embedding_scratch_bytes.restype = C.c_int64
    L.b200_embedding_fwd.restype = L.b200_embedding_bwd.restype = C.c_int
    L.b200_mlp_scratch_bytes.restype = C.c_int64
    L.b200_mlp_create.argtypes = [C.POINTER(C.c_void_p), C.POINTER(MlpDesc)]
    L.b200_mlp_destroy.argtypes = [C.c_void_p]
    L.b200_mlp_forward_backward.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    L.b200_mlp_train.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]
    L.b200_mlp_host_init.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.b200_mlp_host_step.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_float),
                                     C.POINTER(C.c_int), C.c_void_p]
    L.b200_mlp_host_finish.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int), C.c_void_p]
    L.b200_mlp_host_run.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p]
    L.b200_mlp_host_run.restype = C.c_int
    L.b200_mlp_update_clusters.argtypes = [C.c_void_p]
    L.b200_mlp_update_clusters.restype = C.c_int
    for name in ("b200_mlp_create", "b200_mlp_destroy", "b200_mlp_forward_backward", "b200_mlp_train",
                 "b200_mlp_host_init", "b200_mlp_host_step", "b200_mlp_host_finish"):
        getattr(L, name).restype = C.c_int
    L.b200_maxpool_fwd.argtypes = [C.c_void_p] * 3 + [C.c_int] * 8 + [C.c_void_p]
    L.b200_maxpool_bwd.argtypes = [C.c_void_p] * 3 + [C.c_int] * 8 + [C.c_void_p]
    L.b200_avgpool_fwd.argtypes = [C.c_void_p] * 2 + [C.c_int] * 3 + [C.c_void_p]
    L.b200_avgpool_bwd.argtypes = [C.c_void_p] * 2 + [C.c_int] * 3 + [C.c_void_p]
    L.b200_runner_create.argtypes = [C.POINTER(C.c_void_p), C.POINTER(RunnerDesc)]
    L.b200_runner_step.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int)]
    L.b200_runner_finish.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int)]
    L.b200_runner_destroy.argtypes = [C.c_void_p]
    L.b200_last_error.restype = C.c_char_p
    L.b200_launch_count.restype = C.c_uint64
    L.b200_device_info.argtypes = [C.POINTER(C.c_int)]
    L.b200_optimizer_step.argtypes = [C.POINTER(OptHParams), C.c_void_p, C.c_void_p, C.c_int,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                                      C.c_int, C.c_int, C.c_void_p]
    L.b200_ce_scratch_bytes.argtypes = [C.c_int64]
    L.b200_ce_scratch_bytes.restype = C.c_int64
    L.b200_cross_entropy.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_int64, C.c_int,
                                     C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
    L.b200_gemm.argtypes = [C.POINTER(GemmDesc), C.c_void_p]
    L.b200_gemm_plan.argtypes = [C.POINTER(GemmDesc), C.c_int, C.POINTER(C.c_int)]
    L.b200_cast_f32_to_bf16.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
    L.b200_cast_bf16_to_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
    L.b200_stage_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                   C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    L.b200_stage_batch.restype = C.c_int
    L.b200_layernorm_fwd.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_void_p]
    L.b200_layernorm_bwd.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                     C.c_int64, C.c_int, C.c_void_p, C.c_void_p]
    L.b200_colsum.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_void_p, C.c_int,
                              C.c_void_p, C.c_void_p]
    for name in ("b200_layernorm_bwd_scratch_bytes", "b200_colsum_scratch_bytes"):
        getattr(L, name).argtypes = [C.c_int64, C.c_int]
        getattr(L, name).restype = C.c_int64
    L.b200_flash_attn_fwd.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                      C.c_void_p]
    L.b200_flash_attn_fwd.restype = C.c_int
    if hasattr(L, "b200_flash_attn_bwd"):
        L.b200_flash_attn_bwd.argtypes = [C.c_void_p] * 7 + [C.c_int] * 4 + [C.c_void_p]
        L.b200_flash_attn_bwd.restype = C.c_int
    L.b200_causal_softmax_fwd.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_void_p]
    L.b200_causal_softmax_bwd.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_void_p]
    L.b200_causal_softmax_fwd.restype = L.b200_causal_softmax_bwd.restype = C.c_int
    L.b200_batchnorm_scratch_bytes.argtypes = [C.c_int]
    L.b200_batchnorm_scratch_bytes.restype = C.c_int64
    L.b200_batchnorm_fwd.argtypes = [C.c_void_p] * 10 + [C.c_int64, C.c_int, C.c_float, C.c_float, C.c_int,
                                                         C.c_int, C.c_void_p, C.c_void_p]
    L.b200_batchnorm_bwd.argtypes = [C.c_void_p] * 11 + [C.c_int, C.c_int64, C.c_int, C.c_int, C.c_void_p,
                                                          C.c_void_p]
    L.b200_batchnorm_eval_rstd.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_void_p]
    for name in ("b200_batchnorm_fwd", "b200_batchnorm_bwd", "b200_batchnorm_eval_rstd"):
        getattr(L, name).restype = C.c_int
    for name in ("b200_im2col", "b200_col2im"):
        getattr(L, name).argtypes = [C.c_void_p, C.c_void_p] + [C.c_int] * 8 + [C.c_void_p]
        getattr(L, name).restype = C.c_int
    for name in ("b200_layernorm_fwd", "b200_layernorm_bwd", "b200_colsum"):
        getattr(L, name).restype = C.c_int
    for name in ("b200_device_info", "b200_optimizer_step", "b200_cross_entropy", "b200_gemm",
                 "b200_cast_f32_to_bf16", "b200_cast_bf16_to_f32"):
        getattr(L, name).restype = C.c_int
    from . import bindings_comm
    bindings_comm.declare(L)


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def launch_count() -> int:
    return int(lib().b200_launch_count())


# ------------------------------------------------------------------ optimizer
def make_hparams(opt, grad_scale: float) -> OptHParams:
    hp = opt.hp
    return OptHParams(opt.kind_id, hp["lr"], hp["betas"][0], hp["betas"][1], hp["eps"],
                      hp["weight_decay"], hp["momentum"], grad_scale)


def optimizer_step(opt, start: int, end: int, grad_scale: float, *, device_step: bool = False,
                   advance: bool = False) -> None:
    """Fused update of arena slice [start, end).  Host-step mode uses opt.step_count (already
    advanced by the caller); device-step mode reads/advances opt.step_dev (CUDA-graph safe)."""
    hp = make_hparams(opt, grad_scale)
    n = end - start
    g = opt.grad[start:end]
    v = opt.exp_avg_sq[start:end] if opt.kind != "sgd" else None
    sh = opt.shadow[start:end] if opt.shadow is not None else None
    check(lib().b200_optimizer_step(
        C.byref(hp), opt.master[start:end].data_ptr(), g.data_ptr(),
        1 if g.dtype == torch.bfloat16 else 0, opt.exp_avg[start:end].data_ptr(), _ptr(v), _ptr(sh),
        n, opt.step_dev.data_ptr() if device_step else None, opt.step_count,
        1 if advance else 0, stream_ptr(opt.master.device)), "b200_optimizer_step")


# ------------------------------------------------------------------ loss
_ce_scratch = {}


def ce_scratch(device) -> torch.Tensor:
    key = torch.device(device).index
    if key not in _ce_scratch:
        nbytes = int(lib().b200_ce_scratch_bytes(1))
        _ce_scratch[key] = torch.zeros(nbytes, dtype=torch.uint8, device=device)
    return _ce_scratch[key]


def cross_entropy(logits: torch.Tensor, targets: torch.Tensor, dlogits: Optional[torch.Tensor],
                  grad_scale: float, loss_out: torch.Tensor) -> None:
    assert logits.dim() == 2 and logits.stride(1) == 1
    assert logits.dtype in (torch.bfloat16, torch.float32)
    assert targets.dtype == torch.int64 and targets.is_contiguous()
    if dlogits is not None:
        assert dlogits.dtype == logits.dtype and dlogits.stride() == logits.stride()
    check(lib().b200_cross_entropy(
        logits.data_ptr(), 1 if logits.dtype == torch.bfloat16 else 0, logits.stride(0),
        targets.data_ptr(), logits.shape[0], logits.shape[1], _ptr(dlogits), grad_scale,
        loss_out.data_ptr(), ce_scratch(logits.device).data_ptr(), stream_ptr(logits.device)),
        "b200_cross_entropy")


# ------------------------------------------------------------------ embedding
_emb_scratch = {}


def embedding_fwd(idx: torch.Tensor, wte: torch.Tensor, wpe: torch.Tensor) -> torch.Tensor:
    """x[b, t] = wte[idx[b, t]] + wpe[t]  (bf16 tables, int64 ids [B, T])."""
    assert idx.dtype == torch.int64 and idx.dim() == 2 and idx.is_contiguous()
    assert wte.dtype == torch.bfloat16 and wpe.dtype == torch.bfloat16 and wte.is_contiguous() and wpe.is_contiguous()
    Bsz, T = idx.shape
    V, Cc = wte.shape
    out = torch.empty(Bsz, T, Cc, dtype=torch.bfloat16, device=wte.device)
    check(lib().b200_embedding_fwd(idx.data_ptr(), wte.data_ptr(), wpe.data_ptr(), out.data_ptr(), Bsz * T, T, Cc,
                                   V, stream_ptr(wte.device)), "b200_embedding_fwd")
    return out


def embedding_bwd(idx: torch.Tensor, dx: torch.Tensor, g_wte: Optional[torch.Tensor],
                  g_wpe: Optional[torch.Tensor], V: int) -> None:
    """g_wte[v] += sum of dx over tokens with id v (deterministic); g_wpe[t] = sum_b dx[b, t]."""
    Bsz, T = idx.shape
    Cc = dx.shape[-1]
    assert dx.dtype == torch.bfloat16 and dx.is_contiguous() and idx.is_contiguous()
    for g in (g_wte, g_wpe):
        assert g is None or (g.is_contiguous() and g.dtype in (torch.bfloat16, torch.float32))
    key = (torch.device(dx.device).index, Bsz * T)
    sc = _emb_scratch.get(key)
    if sc is None:
        sc = _emb_scratch[key] = torch.empty(int(lib().b200_embedding_scratch_bytes(Bsz * T)), dtype=torch.uint8,
                                             device=dx.device)
    check(lib().b200_embedding_bwd(
        idx.data_ptr(), dx.data_ptr(), _ptr(g_wte), int(g_wte is not None and g_wte.dtype == torch.bfloat16),
        _ptr(g_wpe), int(g_wpe is not None and g_wpe.dtype == torch.bfloat16), Bsz, T,
        g_wpe.shape[0] if g_wpe is not None else T, Cc, V, sc.data_ptr(), stream_ptr(dx.device)),
        "b200_embedding_bwd")


# ------------------------------------------------------------------ GEMM
def gemm(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, *, M: int, N: int, K: int,
         a_mn_major: bool = False, b_mn_major: bool = False, bias: Optional[torch.Tensor] = None,
         epilogue: int = EPI_NONE, aux: Optional[torch.Tensor] = None,
         aux_out: Optional[torch.Tensor] = None, colsum: Optional[torch.Tensor] = None,
         dropout_p: float = 0.0, gate_scale: float = 1.0, seed: int = 0,
         step_dev: Optional[torch.Tensor] = None, layer_id: int = 0,
         targets: Optional[torch.Tensor] = None, ce_grad_scale: float = 1.0,
         loss_out: Optional[torch.Tensor] = None, loss_scale: float = 1.0,
         ce_write_grad: bool = True, tile_n: int = 0, max_ctas: int = 0, split_k: int = 0,
         batch=None, causal: int = 0, lda=None, ldb=None, ldd=None, tile_m: int = 0,
         cta_pair: int = 0, a_conv=None, b_conv=None) -> None:
    """out[M,N] = epilogue(A @ B^T).  A is [M,K] (K-major) or stored [K,M] (a_mn_major);
    B is [N,K] or stored [K,N] (b_mn_major).  Strides are taken from the tensors."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    assert a.stride(-1) == 1 and b.stride(-1) == 1 and out.stride(-1) == 1
    conv = a_conv if a_conv is not None else b_conv     # (R, S, stride, pad): the operand is an NHWC activation
    if conv is not None:
        act = a if a_conv is not None else b
        assert act.dim() == 4 and act.is_contiguous(), "implicit convolution wants a contiguous NHWC tensor"
    elif batch is None:
        assert a.dim() == 2 and b.dim() == 2 and out.dim() == 2
    assert out.dtype in (torch.bfloat16, torch.float32)
    if bias is not None:
        assert bias.dtype == torch.float32
    d = GemmDesc()
    d.M, d.N, d.K = M, N, K
    d.A, d.lda, d.a_mn_major = a.data_ptr(), (lda if lda is not None else a.stride(0)), int(a_mn_major)
    d.B, d.ldb, d.b_mn_major = b.data_ptr(), (ldb if ldb is not None else b.stride(0)), int(b_mn_major)
    if conv is not None:
        d.a_conv, d.b_conv = int(a_conv is not None), int(b_conv is not None)
        d.conv_n, d.conv_h, d.conv_w, d.conv_c = act.shape
        d.conv_r, d.conv_s, d.conv_stride, d.conv_pad = conv
        if a_conv is not None:
            d.lda = 0
        else:
            d.ldb = 0
    d.D, d.ldd, d.d_fp32 = out.data_ptr(), (ldd if ldd is not None else out.stride(0)), int(out.dtype == torch.float32)
    if batch is not None:
        # batch = (n_inner, n_outer, (a_si, a_so), (b_si, b_so), (d_si, d_so)), strides in elements
        d.batch_inner, d.batch_outer = batch[0], batch[1]
        (d.a_stride_inner, d.a_stride_outer), (d.b_stride_inner, d.b_stride_outer), \
            (d.d_stride_inner, d.d_stride_outer) = batch[2], batch[3], batch[4]
    d.causal = causal
    d.tile_m = tile_m
    d.cta_pair = cta_pair
    d.bias = _ptr(bias)
    d.epilogue = epilogue
    d.aux, d.ldaux = _ptr(aux), (aux.stride(0) if aux is not None else 0)
    d.aux_out = _ptr(aux_out)
    d.colsum = _ptr(colsum)
    d.dropout_p, d.gate_scale = dropout_p, gate_scale
    d.seed = seed & 0xFFFFFFFFFFFFFFFF
    d.step_dev = _ptr(step_dev)
    d.layer_id = layer_id
    d.targets = _ptr(targets)
    d.ce_grad_scale = ce_grad_scale
    d.loss_out = _ptr(loss_out)
    d.loss_scale = loss_scale
    d.ce_write_grad = int(ce_write_grad)
    d.tile_n, d.max_ctas, d.split_k = tile_n, max_ctas, split_k
    check(lib().b200_gemm(C.byref(d), stream_ptr(a.device)), "b200_gemm")


def cast_f32_to_bf16(src: torch.Tensor, dst: torch.Tensor) -> None:
    assert src.dtype == torch.float32 and dst.dtype == torch.bfloat16
    assert src.is_contiguous() and dst.is_contiguous() and src.numel() == dst.numel()
    check(lib().b200_cast_f32_to_bf16(src.data_ptr(), dst.data_ptr(), src.numel(),
                                      stream_ptr(src.device)), "b200_cast_f32_to_bf16")


GEMM_KERNELS = ("general", "lean", "m256", "pair")


def gemm_plan(M: int, N: int, K: int, *, sms: int = 148, d_fp32: bool = False, epilogue: int = EPI_NONE,
              bias: bool = False, aux: bool = False, aux_out: bool = False, colsum: bool = False, split_k: int = 0,
              tile_n: int = 0, tile_m: int = 0, cta_pair: int = 0, batch: int = 1, causal: int = 0, ldd=None,
              b_mn_major: bool = False):
    """(kernel family name, N tile) b200_gemm would pick — host logic only, works without a GPU."""
    d = GemmDesc()
    d.M, d.N, d.K = M, N, K
    fake = 0x10000   # only alignment and null-ness of the pointers enter the decision
    d.A = d.B = d.D = fake
    d.lda = d.ldb = (K + 7) // 8 * 8
    d.ldd = ldd if ldd is not None else N
    d.d_fp32, d.epilogue, d.b_mn_major = int(d_fp32), epilogue, int(b_mn_major)
    d.bias = fake if bias else None
    d.aux, d.ldaux = (fake, d.ldd) if aux else (None, 0)
    d.aux_out = fake if aux_out else None
    d.colsum = fake if colsum else None
    d.targets = fake if epilogue == EPI_SOFTMAX_CE else None
    d.split_k, d.tile_n, d.tile_m, d.cta_pair, d.causal = split_k, tile_n, tile_m, cta_pair, causal
    d.batch_inner = batch
    out = (C.c_int * 2)()
    check(lib().b200_gemm_plan(C.byref(d), sms, out), "b200_gemm_plan")
    return GEMM_KERNELS[out[0]], out[1]


def make_zero_list(tensors):
    """(ptr array, count array, n) for stage_batch's zero regions; build once, reuse."""
    n = len(tensors)
    ptrs = (C.c_void_p * max(n, 1))(*[t.data_ptr() for t in tensors])
    counts = (C.c_int * max(n, 1))(*[t.numel() for t in tensors])
    for t in tensors:
        assert t.dtype == torch.float32 and t.is_contiguous()
    return ptrs, counts, n


def stage_batch(x: torch.Tensor, y: torch.Tensor, x_bf16: torch.Tensor, y_out: torch.Tensor,
                zero_list=None) -> None:
    assert x.dtype == torch.float32 and x.is_contiguous() and y.dtype == torch.int64
    ptrs, counts, n = zero_list if zero_list is not None else (None, None, 0)
    check(lib().b200_stage_batch(x.data_ptr(), y.data_ptr(), x_bf16.data_ptr(), y_out.data_ptr(),
                                 x.numel(), y.numel(), ptrs, counts, n, stream_ptr(x.device)),
          "b200_stage_batch")
