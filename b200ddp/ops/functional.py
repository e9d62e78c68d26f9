"""Autograd faces of the sm_100a kernels, used by the transformer model.

Each function has two paths: on a B200 with bf16 tensors it runs our kernels
(``csrc/gemm.cu``, ``norm.cu``, ``loss.cu``) through custom ``autograd.Function``s;
anywhere else (CPU plumbing, fp32) it runs the equivalent torch ops, so models
built on it work under ``--backend gloo`` too.

Gradient routing: when the engine has placed a parameter in its flat arenas it
tags it with ``_b200_grad`` (bf16/fp32 view of the gradient arena),
``_b200_master`` (fp32 master view) and ``_b200_ready`` (bucket bookkeeping).
Backward then writes weight/bias gradients **straight into the arena** from the
wgrad GEMM / column-sum kernel — no ``.grad`` tensor, no accumulate pass — and
signals the bucket, which launches its fused all-reduce+optimizer kernel while
backward is still running.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch
import torch.nn.functional as F

from .. import _native as nat

_scratch = {}


def _B():
    from .._native import bindings
    return bindings


def native_ok(*tensors: torch.Tensor) -> bool:
    return all(t.is_cuda and t.dtype == torch.bfloat16 for t in tensors) and nat.available()


def _scratch_buf(device, nbytes: int) -> torch.Tensor:
    key = (torch.device(device).index, "s")
    buf = _scratch.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = _scratch[key] = torch.empty(max(nbytes, 16 << 20), dtype=torch.uint8, device=device)
    return buf


def _bias_f32(b: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    if b is None:
        return None
    m = getattr(b, "_b200_master", None)
    return m if m is not None else b.float()


def _colsum_into(x2: torch.Tensor, out: torch.Tensor) -> None:
    L = nat.lib()
    rows, cols = x2.shape
    sc = _scratch_buf(x2.device, int(L.b200_colsum_scratch_bytes(rows, cols)))
    nat.check(L.b200_colsum(x2.data_ptr(), x2.stride(0), rows, cols, out.data_ptr(),
                            int(out.dtype == torch.bfloat16), sc.data_ptr(),
                            nat.stream_ptr(x2.device)), "b200_colsum")


_side_streams = {}


class _AsyncBiasGrad:
    """Bias gradient (column sum of dy) on a side stream, so this DRAM-bound reduction overlaps
    the tensor-core-bound dgrad/wgrad GEMMs of the same layer.  `join()` re-joins the streams,
    signals the bucket and returns what autograd should see."""

    def __init__(self, b: torch.Tensor, dy2: torch.Tensor):
        dev = dy2.device
        key = dev.index
        if key not in _side_streams:
            _side_streams[key] = (torch.cuda.Stream(dev), {})
        self.side, scratch = _side_streams[key]
        self.b = b
        self.dst = getattr(b, "_b200_grad", None)
        self.owned = self.dst is not None
        if not self.owned:
            self.dst = torch.empty_like(b)
        L = nat.lib()
        rows, cols = dy2.shape
        need = int(L.b200_colsum_scratch_bytes(rows, cols))
        slot = scratch.get("buf")
        if slot is None or slot.numel() < need:
            slot = scratch["buf"] = torch.empty(max(need, 4 << 20), dtype=torch.uint8, device=dev)
        cur = torch.cuda.current_stream(dev)
        self.side.wait_stream(cur)
        dy2.record_stream(self.side)
        with torch.cuda.stream(self.side):
            nat.check(L.b200_colsum(dy2.data_ptr(), dy2.stride(0), rows, cols, self.dst.data_ptr(),
                                    int(self.dst.dtype == torch.bfloat16), slot.data_ptr(),
                                    nat.stream_ptr(dev)), "b200_colsum")

    def join(self):
        torch.cuda.current_stream(self.dst.device).wait_stream(self.side)
        if self.owned:
            self.b._b200_ready()
            return None
        return self.dst


def _emit_param_grad(p: torch.Tensor, compute_into, signal: bool = True):
    """Runs compute_into(dst) with dst = the arena view when the engine owns p (returns None so
    autograd does not accumulate), else into a fresh tensor that is returned to autograd."""
    dst = getattr(p, "_b200_grad", None)
    if dst is not None:
        compute_into(dst)
        if signal:
            p._b200_ready()
        return None
    g = torch.empty_like(p)
    compute_into(g)
    return g


def wgrad_plan(n_out: int, n_in: int, K: int, bf16_dst: bool, row_stride_ok: bool = True) -> str:
    """How dW[n_out, n_in] = dy^T @ x (reduction over K rows) fills 148 SMs when the output has few
    tiles; constants fitted to B200 measurements (tools/gpt2_gemm_shapes.py):

    direct   the two-SM kernel with 256x128 pair tiles writing bf16 straight into the arena:
             ~0.28 us per k-block, one wave per 74 pair tiles
    split    K slices across the idle SMs, fp32 partials combined with vector atomics into a
             zeroed temp, then one cast: ~0.45 us per k-block spread over all SMs plus ~30 us of
             memset / atomic / cast overhead
    plain    whatever b200_gemm picks (enough tiles, or a short reduction)
    """
    kb = (K + 63) // 64
    tiles = ((n_out + 127) // 128) * ((n_in + 255) // 256)
    direct_ok = bf16_dst and n_out > 128 and n_in > 64 and row_stride_ok
    pair_tiles = ((n_out + 255) // 256) * ((n_in + 127) // 128)
    direct_us = -(-pair_tiles // 74) * kb * 0.28 + 3 if direct_ok else float("inf")
    split_us = tiles * kb * 0.45 / 148 + 30 if (tiles < 100 and K >= 2048) else float("inf")
    if split_us < direct_us:
        return "split"
    return "direct" if direct_ok else "plain"


def _wgrad_into(dy2: torch.Tensor, x2: torch.Tensor, dst: torch.Tensor) -> None:
    # dW[n_out, n_in] = dy^T @ x : both operands read MN-major, nothing is transposed
    n_out, n_in = dst.shape
    K = x2.shape[0]
    if wgrad_plan(n_out, n_in, K, dst.dtype == torch.bfloat16, dst.stride(0) % 8 == 0) == "split":
        tmp = torch.zeros(n_out, n_in, dtype=torch.float32, device=dst.device)
        _B().gemm(dy2, x2, tmp, M=n_out, N=n_in, K=K, a_mn_major=True, b_mn_major=True, split_k=1)
        dst.copy_(tmp)
        return
    _B().gemm(dy2, x2, dst, M=n_out, N=n_in, K=K, a_mn_major=True, b_mn_major=True)


# ------------------------------------------------------------------ linear (+bias, +residual)
class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, residual):
        B = _B()
        K = w.shape[1]
        x2 = x.reshape(-1, K)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        M, N = x2.shape[0], w.shape[0]
        out = torch.empty(M, N, dtype=torch.bfloat16, device=x.device)
        r2 = residual.reshape(M, N) if residual is not None else None
        B.gemm(x2, w, out, M=M, N=N, K=K, bias=_bias_f32(b), epilogue=B.EPI_NONE, aux=r2)
        ctx.save_for_backward(x2)
        ctx.w, ctx.b, ctx.has_res = w, b, residual is not None
        ctx.x_shape = x.shape
        return out.view(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        B = _B()
        (x2,) = ctx.saved_tensors
        w, b = ctx.w, ctx.b
        N, K = w.shape
        dy2 = dy.reshape(-1, N)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        M = dy2.shape[0]
        bias_job = _AsyncBiasGrad(b, dy2) if b is not None else None   # overlaps the GEMMs below
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty(M, K, dtype=torch.bfloat16, device=dy.device)
            B.gemm(dy2, w, dx, M=M, N=K, K=N, b_mn_major=True)  # W stored [N_out, K_in] = [K, N]
            dx = dx.view(ctx.x_shape)
        gw = _emit_param_grad(w, lambda dst: _wgrad_into(dy2, x2, dst))
        gb = bias_job.join() if bias_job is not None else None
        return dx, gw, gb, (dy if ctx.has_res else None)


def _tma_dims_ok(*ws: torch.Tensor) -> bool:
    """Every weight dimension becomes a TMA row pitch in forward, dgrad or wgrad (16-byte rule)."""
    return all(d % 8 == 0 for w in ws for d in w.shape)


def linear(x, w, b=None, residual=None):
    if native_ok(x, w) and _tma_dims_ok(w):
        return _Linear.apply(x, w, b, residual)
    y = F.linear(x, w, b)
    return y + residual if residual is not None else y


# ------------------------------------------------------------------ transformer MLP block
class _MLP(torch.autograd.Function):
    """y = gelu(x Wfc^T + bfc) Wproj^T + bproj + residual, GELU and its backward gate fused
    into GEMM epilogues (the pre-activation is written by the first GEMM's epilogue)."""

    @staticmethod
    def forward(ctx, x, w_fc, b_fc, w_proj, b_proj, residual):
        B = _B()
        C_in = w_fc.shape[1]
        x2 = x.reshape(-1, C_in)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        M, H, C_out = x2.shape[0], w_fc.shape[0], w_proj.shape[0]
        z = torch.empty(M, H, dtype=torch.bfloat16, device=x.device)
        h = torch.empty(M, H, dtype=torch.bfloat16, device=x.device)
        B.gemm(x2, w_fc, h, M=M, N=H, K=C_in, bias=_bias_f32(b_fc), epilogue=B.EPI_GELU, aux_out=z)
        y = torch.empty(M, C_out, dtype=torch.bfloat16, device=x.device)
        r2 = residual.reshape(M, C_out) if residual is not None else None
        B.gemm(h, w_proj, y, M=M, N=C_out, K=H, bias=_bias_f32(b_proj), epilogue=B.EPI_NONE, aux=r2)
        ctx.save_for_backward(x2, z, h)
        ctx.params = (w_fc, b_fc, w_proj, b_proj)
        ctx.has_res = residual is not None
        ctx.x_shape = x.shape
        return y.view(*x.shape[:-1], C_out)

    @staticmethod
    def backward(ctx, dy):
        B = _B()
        x2, z, h = ctx.saved_tensors
        w_fc, b_fc, w_proj, b_proj = ctx.params
        H, C_in = w_fc.shape
        C_out = w_proj.shape[0]
        dy2 = dy.reshape(-1, C_out)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        M = dy2.shape[0]
        bp_job = _AsyncBiasGrad(b_proj, dy2)   # side stream: overlaps the GEMMs below
        # dz = (dy @ Wproj) * gelu'(z): dgrad GEMM with the gate in its epilogue
        dz = torch.empty(M, H, dtype=torch.bfloat16, device=dy.device)
        B.gemm(dy2, w_proj, dz, M=M, N=H, K=C_out, b_mn_major=True, epilogue=B.EPI_GELU_BWD, aux=z)
        bf_job = _AsyncBiasGrad(b_fc, dz)
        g_wp = _emit_param_grad(w_proj, lambda dst: _wgrad_into(dy2, h, dst))
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty(M, C_in, dtype=torch.bfloat16, device=dy.device)
            B.gemm(dz, w_fc, dx, M=M, N=C_in, K=H, b_mn_major=True)
            dx = dx.view(ctx.x_shape)
        g_wf = _emit_param_grad(w_fc, lambda dst: _wgrad_into(dz, x2, dst))
        g_bp = bp_job.join()
        g_bf = bf_job.join()
        return dx, g_wf, g_bf, g_wp, g_bp, (dy if ctx.has_res else None)


def mlp(x, w_fc, b_fc, w_proj, b_proj, residual=None):
    if native_ok(x, w_fc, w_proj) and _tma_dims_ok(w_fc, w_proj):
        return _MLP.apply(x, w_fc, b_fc, w_proj, b_proj, residual)
    y = F.linear(F.gelu(F.linear(x, w_fc, b_fc), approximate="tanh"), w_proj, b_proj)
    return y + residual if residual is not None else y


# ------------------------------------------------------------------ layer norm
class _LayerNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, eps):
        L = nat.lib()
        Cc = x.shape[-1]
        x2 = x.reshape(-1, Cc)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        rows = x2.shape[0]
        y = torch.empty_like(x2)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        nat.check(L.b200_layernorm_fwd(x2.data_ptr(), gamma.data_ptr(), beta.data_ptr(), y.data_ptr(),
                                       mean.data_ptr(), rstd.data_ptr(), rows, Cc, eps,
                                       nat.stream_ptr(x.device)), "b200_layernorm_fwd")
        ctx.save_for_backward(x2, mean, rstd)
        ctx.gamma, ctx.beta = gamma, beta
        ctx.x_shape = x.shape
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        L = nat.lib()
        x2, mean, rstd = ctx.saved_tensors
        gamma, beta = ctx.gamma, ctx.beta
        rows, Cc = x2.shape
        dy2 = dy.reshape(rows, Cc)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = torch.empty_like(x2)
        sc = _scratch_buf(dy.device, int(L.b200_layernorm_bwd_scratch_bytes(rows, Cc)))
        gdst = getattr(gamma, "_b200_grad", None)
        bdst = getattr(beta, "_b200_grad", None)
        owned = gdst is not None and bdst is not None
        if not owned:
            gdst, bdst = torch.empty_like(gamma), torch.empty_like(beta)
        assert gdst.dtype == bdst.dtype
        nat.check(L.b200_layernorm_bwd(dy2.data_ptr(), x2.data_ptr(), gamma.data_ptr(), mean.data_ptr(),
                                       rstd.data_ptr(), dx.data_ptr(), None, gdst.data_ptr(),
                                       bdst.data_ptr(), int(gdst.dtype == torch.bfloat16), rows, Cc,
                                       sc.data_ptr(), nat.stream_ptr(dy.device)), "b200_layernorm_bwd")
        if owned:
            gamma._b200_ready()
            beta._b200_ready()
            return dx.view(ctx.x_shape), None, None, None
        return dx.view(ctx.x_shape), gdst, bdst, None


def layer_norm(x, gamma, beta, eps: float = 1e-5):
    Cc = x.shape[-1]
    if native_ok(x, gamma, beta) and Cc % 256 == 0 and 256 <= Cc <= 2048:
        return _LayerNorm.apply(x, gamma, beta, eps)
    return F.layer_norm(x, (Cc,), gamma, beta, eps)


# ------------------------------------------------------------------ tied LM head + cross entropy
class _LMHeadCE(torch.autograd.Function):
    """loss = mean CE(x @ W^T, targets).  Logits live once, in bf16; the CE kernel overwrites
    them with dlogits; dgrad and wgrad read that buffer (MN-major for wgrad)."""

    @staticmethod
    def forward(ctx, x, w, targets, signal_weight):
        B = _B()
        Cc = w.shape[1]
        V = w.shape[0]
        x2 = x.reshape(-1, Cc)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        M = x2.shape[0]
        ld = (V + 7) // 8 * 8
        logits = torch.empty(M, ld, dtype=torch.bfloat16, device=x.device)
        B.gemm(x2, w, logits, M=M, N=V, K=Cc)
        out = torch.empty(2, dtype=torch.float32, device=x.device)
        lg = logits[:, :V]
        B.cross_entropy(lg, targets.reshape(-1).contiguous(), lg, 1.0 / M, out)
        ctx.save_for_backward(x2, logits)
        ctx.w, ctx.V, ctx.signal = w, V, signal_weight
        ctx.x_shape = x.shape
        return out[0]

    @staticmethod
    def backward(ctx, gout):
        B = _B()
        x2, dlogits = ctx.saved_tensors
        w, V = ctx.w, ctx.V
        M, Cc = x2.shape
        dl = dlogits[:, :V]
        dx = torch.empty(M, Cc, dtype=torch.bfloat16, device=x2.device)
        B.gemm(dl, w, dx, M=M, N=Cc, K=V, b_mn_major=True)
        gw = _emit_param_grad(w, lambda dst: _wgrad_into(dl, x2, dst), signal=ctx.signal)
        # the kernels assume the loss is the root of backward (seed 1.0); honour other seeds
        dx = dx.view(ctx.x_shape) * gout.to(dx.dtype)
        if gw is not None:
            gw = gw * gout.to(gw.dtype)
        return dx, gw, None, None


def lm_head_cross_entropy(x, w, targets, signal_weight: bool = True):
    """Mean token cross-entropy of the (tied) LM head.  `signal_weight=False` when `w` receives
    another gradient contribution later in backward (tied embedding)."""
    if native_ok(x, w):
        return _LMHeadCE.apply(x, w, targets, signal_weight)
    logits = F.linear(x, w).float()
    return F.cross_entropy(logits.view(-1, logits.shape[-1]), targets.reshape(-1))


# ------------------------------------------------------------------ embedding (tied with the head)
class _Embedding(torch.autograd.Function):
    """Token + position embedding on our kernels (csrc/embedding.cu): a gather forward and a
    deterministic (sorted, atomic-free) scatter-add backward that lands on top of the tied LM-head
    gradient already in the arena."""

    @staticmethod
    def forward(ctx, idx, wte, wpe):
        idx = idx.contiguous()
        x = _B().embedding_fwd(idx, wte, wpe)
        ctx.save_for_backward(idx)
        ctx.wte, ctx.wpe = wte, wpe
        return x

    @staticmethod
    def backward(ctx, dx):
        B = _B()
        (idx,) = ctx.saved_tensors
        wte, wpe = ctx.wte, ctx.wpe
        V = wte.shape[0]
        dx = dx.contiguous()

        def wte_grad(dst):  # accumulates on top of the LM-head contribution already in the arena
            B.embedding_bwd(idx, dx, dst, None, V)

        def wpe_grad(dst):
            B.embedding_bwd(idx, dx, None, dst, V)

        owned = getattr(wte, "_b200_grad", None) is not None
        if owned:
            g_wte = _emit_param_grad(wte, wte_grad)
        else:
            g_wte = torch.zeros_like(wte)
            wte_grad(g_wte)
        g_wpe = _emit_param_grad(wpe, wpe_grad)
        return None, g_wte, g_wpe


def embedding(idx, wte, wpe):
    if native_ok(wte, wpe):
        return _Embedding.apply(idx, wte, wpe)
    T = idx.shape[1]
    return F.embedding(idx, wte) + wpe[:T].unsqueeze(0)


# ------------------------------------------------------------------ attention
class _CausalAttention(torch.autograd.Function):
    """softmax(Q K^T / sqrt(d)) V with a causal mask, from our own kernels only: batched causal
    tcgen05 GEMMs (tiles and K ranges outside the triangle are skipped) around the row-softmax
    kernels of csrc/attention.cu.  Q, K, V are read in place from the packed qkv activation and
    dQ, dK, dV are written in place into one packed gradient — no head split/merge copies."""

    @staticmethod
    def forward(ctx, qkv, n_head):
        B = _B()
        L = nat.lib()
        Bs, T, C3 = qkv.shape
        Cc = C3 // 3
        D = Cc // n_head
        if not qkv.is_contiguous():
            qkv = qkv.contiguous()
        q3 = qkv.view(Bs, T, 3, Cc)
        q, k, v = q3[:, :, 0], q3[:, :, 1], q3[:, :, 2]      # views: data_ptr carries the offset
        BH = Bs * n_head
        qkv_b = (D, T * C3)                                   # (head stride, sequence stride)
        s_b = (T * T, n_head * T * T)
        scale = 1.0 / (D ** 0.5)
        S = torch.empty(BH, T, T, dtype=torch.bfloat16, device=qkv.device)
        B.gemm(q, k, S, M=T, N=T, K=D, lda=C3, ldb=C3, ldd=T, causal=1,
               batch=(n_head, Bs, qkv_b, qkv_b, s_b))
        nat.check(L.b200_causal_softmax_fwd(S.data_ptr(), BH, T, scale, nat.stream_ptr(qkv.device)),
                  "b200_causal_softmax_fwd")
        out = torch.empty(Bs, T, Cc, dtype=torch.bfloat16, device=qkv.device)
        B.gemm(S, v, out, M=T, N=D, K=T, lda=T, ldb=C3, ldd=Cc, b_mn_major=True, causal=2,
               batch=(n_head, Bs, s_b, qkv_b, (D, T * Cc)))
        ctx.save_for_backward(qkv, S)
        ctx.n_head = n_head
        return out

    @staticmethod
    def backward(ctx, dout):
        B = _B()
        L = nat.lib()
        qkv, P = ctx.saved_tensors
        n_head = ctx.n_head
        Bs, T, C3 = qkv.shape
        Cc = C3 // 3
        D = Cc // n_head
        BH = Bs * n_head
        if not dout.is_contiguous():
            dout = dout.contiguous()
        q3 = qkv.view(Bs, T, 3, Cc)
        q, k, v = q3[:, :, 0], q3[:, :, 1], q3[:, :, 2]
        dqkv = torch.empty_like(qkv)
        d3 = dqkv.view(Bs, T, 3, Cc)
        dq, dk, dv = d3[:, :, 0], d3[:, :, 1], d3[:, :, 2]
        qkv_b, s_b, o_b = (D, T * C3), (T * T, n_head * T * T), (D, T * Cc)
        scale = 1.0 / (D ** 0.5)
        # dP = dO V^T   (lower triangle only)
        dP = torch.empty(BH, T, T, dtype=torch.bfloat16, device=qkv.device)
        B.gemm(dout, v, dP, M=T, N=T, K=D, lda=Cc, ldb=C3, ldd=T, causal=1,
               batch=(n_head, Bs, o_b, qkv_b, s_b))
        # dV = P^T dO   (queries >= key tile start)
        B.gemm(P, dout, dv, M=T, N=D, K=T, lda=T, ldb=Cc, ldd=C3, a_mn_major=True, b_mn_major=True,
               causal=3, batch=(n_head, Bs, s_b, o_b, qkv_b))
        nat.check(L.b200_causal_softmax_bwd(P.data_ptr(), dP.data_ptr(), BH, T, scale,
                                            nat.stream_ptr(qkv.device)), "b200_causal_softmax_bwd")
        dS = dP
        # dQ = dS K     (keys <= query tile end)
        B.gemm(dS, k, dq, M=T, N=D, K=T, lda=T, ldb=C3, ldd=C3, b_mn_major=True, causal=2,
               batch=(n_head, Bs, s_b, qkv_b, qkv_b))
        # dK = dS^T Q   (queries >= key tile start)
        B.gemm(dS, q, dk, M=T, N=D, K=T, lda=T, ldb=C3, ldd=C3, a_mn_major=True, b_mn_major=True,
               causal=3, batch=(n_head, Bs, s_b, qkv_b, qkv_b))
        return dqkv, None


def flash_attention_forward(qkv: torch.Tensor, n_head: int):
    """Fused causal attention forward (csrc/flash_attention.cu): returns (out [B,T,C], lse [B,H,T])."""
    Bs, T, C3 = qkv.shape
    Cc = C3 // 3
    if not qkv.is_contiguous():
        qkv = qkv.contiguous()
    out = torch.empty(Bs, T, Cc, dtype=torch.bfloat16, device=qkv.device)
    lse = torch.empty(Bs, n_head, T, dtype=torch.float32, device=qkv.device)
    nat.check(nat.lib().b200_flash_attn_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), Bs, T, n_head,
                                            Cc // n_head, nat.stream_ptr(qkv.device)), "b200_flash_attn_fwd")
    return out, lse


class _FlashAttention(torch.autograd.Function):
    """Fused causal attention, forward and backward (csrc/flash_attention.cu): scores and
    probabilities never leave TMEM / shared memory; only O and the per-row log-sum-exp are kept."""

    @staticmethod
    def forward(ctx, qkv, n_head):
        if not qkv.is_contiguous():
            qkv = qkv.contiguous()
        out, lse = flash_attention_forward(qkv, n_head)
        ctx.save_for_backward(qkv, out, lse)
        ctx.n_head = n_head
        return out

    @staticmethod
    def backward(ctx, dout):
        qkv, out, lse = ctx.saved_tensors
        H = ctx.n_head
        Bs, T, C3 = qkv.shape
        D = C3 // 3 // H
        if not dout.is_contiguous():
            dout = dout.contiguous()
        dqkv = torch.empty_like(qkv)
        dq_acc = torch.empty(Bs, H, T, D, dtype=torch.float32, device=qkv.device)
        delta = torch.empty(Bs, H, T, dtype=torch.float32, device=qkv.device)
        nat.check(nat.lib().b200_flash_attn_bwd(
            qkv.data_ptr(), out.data_ptr(), dout.data_ptr(), lse.data_ptr(), dqkv.data_ptr(),
            dq_acc.data_ptr(), delta.data_ptr(), Bs, T, H, D, nat.stream_ptr(qkv.device)),
            "b200_flash_attn_bwd")
        return dqkv, None


def causal_attention(qkv: torch.Tensor, n_head: int) -> torch.Tensor:
    """qkv [B, T, 3C] -> [B, T, C].  On a B200: the fused flash-style tcgen05 kernels (head dim 64),
    or the batched-GEMM + row-softmax path (B200DDP_ATTENTION=gemm, other head dims); both are
    ours.  B200DDP_ATTENTION=sdpa selects torch's library kernel for comparison; torch SDPA on CPU."""
    Bsz, T, C3 = qkv.shape
    Cc = C3 // 3
    D = Cc // n_head
    mode = os.environ.get("B200DDP_ATTENTION", "flash")   # flash | gemm | sdpa
    if native_ok(qkv) and D % 64 == 0 and T % 128 == 0 and mode != "sdpa":
        if D == 64 and mode == "flash":
            if not (torch.is_grad_enabled() and qkv.requires_grad):
                return flash_attention_forward(qkv, n_head)[0]
            return _FlashAttention.apply(qkv, n_head)
        return _CausalAttention.apply(qkv, n_head)   # batched causal GEMMs + row softmax kernels
    q, k, v = qkv.view(Bsz, T, 3, n_head, D).permute(2, 0, 3, 1, 4)
    y = F.scaled_dot_product_attention(q, k, v, is_causal=True)
    return y.transpose(1, 2).reshape(Bsz, T, Cc)


# ------------------------------------------------------------------ stand-alone cross entropy
class _CrossEntropy(torch.autograd.Function):
    """Mean CE over rows with our fused kernel: the forward pass already produces dlogits."""

    @staticmethod
    def forward(ctx, logits, targets):
        lg = logits if logits.stride(-1) == 1 else logits.contiguous()
        dl = torch.empty_like(lg)
        out = torch.empty(2, dtype=torch.float32, device=lg.device)
        _B().cross_entropy(lg, targets.contiguous(), dl, 1.0 / lg.shape[0], out)
        ctx.save_for_backward(dl)
        return out[0]

    @staticmethod
    def backward(ctx, gout):
        (dl,) = ctx.saved_tensors
        return dl * gout.to(dl.dtype), None


def cross_entropy(logits, targets):
    """F.cross_entropy(mean) for [rows, classes] logits; fused fwd+bwd kernel on a B200."""
    if (logits.is_cuda and logits.dim() == 2 and logits.dtype in (torch.bfloat16, torch.float32)
            and targets.dtype == torch.int64 and nat.available()):
        return _CrossEntropy.apply(logits, targets)
    return F.cross_entropy(logits.float(), targets)


# ------------------------------------------------------------------ convolution (ResNet-50)
def _im2col(x_nhwc: torch.Tensor, R: int, S: int, stride: int, pad: int) -> torch.Tensor:
    N, H, W, Cc = x_nhwc.shape
    P = (H + 2 * pad - R) // stride + 1
    Q = (W + 2 * pad - S) // stride + 1
    col = torch.empty(N * P * Q, R * S * Cc, dtype=torch.bfloat16, device=x_nhwc.device)
    nat.check(nat.lib().b200_im2col(x_nhwc.data_ptr(), col.data_ptr(), N, H, W, Cc, R, S, stride, pad,
                                    nat.stream_ptr(x_nhwc.device)), "b200_im2col")
    return col


class _Conv2d(torch.autograd.Function):
    """Convolution on the tcgen05 GEMM.  `x` is a logical-NCHW tensor in channels-last memory
    (i.e. NHWC storage); weights keep the torch [Cout, Cin, R, S] layout in the arenas and are
    re-laid to [Cout, R*S*Cin] per call (a few tens of MB per step for ResNet-50)."""

    @staticmethod
    def forward(ctx, x, w, stride, pad):
        B = _B()
        Cout, Cin, R, S = w.shape
        xh = x.permute(0, 2, 3, 1)
        if not xh.is_contiguous():
            xh = xh.contiguous()
        cpad = (-Cin) % 8
        if cpad:  # stem: 3 input channels -> 8
            xh = F.pad(xh, (0, cpad))
        w2 = w.permute(0, 2, 3, 1)
        if cpad:
            w2 = F.pad(w2, (0, cpad))
        w2 = w2.reshape(Cout, -1).contiguous()
        N, H, W_, Cp = xh.shape
        direct = R == 1 and S == 1 and stride == 1 and pad == 0
        # Implicit GEMM: every other convolution whose channel count fills 64-wide k-blocks reads the
        # activation through a TMA im2col-mode tensor map — forward and weight gradient — so no
        # gathered [N*P*Q, R*S*C] matrix is ever written (round 1: 1.0 ms of im2col per step and
        # 2.2 GB of saved operands at batch 64).  The 3-channel stem keeps the explicit gather.
        implicit = (not direct) and Cp % 64 == 0 and os.environ.get("B200DDP_NO_IMPLICIT_CONV") is None
        P = (H + 2 * pad - R) // stride + 1
        Q = (W_ + 2 * pad - S) // stride + 1
        y = torch.empty(N * P * Q, Cout, dtype=torch.bfloat16, device=x.device)
        if implicit:
            a = xh
            B.gemm(xh, w2, y, M=N * P * Q, N=Cout, K=w2.shape[1], a_conv=(R, S, stride, pad))
        else:
            a = xh.view(N * H * W_, Cp) if direct else _im2col(xh, R, S, stride, pad)
            B.gemm(a, w2, y, M=N * P * Q, N=Cout, K=w2.shape[1])
        # explicit path: the gathered operand is kept for the weight gradient
        ctx.save_for_backward(xh, w2, a)
        ctx.w, ctx.geom = w, (R, S, stride, pad, direct, cpad, Cin, P, Q)
        ctx.implicit = implicit
        return y.view(N, P, Q, Cout).permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy):
        B = _B()
        xh, w2, a = ctx.saved_tensors
        w = ctx.w
        R, S, stride, pad, direct, cpad, Cin, P, Q = ctx.geom
        N, H, W_, Cp = xh.shape
        Cout, Kt = w2.shape
        M = N * P * Q
        dyh = dy.permute(0, 2, 3, 1)
        if not dyh.is_contiguous():
            dyh = dyh.contiguous()
        dy2 = dyh.view(M, Cout)
        dx = None
        if ctx.needs_input_grad[0] and ctx.implicit and stride == 1 and Cout % 64 == 0 and not cpad:
            # Data gradient of a stride-1 convolution = convolution of dY with the filter mirrored in
            # (r, s) and transposed in (Cout, Cin), padding R-1-pad: the same implicit-GEMM forward
            # kernel, no [M, R*S*C] column matrix and no col2im pass.
            wf = w.flip(2, 3).permute(1, 2, 3, 0).reshape(Cin, R * S * Cout).contiguous()
            dxh = torch.empty(N * H * W_, Cin, dtype=torch.bfloat16, device=dy.device)
            B.gemm(dyh, wf, dxh, M=N * H * W_, N=Cin, K=R * S * Cout, a_conv=(R, S, 1, R - 1 - pad))
            dx = dxh.view(N, H, W_, Cin).permute(0, 3, 1, 2)
        elif ctx.needs_input_grad[0]:
            dcol = torch.empty(M, Kt, dtype=torch.bfloat16, device=dy.device)
            B.gemm(dy2, w2, dcol, M=M, N=Kt, K=Cout, b_mn_major=True)  # W2 stored [Cout, Kt] = [K, N]
            if direct:
                dxh = dcol.view(N, H, W_, Cp)
            else:
                dxh = torch.empty(N, H, W_, Cp, dtype=torch.bfloat16, device=dy.device)
                nat.check(nat.lib().b200_col2im(dcol.data_ptr(), dxh.data_ptr(), N, H, W_, Cp, R, S, stride,
                                                pad, nat.stream_ptr(dy.device)), "b200_col2im")
            if cpad:
                dxh = dxh[..., :Cin]
            dx = dxh.permute(0, 3, 1, 2)
        # reduction over all N*P*Q pixels into a small [Cout, R*S*Cin] output: split-K across the SMs,
        # or the two-SM kernel straight to bf16 when the output has enough tiles (_wgrad_into)
        garena = getattr(w, "_b200_grad", None)
        dw2 = torch.empty(Cout, Kt, dtype=garena.dtype if garena is not None else torch.bfloat16, device=dy.device)
        if ctx.implicit:
            # dW = dY^T @ im2col(x) with the im2col view as MN-major B operand; the output is small
            # ([Cout, R*S*C]) and the reduction long (all pixels): split-K into an fp32 buffer
            tmp = torch.zeros(Cout, Kt, dtype=torch.float32, device=dy.device)
            B.gemm(dy2, xh, tmp, M=Cout, N=Kt, K=M, a_mn_major=True, b_conv=(R, S, stride, pad), split_k=1)
            dw2.copy_(tmp)
        else:
            _wgrad_into(dy2, a, dw2)
        dw = dw2.view(Cout, R, S, Cp)[..., :Cin].permute(0, 3, 1, 2)  # back to [Cout, Cin, R, S]

        def emit(dst):
            dst.copy_(dw)
        gw = _emit_param_grad(w, emit)
        return dx, gw, None, None


def conv2d(x, w, stride: int = 1, pad: int = 0):
    """Bias-free 2-D convolution (square stride/padding).  tcgen05 GEMM path for bf16 CUDA
    tensors with Cout a multiple of 8; torch otherwise."""
    if native_ok(x, w) and w.shape[0] % 8 == 0:
        return _Conv2d.apply(x, w, stride, pad)
    return F.conv2d(x, w, None, stride, pad)


# ------------------------------------------------------------------ batch norm (+ReLU, +residual)
def _f32_param(p: torch.Tensor) -> torch.Tensor:
    m = getattr(p, "_b200_master", None)
    return m if m is not None else p.float()


class _BatchNormAct(torch.autograd.Function):
    """y = relu?(BN(x) (+ residual)) on channels-last bf16 activations: 2 passes forward,
    2 backward (csrc/batchnorm.cu); running statistics are updated in the statistics pass."""

    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, training, momentum, eps, relu, residual):
        L = nat.lib()
        xh = x.permute(0, 2, 3, 1)
        if not xh.is_contiguous():
            xh = xh.contiguous()
        N, H, W, Cc = xh.shape
        rows = N * H * W
        rh = None
        if residual is not None:
            rh = residual.permute(0, 2, 3, 1)
            if not rh.is_contiguous():
                rh = rh.contiguous()
        y = torch.empty_like(xh)
        gamma, beta = _f32_param(weight), _f32_param(bias)
        sc = _scratch_buf(x.device, int(L.b200_batchnorm_scratch_bytes(Cc)))
        if training:
            mean = torch.empty(Cc, dtype=torch.float32, device=x.device)
            rstd = torch.empty(Cc, dtype=torch.float32, device=x.device)
        else:
            mean = running_mean
            rstd = torch.empty(Cc, dtype=torch.float32, device=x.device)
            nat.check(L.b200_batchnorm_eval_rstd(running_var.data_ptr(), rstd.data_ptr(), Cc, eps,
                                                 nat.stream_ptr(x.device)), "b200_batchnorm_eval_rstd")
        mask = None
        if relu and training and any(ctx.needs_input_grad[:3]):
            mask = torch.empty(rows * (Cc // 8), dtype=torch.uint8, device=x.device)
        nat.check(L.b200_batchnorm_fwd(
            xh.data_ptr(), rh.data_ptr() if rh is not None else None, y.data_ptr(),
            mask.data_ptr() if mask is not None else None, gamma.data_ptr(),
            beta.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
            running_mean.data_ptr() if training else None, running_var.data_ptr() if training else None,
            rows, Cc, eps, momentum, int(training), int(relu), sc.data_ptr(), nat.stream_ptr(x.device)),
            "b200_batchnorm_fwd")
        ctx.save_for_backward(xh, mask if mask is not None else (y if relu else xh), mean, rstd)
        ctx.has_mask = mask is not None
        ctx.weight, ctx.bias = weight, bias
        ctx.relu, ctx.has_res = relu, residual is not None
        return y.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy):
        L = nat.lib()
        xh, y, mean, rstd = ctx.saved_tensors
        weight, bias = ctx.weight, ctx.bias
        N, H, W, Cc = xh.shape
        rows = N * H * W
        dyh = dy.permute(0, 2, 3, 1)
        if not dyh.is_contiguous():
            dyh = dyh.contiguous()
        dx = torch.empty_like(xh)
        dres = torch.empty_like(xh) if ctx.has_res else None
        gdst, bdst = getattr(weight, "_b200_grad", None), getattr(bias, "_b200_grad", None)
        owned = gdst is not None and bdst is not None
        if not owned:
            gdst, bdst = torch.empty_like(weight), torch.empty_like(bias)
        sc = _scratch_buf(dy.device, int(L.b200_batchnorm_scratch_bytes(Cc)))
        nat.check(L.b200_batchnorm_bwd(
            dyh.data_ptr(), xh.data_ptr(), None if ctx.has_mask else y.data_ptr(),
            y.data_ptr() if ctx.has_mask else None, mean.data_ptr(), rstd.data_ptr(),
            _f32_param(weight).data_ptr(), dx.data_ptr(), dres.data_ptr() if dres is not None else None,
            gdst.data_ptr(), bdst.data_ptr(), int(gdst.dtype == torch.bfloat16), rows, Cc, int(ctx.relu),
            sc.data_ptr(), nat.stream_ptr(dy.device)), "b200_batchnorm_bwd")
        if owned:
            weight._b200_ready()
            bias._b200_ready()
            gw = gb = None
        else:
            gw, gb = gdst, bdst
        return (dx.permute(0, 3, 1, 2), gw, gb, None, None, None, None, None, None,
                dres.permute(0, 3, 1, 2) if dres is not None else None)


def batch_norm_act(x, weight, bias, running_mean, running_var, training: bool, momentum: float, eps: float,
                   relu: bool = False, residual=None):
    """BatchNorm2d on a logical-NCHW / channels-last tensor, optionally fused with a residual
    add and ReLU.  Own kernels for bf16 CUDA tensors with power-of-two channels, torch otherwise."""
    Cc = x.shape[1]
    if (x.is_cuda and x.dtype == torch.bfloat16 and nat.available() and 8 <= Cc <= 2048
            and (Cc & (Cc - 1)) == 0 and running_mean is not None and running_mean.dtype == torch.float32):
        return _BatchNormAct.apply(x, weight, bias, running_mean, running_var, training, momentum, eps,
                                   relu, residual)
    y = F.batch_norm(x, running_mean, running_var, weight.float(), bias.float(), training, momentum, eps)
    if residual is not None:
        y = y + residual
    return F.relu(y) if relu else y


# ------------------------------------------------------------------ pooling (channels-last bf16)
class _MaxPool2d(torch.autograd.Function):
    """Max pooling on NHWC storage (csrc/pool.cu): the forward keeps one byte per output element
    (the winning tap), the backward is a gather over the windows that contain each input pixel."""

    @staticmethod
    def forward(ctx, x, k, stride, pad):
        L = nat.lib()
        xh = x.permute(0, 2, 3, 1)
        if not xh.is_contiguous():
            xh = xh.contiguous()
        N, H, W_, Cc = xh.shape
        P = (H + 2 * pad - k) // stride + 1
        Q = (W_ + 2 * pad - k) // stride + 1
        y = torch.empty(N, P, Q, Cc, dtype=torch.bfloat16, device=x.device)
        idx = torch.empty(N, P, Q, Cc, dtype=torch.uint8, device=x.device)
        nat.check(L.b200_maxpool_fwd(xh.data_ptr(), y.data_ptr(), idx.data_ptr(), N, H, W_, Cc, k, k, stride, pad,
                                     nat.stream_ptr(x.device)), "b200_maxpool_fwd")
        ctx.save_for_backward(idx)
        ctx.geom = (N, H, W_, Cc, k, stride, pad)
        return y.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy):
        L = nat.lib()
        (idx,) = ctx.saved_tensors
        N, H, W_, Cc, k, stride, pad = ctx.geom
        dyh = dy.permute(0, 2, 3, 1)
        if not dyh.is_contiguous():
            dyh = dyh.contiguous()
        dx = torch.empty(N, H, W_, Cc, dtype=torch.bfloat16, device=dy.device)
        nat.check(L.b200_maxpool_bwd(dyh.data_ptr(), idx.data_ptr(), dx.data_ptr(), N, H, W_, Cc, k, k, stride, pad,
                                     nat.stream_ptr(dy.device)), "b200_maxpool_bwd")
        return dx.permute(0, 3, 1, 2), None, None, None


def max_pool2d(x, kernel_size: int, stride: int, padding: int = 0):
    if native_ok(x) and x.dim() == 4 and x.shape[1] % 8 == 0 and 2 * padding <= kernel_size and kernel_size <= 15:
        return _MaxPool2d.apply(x, kernel_size, stride, padding)
    return F.max_pool2d(x, kernel_size, stride, padding)


class _GlobalAvgPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        L = nat.lib()
        xh = x.permute(0, 2, 3, 1)
        if not xh.is_contiguous():
            xh = xh.contiguous()
        N, H, W_, Cc = xh.shape
        y = torch.empty(N, Cc, dtype=torch.bfloat16, device=x.device)
        nat.check(L.b200_avgpool_fwd(xh.data_ptr(), y.data_ptr(), N, H * W_, Cc, nat.stream_ptr(x.device)),
                  "b200_avgpool_fwd")
        ctx.geom = (N, H, W_, Cc)
        return y

    @staticmethod
    def backward(ctx, dy):
        L = nat.lib()
        N, H, W_, Cc = ctx.geom
        dy2 = dy.contiguous()
        dx = torch.empty(N, H, W_, Cc, dtype=torch.bfloat16, device=dy.device)
        nat.check(L.b200_avgpool_bwd(dy2.data_ptr(), dx.data_ptr(), N, H * W_, Cc, nat.stream_ptr(dy.device)),
                  "b200_avgpool_bwd")
        return dx.permute(0, 3, 1, 2)


def global_avg_pool(x):
    """[N, C, H, W] -> [N, C] (mean over the spatial positions)."""
    if native_ok(x) and x.dim() == 4 and x.shape[1] % 8 == 0:
        return _GlobalAvgPool.apply(x)
    return torch.flatten(F.adaptive_avg_pool2d(x, 1), 1)
