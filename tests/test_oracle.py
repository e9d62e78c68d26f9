"""Host numerics oracle (oracle/oracle.cpp) against torch on the CPU.

The oracle states what each sm_100a kernel computes; these tests pin the oracle itself to torch, so
kernel-vs-oracle comparisons on a GPU box and the shared optimizer code are anchored on a box
without one.  ("CPU C++ oracle of the fused kernel vs torch.optim".)
"""
import time

import pytest
import torch
import torch.nn.functional as F

from b200ddp._native import oracle as O


def bf(x):
    return x.to(torch.bfloat16)


def rel_err(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


@pytest.mark.parametrize("kind,kw", [
    ("adam", dict(lr=1e-3)), ("adam", dict(lr=3e-3, weight_decay=0.01)),
    ("adamw", dict(lr=1e-3, weight_decay=0.05, betas=(0.8, 0.95))),
    ("sgd", dict(lr=0.05, momentum=0.9)), ("sgd", dict(lr=0.1, momentum=0.9, weight_decay=1e-4)),
])
def test_shared_optimizer_update_matches_torch_optim_to_rounding(kind, kw):
    """csrc/optim_math.h — the code optim.cu and comm.cu compile — against torch.optim, 6 steps.
    The shared code fuses multiply-adds (one rounding, like torch's CUDA kernels); torch's CPU
    kernels round twice, so agreement is to the last bit or two, not bit-for-bit."""
    torch.manual_seed(0)
    n = 4099
    p0 = torch.randn(n)
    ref_p = p0.clone().requires_grad_(True)
    tkw = dict(kw)
    if kind == "adam":
        opt = torch.optim.Adam([ref_p], foreach=False, **tkw)
    elif kind == "adamw":
        opt = torch.optim.AdamW([ref_p], foreach=False, **tkw)
    else:
        opt = torch.optim.SGD([ref_p], foreach=False, **tkw)
    p, m = p0.clone(), torch.zeros(n)
    v = torch.zeros(n) if kind != "sgd" else None
    okw = dict(lr=kw["lr"], weight_decay=kw.get("weight_decay", 0.0), momentum=kw.get("momentum", 0.0),
               betas=kw.get("betas", (0.9, 0.999)))
    for t in range(1, 7):
        g = torch.randn(n) * (0.1 * t)
        ref_p.grad = g.clone()
        opt.step()
        O.optimizer_step(kind, p, g, m, v, t, **okw)
        assert torch.allclose(p, ref_p.detach(), rtol=3e-6, atol=3e-7), (kind, t, (p - ref_p.detach()).abs().max())
    st = opt.state[ref_p]
    if kind == "sgd":
        assert torch.allclose(m, st["momentum_buffer"], rtol=3e-6, atol=3e-7)
    else:
        assert torch.allclose(m, st["exp_avg"], rtol=3e-6, atol=3e-7)
        # the C ABI carries the betas as fp32: 1 - fp32(0.999) differs from 1 - 0.999 by 1.3e-5 relative
        assert torch.allclose(v, st["exp_avg_sq"], rtol=5e-5, atol=1e-9)


def test_optimizer_grad_scale_and_bf16_shadow():
    torch.manual_seed(1)
    n = 1000
    p, g, m, v = torch.randn(n), torch.randn(n), torch.zeros(n), torch.zeros(n)
    q, m2, v2 = p.clone(), m.clone(), v.clone()
    shadow = torch.empty(n, dtype=torch.bfloat16)
    O.optimizer_step("adam", p, g * 8, m, v, 1, lr=1e-2, grad_scale=0.125, shadow=shadow)
    O.optimizer_step("adam", q, g, m2, v2, 1, lr=1e-2)
    assert torch.equal(p, q) and torch.equal(shadow, p.to(torch.bfloat16))


def test_ordered_allreduce():
    torch.manual_seed(2)
    parts = [torch.randn(777) for _ in range(8)]
    want = parts[0].clone()
    for t in parts[1:]:
        want += t
    assert torch.equal(O.allreduce_sum(parts, 0.125), want * 0.125)


@pytest.mark.parametrize("epi", ["none", "bias_res", "relu", "gelu", "gelu_bwd"])
def test_gemm_epilogues(epi):
    torch.manual_seed(3)
    M, N, K = 37, 24, 72
    a, b = bf(torch.randn(M, K)), bf(torch.randn(N, K) * 0.2)
    acc = a.float() @ b.float().t()
    bias = torch.randn(N)
    if epi == "none":
        d, _ = O.gemm(a, b)
        want = acc
    elif epi == "bias_res":
        r = bf(torch.randn(M, N))
        d, _ = O.gemm(a, b, bias=bias, aux=r)
        want = acc + bias + r.float()
    elif epi == "relu":
        d, _ = O.gemm(a, b, bias=bias, epilogue=O.EPI_RELU)
        want = F.relu(acc + bias)
    elif epi == "gelu":
        d, z = O.gemm(a, b, bias=bias, epilogue=O.EPI_GELU, want_aux_out=True)
        want = F.gelu(acc + bias, approximate="tanh")
        assert rel_err(z, acc + bias) < 4e-3
    else:
        z = bf(torch.randn(M, N)).requires_grad_(False)
        d, _ = O.gemm(a, b, epilogue=O.EPI_GELU_BWD, aux=z)
        zz = z.float().requires_grad_(True)
        F.gelu(zz, approximate="tanh").sum().backward()
        want = acc * zz.grad
    assert rel_err(d, want) < 4e-3, rel_err(d, want)


def test_cross_entropy_forward_backward():
    torch.manual_seed(4)
    rows, cols = 50, 13
    logits = torch.randn(rows, cols) * 3
    targets = torch.randint(0, cols, (rows,))
    loss, correct, dl = O.cross_entropy(logits, targets, 1.0 / rows)
    lt = logits.clone().requires_grad_(True)
    want = F.cross_entropy(lt, targets)
    want.backward()
    assert abs(loss - want.item()) < 1e-5
    assert correct == float((logits.argmax(1) == targets).sum())
    assert torch.allclose(dl, lt.grad, atol=1e-6)


def test_layernorm_forward_backward():
    torch.manual_seed(5)
    rows, Cc = 33, 256
    x, gamma, beta = bf(torch.randn(rows, Cc) * 2 + 0.5), bf(torch.rand(Cc) + 0.5), bf(torch.randn(Cc) * 0.1)
    y, mean, rstd = O.layernorm_fwd(x, gamma, beta)
    xr = x.float().requires_grad_(True)
    gr, br = gamma.float().requires_grad_(True), beta.float().requires_grad_(True)
    yr = F.layer_norm(xr, (Cc,), gr, br, 1e-5)
    assert rel_err(y, yr) < 4e-3
    assert torch.allclose(mean, xr.detach().mean(1), atol=1e-5)
    dy = bf(torch.randn(rows, Cc))
    yr.backward(dy.float())
    dx, dg, db = O.layernorm_bwd(dy, x, gamma, mean, rstd)
    assert rel_err(dx, xr.grad) < 5e-3 and rel_err(dg, gr.grad) < 1e-4 and rel_err(db, br.grad) < 1e-5


@pytest.mark.parametrize("relu,res", [(False, False), (True, True)])
def test_batchnorm_training_forward(relu, res):
    torch.manual_seed(6)
    N, Cc, H = 4, 16, 5
    x = bf(torch.randn(N, H, H, Cc) * 1.5 + 0.3)
    r = bf(torch.randn(N, H, H, Cc)) if res else None
    gamma, beta = torch.rand(Cc) + 0.5, torch.randn(Cc) * 0.1
    rm, rv = torch.zeros(Cc), torch.ones(Cc)
    y = O.batchnorm_fwd(x.view(-1, Cc), gamma, beta, rm, rv, relu=relu, residual=r.view(-1, Cc) if res else None)
    rm2, rv2 = torch.zeros(Cc), torch.ones(Cc)
    want = F.batch_norm(x.float().permute(0, 3, 1, 2), rm2, rv2, gamma, beta, True, 0.1, 1e-5).permute(0, 2, 3, 1)
    if res:
        want = want + r.float()
    if relu:
        want = F.relu(want)
    assert rel_err(y.view(N, H, H, Cc), want) < 5e-3
    assert torch.allclose(rm, rm2, atol=1e-5) and torch.allclose(rv, rv2, atol=1e-4)


@pytest.mark.parametrize("k,stride,pad,H", [(3, 1, 1, 7), (3, 2, 1, 8), (1, 2, 0, 6), (7, 2, 3, 12)])
def test_im2col_matches_unfold_and_col2im_is_its_adjoint(k, stride, pad, H):
    torch.manual_seed(7)
    N, Cc = 2, 8
    x = bf(torch.randn(N, H, H, Cc))
    col = O.im2col(x, k, k, stride, pad)
    unf = F.unfold(x.float().permute(0, 3, 1, 2), k, padding=pad, stride=stride)        # [N, C*k*k, L]
    L = unf.shape[-1]
    want = unf.view(N, Cc, k * k, L).permute(0, 3, 2, 1).reshape(N * L, k * k * Cc)     # (r,s,c) order
    assert torch.equal(col.float(), want)
    dcol = bf(torch.randn_like(col.float()))
    dx = O.col2im(dcol, (N, H, H, Cc), k, k, stride, pad)
    fold_in = dcol.float().view(N, L, k * k, Cc).permute(0, 3, 2, 1).reshape(N, Cc * k * k, L)
    want_dx = F.fold(fold_in, (H, H), k, padding=pad, stride=stride).permute(0, 2, 3, 1)
    assert rel_err(dx, want_dx) < 4e-3


@pytest.mark.parametrize("k,stride,pad,H", [(3, 2, 1, 9), (2, 2, 0, 8), (3, 1, 1, 5)])
def test_pooling(k, stride, pad, H):
    torch.manual_seed(10)
    N, Cc = 2, 8
    x = bf(torch.randn(N, H, H, Cc))
    y, idx = O.maxpool_fwd(x, k, stride, pad)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    yr = F.max_pool2d(xr, k, stride, pad)
    assert torch.equal(y.float(), yr.detach().permute(0, 2, 3, 1))
    dy = bf(torch.randn(*y.shape))
    yr.backward(dy.float().permute(0, 3, 1, 2))
    dx = O.maxpool_bwd(dy, idx, x.shape, k, stride, pad)
    assert rel_err(dx, xr.grad.permute(0, 2, 3, 1)) < 4e-3
    assert rel_err(O.avgpool_fwd(x), x.float().mean((1, 2))) < 4e-3


def test_causal_attention_matches_sdpa():
    torch.manual_seed(8)
    B, T, H, D = 2, 40, 3, 16
    qkv = bf(torch.randn(B, T, 3 * H * D))
    out, lse = O.causal_attention(qkv, H)
    q, k, v = qkv.float().view(B, T, 3, H, D).permute(2, 0, 3, 1, 4)
    want = F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(B, T, H * D)
    assert rel_err(out, want) < 8e-3
    s = (q @ k.transpose(-1, -2)) / D ** 0.5
    mask = torch.ones(T, T, dtype=torch.bool).tril()
    assert torch.allclose(lse, torch.logsumexp(s.masked_fill(~mask, float("-inf")), -1), atol=1e-4)


def test_timed_cpu_baseline_of_one_simplenet_step():
    """The reference model's step (784-256-256-10, batch 64) through the oracle: forward GEMMs with
    fused ReLU, cross-entropy, and the Adam update of all 269,322 parameters.  Printed, not asserted:
    the CPU figure that the 50 us GPU step is compared with in DESIGN.md."""
    torch.manual_seed(9)
    x = bf(torch.randn(64, 784))
    ws = [bf(torch.randn(256, 784) * 0.03), bf(torch.randn(256, 256) * 0.06), bf(torch.randn(10, 256) * 0.06)]
    bs = [torch.zeros(256), torch.zeros(256), torch.zeros(10)]
    n = sum(w.numel() + b.numel() for w, b in zip(ws, bs))
    assert n == 269_322
    p, g, m, v = torch.randn(n), torch.randn(n), torch.zeros(n), torch.zeros(n)
    t0 = time.perf_counter()
    h1, _ = O.gemm(x, ws[0], bias=bs[0], epilogue=O.EPI_RELU)
    h2, _ = O.gemm(h1, ws[1], bias=bs[1], epilogue=O.EPI_RELU)
    lg, _ = O.gemm(h2, ws[2], bias=bs[2])
    loss, _, _ = O.cross_entropy(lg.float(), torch.randint(0, 10, (64,)), 1 / 64)
    O.optimizer_step("adam", p, g, m, v, 1, lr=1e-3)
    ms = (time.perf_counter() - t0) * 1e3
    assert loss == loss and ms > 0
    print(f"\n[oracle] forward + CE + Adam of one SimpleNet step on the host: {ms:.1f} ms (single thread)")
