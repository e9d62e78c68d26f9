#!/usr/bin/env python3
"""Achieved bandwidth / throughput of every kernel family at a representative size, as a
fraction of the measured peaks in MEASURED_PEAKS.json ("of measured").  Single GPU.

Timing: CUDA events, 3 warm-up + 10 timed iterations; inputs are larger than L2 (126 MB) or
rotated, except where the working set is inherently L2-resident (noted)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.nn.functional as F
from b200ddp import _native as nat
from b200ddp._native import bindings as B
from b200ddp.ops import functional as Fn
assert nat.available()
dev = torch.device("cuda")
try:
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    HBM, TF, tag = peaks["hbm_gbs"], peaks["bf16_tflops"], "measured"
except Exception:
    HBM, TF, tag = 6650.0, 1590.0, "fallback"
bf = torch.bfloat16


def timeit(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3  # us


rows = []
def mem(name, nbytes, us, note=""):
    gbs = nbytes / us / 1e3
    rows.append((name, f"{us:9.1f} us", f"{gbs:7.0f} GB/s", f"{gbs / HBM:5.2f} of {tag} HBM copy", note))
def flop(name, fl, us, note=""):
    tf = fl / us / 1e6
    rows.append((name, f"{us:9.1f} us", f"{tf:7.0f} TFLOP/s", f"{tf / TF:5.2f} of {tag} cuBLAS bf16", note))

# ---- GEMM
for M, N, K in ((8192, 8192, 8192), (4096, 4096, 4096), (8192, 3072, 768), (8192, 768, 3072)):
    a = torch.randn(M, K, device=dev).to(bf); b = torch.randn(N, K, device=dev).to(bf)
    out = torch.empty(M, N, dtype=bf, device=dev)
    flop(f"gemm {M}x{N}x{K}", 2.0 * M * N * K, timeit(lambda: B.gemm(a, b, out, M=M, N=N, K=K)))
# ---- optimizer (28 B/param fp32 grads + 2 B shadow)
from b200ddp.parallel.engine import DataParallel
class Big(torch.nn.Module):
    def __init__(self):
        super().__init__(); self.w = torch.nn.Parameter(torch.zeros(64 << 20, device=dev))
eng = DataParallel(Big(), backend="b200", compute_dtype=bf, wire_dtype=torch.float32, use_program=False)
eng.optimizer.step_count = 1
n = eng.layout.total
mem("fused Adam (fp32 g, bf16 shadow), 64 M params", n * 30, timeit(lambda: B.optimizer_step(eng.optimizer, 0, n, 1.0)))
del eng
# ---- cross entropy (LM head shape): 2 reads + 1 write of the logits
rowsN, V = 8192, 50257
lg = torch.randn(rowsN, 50264, device=dev).to(bf)[:, :V]
tg = torch.randint(0, V, (rowsN,), device=dev); o2 = torch.zeros(2, device=dev)
mem("cross-entropy fwd+bwd 8192x50257 bf16 (in place)", rowsN * V * 2 * 3, timeit(lambda: B.cross_entropy(lg, tg, lg, 1.0 / rowsN, o2)))
del lg
# ---- layernorm
x = torch.randn(65536, 768, device=dev).to(bf).requires_grad_(True)
g = torch.ones(768, device=dev, dtype=bf).requires_grad_(True); bb = torch.zeros(768, device=dev, dtype=bf).requires_grad_(True)
mem("layernorm fwd 65536x768", 65536 * 768 * 2 * 2, timeit(lambda: Fn.layer_norm(x.detach(), g.detach(), bb.detach())))
y = Fn.layer_norm(x, g, bb); dy = torch.randn_like(y)
mem("layernorm bwd 65536x768", 65536 * 768 * 2 * 3, timeit(lambda: torch.autograd.grad(y, (x, g, bb), dy, retain_graph=True)))
del x, y, dy
# ---- batch norm (+relu+residual)
xc = torch.randn(64, 256, 56, 56, device=dev).to(bf).contiguous(memory_format=torch.channels_last).requires_grad_(True)
rc = torch.randn_like(xc).requires_grad_(True)
w = torch.ones(256, device=dev, dtype=bf).requires_grad_(True); b_ = torch.zeros(256, device=dev, dtype=bf).requires_grad_(True)
rm, rv = torch.zeros(256, device=dev), torch.ones(256, device=dev)
nb = xc.numel() * 2
mem("batchnorm+add+relu fwd 64x256x56x56", nb * 4, timeit(lambda: Fn.batch_norm_act(xc.detach(), w.detach(), b_.detach(), rm, rv, True, 0.1, 1e-5, True, rc.detach())), "stats pass + apply pass")
yb = Fn.batch_norm_act(xc, w, b_, rm, rv, True, 0.1, 1e-5, True, rc); dyb = torch.randn_like(yb)
mem("batchnorm+add+relu bwd 64x256x56x56", nb * 8, timeit(lambda: torch.autograd.grad(yb, (xc, rc, w, b_), dyb, retain_graph=True)), "reduce pass (3 reads) + apply pass (3 reads, 2 writes)")
del xc, rc, yb, dyb
# ---- im2col / col2im
xi = torch.randn(64, 56, 56, 64, device=dev).to(bf)
col = Fn._im2col(xi, 3, 3, 1, 1)
mem("im2col 64x56x56x64, 3x3", xi.numel() * 2 + col.numel() * 2, timeit(lambda: Fn._im2col(xi, 3, 3, 1, 1)))
dx = torch.empty_like(xi)
mem("col2im 64x56x56x64, 3x3", xi.numel() * 2 + col.numel() * 2, timeit(lambda: nat.check(nat.lib().b200_col2im(col.data_ptr(), dx.data_ptr(), 64, 56, 56, 64, 3, 3, 1, 1, nat.stream_ptr(dev)), "col2im")))
del xi, col, dx
# ---- column sum
z = torch.randn(65536, 3072, device=dev).to(bf); oc = torch.empty(3072, device=dev, dtype=bf)
mem("colsum (bias grad) 65536x3072", z.numel() * 2, timeit(lambda: Fn._colsum_into(z, oc)))
del z
# ---- attention (GPT-2 small layer)
qkv = torch.randn(8, 1024, 3 * 768, device=dev).to(bf).requires_grad_(True)
att_fl = 4.0 * 8 * 12 * 1024 * 1024 * 64 / 2  # causal: half of QK^T + PV
flop("flash attention fwd B8 H12 T1024 D64", att_fl, timeit(lambda: Fn.flash_attention_forward(qkv.detach(), 12)))
oa = Fn.causal_attention(qkv, 12); da = torch.randn_like(oa)
flop("flash attention bwd B8 H12 T1024 D64", att_fl * 2.5, timeit(lambda: torch.autograd.grad(oa, qkv, da, retain_graph=True)))
w0 = max(len(r[0]) for r in rows)
for r in rows:
    print(f"{r[0]:<{w0}}  {r[1]}  {r[2]}  {r[3]}  {r[4]}")
