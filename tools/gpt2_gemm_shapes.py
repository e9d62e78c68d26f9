"""Every GEMM of one GPT-2-small training step (B=8, T=1024), timed alone with the operand
layouts and epilogues the autograd functions use, against torch.matmul on the same box.
Buffers rotate through a pool larger than L2 so no call finds its operands cached."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200ddp import _native as nat
from b200ddp._native import bindings as B

assert nat.available()
dev = torch.device("cuda")
POOL = 4


def timeit(fn, reps=12):
    for i in range(3):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def bf(*shape):
    return [torch.randn(*shape, device=dev).to(torch.bfloat16) * 0.05 for _ in range(POOL)]


T = 8192
C = 768
tiles = (("auto", {}), ("128x256", dict(tile_n=256, tile_m=128)), ("128x128", dict(tile_n=128, tile_m=128)),
         ("256x256", dict(tile_n=256, tile_m=256)), ("2sm256", dict(cta_pair=1, tile_n=256)), ("2sm128", dict(cta_pair=1, tile_n=128)))
total = {k: 0.0 for k, _ in tiles}
total["cublas"] = 0.0
per_layer = 12


def run(name, M, N, K, a, b, kw, count, ref):
    fl = 2.0 * M * N * K
    ld = (N + 7) // 8 * 8
    outs = [torch.empty(M, ld, dtype=torch.bfloat16, device=dev) for _ in range(POOL)]
    row = [f"{name:<22s} {M}x{N}x{K}"]
    for tname, tkw in tiles:
        if tname == "256x256" and (kw.get("epilogue") or M < 512):
            pass
        try:
            ms = timeit(lambda i: B.gemm(a[i % POOL], b[i % POOL], outs[i % POOL], M=M, N=N, K=K, **kw, **tkw))
            row.append(f"{tname} {ms*1e3:7.1f}us {fl/ms/1e9:6.0f}TF")
            total[tname] += ms * count
        except Exception as e:  # tile not valid for this problem
            row.append(f"{tname}     n/a")
            total[tname] += float("nan")
    ms = timeit(lambda i: ref(i % POOL))
    total["cublas"] += ms * count
    row.append(f"cublas {ms*1e3:7.1f}us {fl/ms/1e9:6.0f}TF")
    print("  ".join(row), flush=True)


for lname, N, K, extra in (("qkv", 2304, C, {}), ("attn_proj", C, C, {"res": True}), ("fc", 3072, C, {"gelu": True}),
                           ("fc2", C, 3072, {"res": True}), ("lm_head", 50257, C, {"head": True})):
    count = 1 if extra.get("head") else per_layer
    x = bf(T, K)
    w = bf(N, K)
    dy = bf(T, (N + 7) // 8 * 8)
    dyv = [d[:, :N] for d in dy]
    bias = torch.zeros(N, dtype=torch.float32, device=dev)
    # forward: x[T,K] * w[N,K]^T
    kw = {}
    if not extra.get("head"):
        kw["bias"] = bias
    if extra.get("gelu"):
        kw["epilogue"] = B.EPI_GELU
        kw["aux_out"] = torch.empty(T, N, dtype=torch.bfloat16, device=dev)
    if extra.get("res"):
        kw["aux"] = torch.zeros(T, N, dtype=torch.bfloat16, device=dev)
    run(lname + " fwd", T, N, K, x, w, kw, count, lambda i: torch.matmul(x[i], w[i].t()))
    # dgrad: dy[T,N] * w[N,K]  (B MN-major)
    kw = dict(b_mn_major=True)
    if lname == "fc2":
        kw["epilogue"] = B.EPI_GELU_BWD
        kw["aux"] = torch.zeros(T, K, dtype=torch.bfloat16, device=dev)
    run(lname + " dgrad", T, K, N, dyv, w, kw, count, lambda i: torch.matmul(dyv[i], w[i]))
    # wgrad: dy^T[N,T] * x[T,K]  (both MN-major), the function's own split-K policy
    from b200ddp.ops.functional import _wgrad_into
    dsts = [torch.empty(N, K, dtype=torch.bfloat16, device=dev) for _ in range(POOL)]
    fl = 2.0 * T * N * K
    ms = timeit(lambda i: _wgrad_into(dyv[i % POOL], x[i % POOL], dsts[i % POOL]))
    msr = timeit(lambda i: torch.matmul(dyv[i % POOL].t(), x[i % POOL]))
    for k in total:
        total[k] += (msr if k == "cublas" else ms) * count
    extra_cols = []
    for tname, tkw in (("2sm256", dict(cta_pair=1, tile_n=256)), ("2sm128", dict(cta_pair=1, tile_n=128))):
        m2 = timeit(lambda i: B.gemm(dyv[i % POOL], x[i % POOL], dsts[i % POOL], M=N, N=K, K=T, a_mn_major=True,
                                     b_mn_major=True, **tkw))
        extra_cols.append(f"{tname} {m2*1e3:7.1f}us {fl/m2/1e9:6.0f}TF")
    print(f"{lname + ' wgrad':<22s} {N}x{K}x{T}  policy {ms*1e3:7.1f}us {fl/ms/1e9:6.0f}TF  " + "  ".join(extra_cols) +
          f"  cublas {msr*1e3:7.1f}us {fl/msr/1e9:6.0f}TF", flush=True)

print("per-step GEMM time (ms):", {k: round(v, 3) for k, v in total.items()})
