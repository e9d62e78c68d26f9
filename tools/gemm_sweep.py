"""GEMM throughput for a few shapes/tiles vs torch.matmul (cuBLAS) on the same box."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200ddp import _native as nat
from b200ddp._native import bindings as B
assert nat.available()
dev = torch.device("cuda")
def timeit(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
shapes = [(4096, 4096, 4096), (8192, 8192, 8192), (8192, 3072, 768), (8192, 768, 3072), (8192, 2304, 768),
          (768, 3072, 8192), (8192, 50257, 768)]
for M, N, K in shapes:
    a = torch.randn(M, K, device=dev).to(torch.bfloat16); b = torch.randn(N, K, device=dev).to(torch.bfloat16)
    Np = (N + 7) // 8 * 8
    out = torch.empty(M, Np, dtype=torch.bfloat16, device=dev)
    fl = 2.0 * M * N * K
    row = [f"{M}x{N}x{K}"]
    for name, kw in (("auto", {}), ("128x256", dict(tile_n=256, tile_m=128)), ("256x256", dict(tile_n=256, tile_m=256)), ("2sm", dict(cta_pair=1))):
        ms = timeit(lambda: B.gemm(a, b, out, M=M, N=N, K=K, **kw))
        row.append(f"{name} {ms*1e3:8.1f}us {fl/ms/1e9:7.1f}TF")
    B.gemm(a, b, out, M=M, N=N, K=K, tile_n=256, tile_m=256)
    ref = torch.matmul(a, b.t())
    err = ((out[:, :N].float() - ref.float()).norm() / ref.float().norm()).item()
    row.append(f"relerr256 {err:.1e}")
    ms = timeit(lambda: torch.matmul(a, b.t()))
    row.append(f"cublas {ms*1e3:8.1f}us {fl/ms/1e9:7.1f}TF")
    print("  ".join(row), flush=True)
