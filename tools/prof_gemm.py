"""Runs the tcgen05 GEMM a few times (target for ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200ddp import _native as nat
from b200ddp._native import bindings as B
assert nat.available()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
dev = torch.device("cuda")
a = torch.randn(n, n, device=dev).to(torch.bfloat16)
b = torch.randn(n, n, device=dev).to(torch.bfloat16)
out = torch.empty(n, n, dtype=torch.bfloat16, device=dev)
for _ in range(reps):
    B.gemm(a, b, out, M=n, N=n, K=n)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    B.gemm(a, b, out, M=n, N=n, K=n)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print(f"gemm {n}^3 bf16: {ms:.4f} ms, {2*n**3/ms/1e9:.1f} TFLOP/s")
