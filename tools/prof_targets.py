"""Small single-GPU targets for ncu: `python tools/prof_targets.py gemm256|flash|step`."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200ddp import _native as nat
from b200ddp._native import bindings as B
assert nat.available()
dev = torch.device("cuda")
what = sys.argv[1]
if what == "gemm256":
    n = 8192
    a = torch.randn(n, n, device=dev).to(torch.bfloat16); b = torch.randn(n, n, device=dev).to(torch.bfloat16)
    out = torch.empty(n, n, dtype=torch.bfloat16, device=dev)
    for _ in range(4):
        B.gemm(a, b, out, M=n, N=n, K=n)
elif what == "flash":
    from b200ddp.ops import functional as Fn
    qkv = torch.randn(8, 1024, 3 * 768, device=dev).to(torch.bfloat16)
    for _ in range(4):
        Fn.flash_attention_forward(qkv, 12)
elif what == "step":
    from b200ddp import DataParallel, SimpleNet
    eng = DataParallel(SimpleNet().to(dev), backend="b200", compute_dtype=torch.bfloat16)
    x = torch.randn(64, 784, device=dev); y = torch.randint(0, 10, (64,), device=dev)
    for _ in range(12):
        eng.train_step(x, y)
torch.cuda.synchronize()
print("OK", what)
