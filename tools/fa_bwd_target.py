import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200ddp.ops import functional as Fn
dev = torch.device("cuda")
qkv = torch.randn(8, 1024, 3 * 768, device=dev).to(torch.bfloat16).requires_grad_(True)
out = Fn.causal_attention(qkv, 12)
d = torch.randn_like(out)
for _ in range(3):
    torch.autograd.grad(out, qkv, d, retain_graph=True)
torch.cuda.synchronize()
print("OK")
