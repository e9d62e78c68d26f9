"""LayerNorm backward at the GPT-2 shape (8192 x 768) and the roofline shape (65536 x 768), CUDA-event timed."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200ddp.ops import functional as Fn
dev = torch.device("cuda")
for rows in (8192, 65536):
    # rotate over enough distinct tensors that nothing comes from L2
    k = max(2, int(300e6 // (rows * 768 * 2 * 3)))
    xs = [torch.randn(rows, 768, device=dev).to(torch.bfloat16).requires_grad_(True) for _ in range(k)]
    w = torch.ones(768, device=dev, dtype=torch.bfloat16, requires_grad=True)
    b = torch.zeros(768, device=dev, dtype=torch.bfloat16, requires_grad=True)
    ys = [Fn.layer_norm(x, w, b, 1e-5) for x in xs]
    dy = torch.randn(rows, 768, device=dev).to(torch.bfloat16)
    for i in range(k):
        torch.autograd.grad(ys[i], (xs[i], w, b), dy, retain_graph=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        for i in range(k):
            torch.autograd.grad(ys[i], (xs[i], w, b), dy, retain_graph=True)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (reps * k)
    print(f"layernorm bwd {rows} x 768: {us:7.1f} us  ({rows * 768 * 2 * 3 / us / 1e3:6.0f} GB/s)")
