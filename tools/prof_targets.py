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
elif what == "gemm_qkv":
    # short-K, epilogue-heavy: GPT-2 qkv forward (bias) and attn_proj forward (bias + residual)
    a = torch.randn(8192, 768, device=dev).to(torch.bfloat16); b = torch.randn(2304, 768, device=dev).to(torch.bfloat16)
    out = torch.empty(8192, 2304, dtype=torch.bfloat16, device=dev)
    bias = torch.zeros(2304, dtype=torch.float32, device=dev)
    for _ in range(4):
        B.gemm(a, b, out, M=8192, N=2304, K=768, bias=bias)
elif what == "gemm_shortk":
    # plain epilogue, K = 768, many tiles per SM: steady-state mainloop vs epilogue
    a = torch.randn(8192, 768, device=dev).to(torch.bfloat16); b = torch.randn(8192, 768, device=dev).to(torch.bfloat16)
    out = torch.empty(8192, 8192, dtype=torch.bfloat16, device=dev)
    for _ in range(4):
        B.gemm(a, b, out, M=8192, N=8192, K=768)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        B.gemm(a, b, out, M=8192, N=8192, K=768)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"8192x8192x768 plain: {ms*1e3:.1f} us  {2*8192*8192*768/ms/1e9:.0f} TFLOP/s")
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
elif what == "mlpstep":
    from b200ddp import DataParallel, SimpleNet
    eng = DataParallel(SimpleNet().to(dev), backend="b200", compute_dtype=torch.bfloat16)
    xs = torch.randn(8, 64, 784, device=dev); ys = torch.randint(0, 10, (8, 64), device=dev)
    for _ in range(3):      # three launches of 24 steps each (the update rides in the step launch)
        eng.train_steps(xs, ys, 0, 24)
elif what in ("gpt2step", "rn50step"):
    from b200ddp import DataParallel
    torch.manual_seed(0)
    if what == "rn50step":
        from b200ddp.models.resnet50 import resnet50
        net = resnet50().to(dev)
        x = torch.randn(32, 3, 224, 224, device=dev); y = torch.randint(0, 1000, (32,), device=dev)
    else:
        from b200ddp.models.gpt2 import GPT2
        net = GPT2().to(dev)
        t = torch.randint(0, 50257, (8, 1025), device=dev); x, y = t[:, :-1].contiguous(), t[:, 1:].contiguous()
    eng = DataParallel(net, optimizer="adamw", lr=3e-4, backend="b200", compute_dtype=torch.bfloat16, step_graph=False)
    for _ in range(2):
        eng.train_step(x, y)
torch.cuda.synchronize()
print("OK", what)
