import os, sys
sys.path.insert(0, "/root/repo")
import torch
from torch.profiler import ProfilerActivity, profile
from b200ddp.ops import functional as Fn
dev = torch.device("cuda")
qkv = torch.randn(8, 1024, 3 * 768, device=dev).to(torch.bfloat16)
for _ in range(3): Fn.flash_attention_forward(qkv, 12)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(5): Fn.flash_attention_forward(qkv, 12)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=5, max_name_column_width=60))

if os.environ.get("B200DDP_FA_TRACE"):
    import ctypes as C
    from b200ddp import _native as nat
    buf = (C.c_longlong * 128)()
    nat.lib().b200_flash_debug_trace.argtypes = [C.c_void_p]
    nat.check(nat.lib().b200_flash_debug_trace(buf), "trace")
    print("step: wait S | tmem ld | max | wait O(i-1) | exp+store | arrive   (cycles; block 0 = heaviest pair, tile 0, thread 0)")
    for i in range(7):
        r = [buf[i * 8 + k] for k in range(7)]
        print(i, [r[k + 1] - r[k] if r[k + 1] and r[k] else None for k in range(6)], "step total", (buf[(i + 1) * 8] - r[0]) if buf[(i + 1) * 8] else None)

if os.environ.get("B200DDP_FA_TRACE_BWD"):
    import ctypes as C
    from b200ddp import _native as nat
    q2 = qkv.clone().requires_grad_(True)
    out = Fn.causal_attention(q2, 12)
    for _ in range(2):
        torch.autograd.grad(out, q2, torch.randn_like(out), retain_graph=True)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(3): torch.autograd.grad(out, q2, torch.randn_like(out), retain_graph=True)
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=8, max_name_column_width=50))
    buf = (C.c_longlong * 128)()
    nat.lib().b200_flash_debug_trace.argtypes = [C.c_void_p]
    nat.check(nat.lib().b200_flash_debug_trace(buf), "trace")
    print("bwd step: wait S,dP | P,dS compute | wait dQ | dQ atomics   (cycles; key block 0, thread 0)")
    for i in range(8):
        r = [buf[i * 8 + k] for k in range(5)]
        print(i, [r[k + 1] - r[k] for k in range(4)], "step total", (buf[(i + 1) * 8] - r[0]) if i < 7 else None)
