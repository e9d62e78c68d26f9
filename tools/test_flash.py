import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from b200ddp import _native as nat
assert nat.available()
L = nat.lib(); dev = torch.device("cuda")
def rel(a, b): return ((a.float()-b.float()).norm()/b.float().norm()).item()
for (Bs, H, T) in ((1, 1, 128), (2, 4, 256), (2, 12, 1024)):
    D = 64; Cc = H*D
    torch.manual_seed(0)
    qkv = torch.randn(Bs, T, 3*Cc, device=dev).to(torch.bfloat16)
    out = torch.empty(Bs, T, Cc, dtype=torch.bfloat16, device=dev)
    lse = torch.empty(Bs, H, T, dtype=torch.float32, device=dev)
    nat.check(L.b200_flash_attn_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), Bs, T, H, D, nat.stream_ptr(dev)), "fwd")
    torch.cuda.synchronize()
    q, k, v = qkv.float().view(Bs, T, 3, H, D).permute(2, 0, 3, 1, 4)
    ref = F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(Bs, T, Cc)
    s = (q @ k.transpose(-1, -2)) / 8.0
    mask = torch.ones(T, T, device=dev, dtype=torch.bool).tril()
    lse_ref = torch.logsumexp(s.masked_fill(~mask, float("-inf")), dim=-1)
    print(f"B{Bs} H{H} T{T}: out rel err {rel(out, ref):.2e}  lse max abs err {(lse-lse_ref).abs().max().item():.2e}", flush=True)
    if T == 1024:
        for _ in range(3): L.b200_flash_attn_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), Bs, T, H, D, nat.stream_ptr(dev))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(10): L.b200_flash_attn_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), Bs, T, H, D, nat.stream_ptr(dev))
        e1.record(); torch.cuda.synchronize()
        print(f"   fwd {e0.elapsed_time(e1)/10*1e3:.1f} us per call (B{Bs} H{H} T{T})")
