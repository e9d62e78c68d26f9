"""Per-layer gradient error of the fused MLP program vs (a) fp32 autograd with bf16 weights,
(b) an emulation that rounds to bf16 at the same points as the program."""
import copy, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from b200ddp.models.simplenet import SimpleNet
from b200ddp.parallel.engine import DataParallel

def rel(a, b): return ((a.float()-b.float()).norm()/b.float().norm()).item()
r = lambda t: t.to(torch.bfloat16).float()
dev = torch.device("cuda")
torch.manual_seed(7)
for bsz in (64, 34, 256):
    net = SimpleNet(dropout=0.0).to(dev)
    W = [r(net.layers[i].weight.detach()) for i in (0, 3, 6)]
    b = [net.layers[i].bias.detach().clone() for i in (0, 3, 6)]
    eng = DataParallel(net, backend="b200", compute_dtype=torch.bfloat16, lr=1e-3)
    x = torch.randn(bsz, 784, device=dev); y = torch.randint(0, 10, (bsz,), device=dev)
    xr = r(x)
    # (a) fp32 chain
    Wa = [w.clone().requires_grad_(True) for w in W]; ba = [t.clone().requires_grad_(True) for t in b]
    h1 = torch.relu(xr @ Wa[0].t() + ba[0]); h2 = torch.relu(h1 @ Wa[1].t() + ba[1])
    la = F.cross_entropy(h2 @ Wa[2].t() + ba[2], y); la.backward()
    # (b) emulated rounding
    h1e = r(torch.relu(xr @ W[0].t() + b[0])); h2e = r(torch.relu(h1e @ W[1].t() + b[1]))
    lg = h2e @ W[2].t() + b[2]
    dl = r((torch.softmax(lg, 1) - F.one_hot(y, 10).float()) / bsz)
    gW3 = dl.t() @ h2e; gb3 = dl.sum(0)
    dz2 = r((dl @ W[2]) * (h2e > 0)); gW2 = dz2.t() @ h1e; gb2 = dz2.sum(0)
    dz1 = r((dz2 @ W[1]) * (h1e > 0)); gW1 = dz1.t() @ xr; gb1 = dz1.sum(0)
    emu = {"layers.0.weight": gW1, "layers.0.bias": gb1, "layers.3.weight": gW2, "layers.3.bias": gb2,
           "layers.6.weight": gW3, "layers.6.bias": gb3}
    fp = dict(zip(emu.keys(), [Wa[0].grad, ba[0].grad, Wa[1].grad, ba[1].grad, Wa[2].grad, ba[2].grad]))
    loss = eng.train_step(x, y); torch.cuda.synchronize()
    print(f"bsz={bsz} loss ours {loss.item():.5f} fp32 {la.item():.5f}")
    for s in eng.layout.slots:
        g = eng.grad[s.offset:s.offset+s.numel].view(s.shape)
        print(f"  {s.name:18s} vs fp32 {rel(g, fp[s.name]):.5f}  vs emu {rel(g, emu[s.name]):.6f}  emu vs fp32 {rel(emu[s.name], fp[s.name]):.5f}")
