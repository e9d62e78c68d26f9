"""Multi-rank GPU checks, one process per GPU (launched by tests/test_gpu_multi.py or by hand:
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/multi_gpu_worker.py
Numerics of the peer-memory kernels vs NCCL / torch ops; replicas stay bit-identical."""
import copy
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

from b200ddp import DataParallel, SimpleNet
from b200ddp import _native as nat
from b200ddp.parallel import symm as symm_mod
from b200ddp.parallel.process_group import setup_distributed
from b200ddp.parallel.symm import SymmetricArena


def log(*a):
    if dist.get_rank() == 0:
        print(*a, flush=True)


def rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-20)).item()


def check_plain_allreduce(arena, rank, world, dev):
    for dtype in (torch.float32, torch.bfloat16):
        for n in (1024, 1 << 20, 12 * (1 << 20) + 64):
            buf = arena.alloc(n, dtype)
            for algo in (0, 1, 2) if arena.nvls else (0, 1):
                torch.manual_seed(1000 * rank + n % 997 + algo)
                src = torch.randn(n, device=dev).to(dtype)
                buf.copy_(src)
                want = src.float().clone()
                dist.all_reduce(want)  # NCCL fp32 sum of the (possibly bf16-rounded) inputs
                want = want * 0.5
                out = torch.empty_like(buf) if algo == 0 else None
                torch.cuda.synchronize(); dist.barrier()
                arena.allreduce_(buf, 0, n, algo=algo, scale=0.5, out=out)
                torch.cuda.synchronize()
                got = out if algo == 0 else buf
                tol = 1e-6 if dtype == torch.float32 else 4e-3
                err = rel(got, want)
                assert err < tol, (dtype, n, algo, err)
                # identical bits on every rank
                ref = got.clone(); dist.broadcast(ref, src=0)
                assert torch.equal(ref, got), ("ranks differ", dtype, n, algo)
                dist.barrier()
    log(f"plain all-reduce: one-shot, two-shot{' and NVLS multimem' if arena.nvls else ''} match NCCL "
        f"(fp32, bf16), ranks bit-identical")


def check_fused_optimizer(rank, world, dev, compute_dtype, two_shot, kind):
    """Engine on the b200 backend vs NCCL all-reduce + the torch-op optimizer."""
    symm_mod.ONE_SHOT_MAX_BYTES = 0 if two_shot else 1 << 40
    torch.manual_seed(5)
    net = SimpleNet().to(dev)
    eng = DataParallel(copy.deepcopy(net), optimizer=kind, lr=1e-2, backend="b200",
                       compute_dtype=compute_dtype, use_program=False)
    ref = DataParallel(copy.deepcopy(net), optimizer=kind, lr=1e-2, backend="nccl",
                       compute_dtype=compute_dtype, use_program=False)
    assert torch.equal(eng.master, ref.master)
    nvls = bool(two_shot and eng.symm.nvls and world > 2)
    nvls_bf16 = nvls and compute_dtype == torch.bfloat16
    # in-switch reduction order is the hardware's: fp32 differs in the last bits (amplified by
    # Adam's normalisation), bf16 additionally rounds the sum
    rtol, atol = (5e-3, 2e-3) if nvls_bf16 else ((1e-4, 1e-5) if nvls else (3e-6, 2e-6))
    for step in range(3):
        torch.manual_seed(77 * rank + step)
        g = torch.randn(eng.layout.total, device=dev)
        eng.grad.copy_(g); ref.grad.copy_(g)
        # reference: NCCL sum, then the torch-op update
        # reference sum in the kernels' order (rank 0..W-1, fp32): gather then add, so the P2P
        # paths can be held to a tight tolerance whatever algorithm NCCL would pick
        parts = [torch.empty_like(ref.grad) for _ in range(world)]
        dist.all_gather(parts, ref.grad)
        gsum = parts[0].float().clone()
        for part in parts[1:]:
            gsum += part.float()
        if nvls_bf16:  # the switch returns the fp32-accumulated sum rounded to bf16
            gsum = gsum.to(torch.bfloat16).float()
        ref.grad.copy_(gsum) if ref.grad.dtype == torch.float32 else None
        ref.optimizer.step_count += 1
        if ref.grad.dtype == torch.float32:
            ref.optimizer.step(0, ref.layout.total, grad_scale=1.0 / world, advance=False)
        else:
            tmp = ref.optimizer.grad
            ref.optimizer.grad = gsum
            ref.optimizer.step(0, ref.layout.total, grad_scale=1.0 / world, advance=False)
            ref.optimizer.grad = tmp
        eng.begin_backward()
        eng.finish_step()
        torch.cuda.synchronize()
    eng.consolidate_state()
    torch.cuda.synchronize()
    assert torch.allclose(eng.master, ref.master, rtol=rtol, atol=atol), rel(eng.master, ref.master)
    # in-switch bf16 reduction rounds more than one fp32 accumulation would (hardware property,
    # same class of error as an NCCL bf16 all-reduce): the moments get a looser bound
    # (the SGD momentum buffer holds raw gradient sums, so it sees the full in-switch rounding)
    assert torch.allclose(eng.optimizer.exp_avg, ref.optimizer.exp_avg, rtol=5e-2 if nvls_bf16 else rtol,
                          atol=5e-2 if nvls_bf16 else atol)
    if eng.shadow is not None:
        assert torch.equal(eng.shadow, eng.master.to(torch.bfloat16))
    got = [torch.empty_like(eng.master) for _ in range(world)]
    dist.all_gather(got, eng.master)
    assert all(torch.equal(got[0], t) for t in got), "replicas diverged"
    if eng.shadow is not None:
        gs = [torch.empty_like(eng.shadow) for _ in range(world)]
        dist.all_gather(gs, eng.shadow)
        assert all(torch.equal(gs[0], t) for t in gs), "shadow replicas diverged"
    sd = eng.optimizer.state_dict()
    if kind != "sgd":
        assert float(sd["state"][0]["step"]) == 3.0
    algo = ("NVLS" if nvls else "two-shot") if two_shot else "one-shot"
    log(f"fused all-reduce+{kind}: {algo} "
        f"{str(compute_dtype).split('.')[-1]} matches NCCL + torch-op update; replicas identical")


def check_program(rank, world, dev, sharded, fused=True):
    """The SimpleNet step program at world > 1.  `sharded` forces the two-shot / NVLS kernels
    (whose fp32 master lives on the shard owner); otherwise the size rule decides.  Replicas must
    compute with identical numbers EVERY step, biases included, without consolidate_state()."""
    symm_mod.ONE_SHOT_MAX_BYTES = 0 if sharded else 2 << 20
    if not fused:
        os.environ["B200DDP_NO_FUSED_MLP"] = "1"
    else:
        os.environ.pop("B200DDP_NO_FUSED_MLP", None)
    torch.manual_seed(100 + rank)  # different init per rank: ctor broadcast must fix it
    eng = DataParallel(SimpleNet().to(dev), backend="b200", compute_dtype=torch.bfloat16, lr=1e-3)
    os.environ.pop("B200DDP_NO_FUSED_MLP", None)
    assert eng.program is not None and (eng.program.fused is not None) == fused
    torch.manual_seed(200 + rank)
    x = torch.randn(64, 784, device=dev); y = torch.randint(0, 10, (64,), device=dev)
    torch.manual_seed(4242)
    xc = torch.randn(64, 784, device=dev)          # the same probe batch on every rank
    losses = []
    for step in range(40):
        losses.append(eng.train_step(x, y))
        if step in (0, 1, 7, 39):
            # biases read by forward must be current on every rank after every step
            lg = eng.program.logits(xc)
            allg = [torch.empty_like(lg) for _ in range(world)]
            dist.all_gather(allg, lg)
            assert all(torch.equal(allg[0], t) for t in allg), f"replica logits differ after step {step + 1}"
            for lay in eng.program.layers:
                bs = [torch.empty_like(lay["b"]) for _ in range(world)]
                dist.all_gather(bs, lay["b"].contiguous())
                assert all(torch.equal(bs[0], t) for t in bs), f"fp32 biases differ after step {step + 1}"
    torch.cuda.synchronize()
    l0, l1 = losses[0].item(), losses[-1].item()
    assert l1 < 0.6 * l0, (l0, l1)
    gs = [torch.empty_like(eng.shadow) for _ in range(world)]
    dist.all_gather(gs, eng.shadow)
    assert all(torch.equal(gs[0], t) for t in gs), "bf16 replicas diverged under the step program"
    algos = eng.program.fused.algos if fused else None
    eng.consolidate_state()  # two-shot/NVLS buckets keep the fp32 master on the shard owner
    got = [torch.empty_like(eng.master) for _ in range(world)]
    dist.all_gather(got, eng.master)
    assert all(torch.equal(got[0], t) for t in got), "replicas diverged under the step program"
    assert eng.optimizer.step_dev[0].item() == 40
    log(f"{'fused cluster-kernel' if fused else 'GEMM-per-layer'} step program at world {world} "
        f"({'sharded' if sharded else 'size rule'}; algos {algos}): loss {l0:.3f} -> {l1:.3f}, "
        f"replicas bit-identical every step")


def check_program_pool(rank, world, dev):
    """engine.train_steps (one native call, many steps per launch) == the same steps one by one."""
    symm_mod.ONE_SHOT_MAX_BYTES = 2 << 20
    torch.manual_seed(300)
    net = SimpleNet().to(dev)
    a = DataParallel(copy.deepcopy(net), backend="b200", compute_dtype=torch.bfloat16, lr=1e-3)
    b = DataParallel(copy.deepcopy(net), backend="b200", compute_dtype=torch.bfloat16, lr=1e-3)
    torch.manual_seed(400 + rank)
    xs = torch.randn(6, 64, 784, device=dev); ys = torch.randint(0, 10, (6, 64), device=dev)
    a.train_steps(xs, ys, 0, 50)
    for i in range(50):
        b.train_step(xs[i % 6], ys[i % 6])
    torch.cuda.synchronize()
    assert torch.equal(a.shadow, b.shadow), rel(a.shadow, b.shadow)
    assert a.optimizer.step_count == 50 and a.optimizer.step_dev[0].item() == 50
    log("train_steps over a device pool == 50 single steps (bit-identical), back-to-back no-exit-wait kernels")


def check_host_run(rank, world, dev):
    """host_stepper.run_pool (pinned host batches, the kernel waits for every step's own copy, many
    steps per launch) == engine.train_steps on the same batches resident on the device."""
    symm_mod.ONE_SHOT_MAX_BYTES = 2 << 20
    torch.manual_seed(500)
    net = SimpleNet().to(dev)
    a = DataParallel(copy.deepcopy(net), backend="b200", compute_dtype=torch.bfloat16, lr=1e-3)
    b = DataParallel(copy.deepcopy(net), backend="b200", compute_dtype=torch.bfloat16, lr=1e-3)
    torch.manual_seed(600 + rank)
    hx = torch.randn(5, 64, 784).pin_memory(); hy = torch.randint(0, 10, (5, 64)).pin_memory()
    a.train_steps(hx.to(dev), hy.to(dev), 0, 70)
    losses = b.host_stepper(slots=4).run_pool(hx, hy, 0, 70)
    torch.cuda.synchronize()
    assert len(losses) == 70 and all(l == l for l in losses)
    assert torch.equal(a.shadow, b.shadow), rel(a.shadow, b.shadow)
    assert abs(losses[-1] - a.last_loss().item()) == 0.0
    gs = [torch.empty_like(b.shadow) for _ in range(world)]
    dist.all_gather(gs, b.shadow)
    assert all(torch.equal(gs[0], t) for t in gs), "replicas diverged under host_stepper.run_pool"
    log("host_stepper.run_pool over pinned batches == train_steps on the device (bit-identical), replicas identical")


def check_buffer_broadcast(rank, world, dev):
    """K4: module buffers follow rank 0 before every training forward — our broadcast kernel on the
    peer-mapped buffer arenas (captured in the step graph), no NCCL call on the step."""
    import torch.nn as nn

    class Net(nn.Module):
        def __init__(self):
            super().__init__()
            self.fc = nn.Linear(32, 16)
            self.bn = nn.BatchNorm1d(16)
            self.out = nn.Linear(16, 4)

        def forward(self, x):
            return self.out(torch.relu(self.bn(self.fc(x))))

    torch.manual_seed(7)
    eng = DataParallel(Net().to(dev), backend="b200", compute_dtype=torch.float32, use_program=False)
    assert eng._buffers_in_graph and len(eng._buffer_arenas) == 2      # fp32 statistics + int64 counter
    for step in range(6):       # 2 eager steps, then the captured step is replayed
        torch.manual_seed(1000 * rank + step)     # different data per rank -> different local statistics
        x = torch.randn(24, 32, device=dev) * (1 + rank); y = torch.randint(0, 4, (24,), device=dev)
        # what rank 0 holds BEFORE the step is what every rank must normalise with during it
        eng.train_step(x, y)
        torch.cuda.synchronize()
    assert len(eng._step_graphs) == 1
    # after a step every rank holds "rank 0's pre-step buffers updated with ITS OWN batch": they differ ...
    rm = eng.module.bn.running_mean.clone()
    allm = [torch.empty_like(rm) for _ in range(world)]
    dist.all_gather(allm, rm)
    assert not torch.equal(allm[0], allm[-1])
    # ... and the next forward starts from rank 0's again
    eng._sync_buffers()
    torch.cuda.synchronize()
    rm = eng.module.bn.running_mean.clone()
    nb = eng.module.bn.num_batches_tracked.clone()
    dist.all_gather(allm, rm)
    assert all(torch.equal(allm[0], t) for t in allm)
    alln = [torch.empty_like(nb) for _ in range(world)]
    dist.all_gather(alln, nb)
    assert all(torch.equal(alln[0], t) for t in alln) and int(nb) == 6
    got = [torch.empty_like(eng.master) for _ in range(world)]
    dist.all_gather(got, eng.master)
    assert all(torch.equal(got[0], t) for t in got), "replicas diverged with broadcast buffers"
    log("module buffers (fp32 statistics, int64 counter) broadcast from rank 0 by our kernel inside the step graph")


def check_autograd_overlap(rank, world, dev):
    """Generic autograd path with bucket hooks on the b200 backend vs the nccl backend."""
    symm_mod.ONE_SHOT_MAX_BYTES = 64 * 1024  # mix of one-shot and two-shot buckets
    torch.manual_seed(9)
    net = SimpleNet(dropout=0.0).to(dev)
    a = DataParallel(copy.deepcopy(net), backend="b200", compute_dtype=torch.float32,
                     bucket_cap_mb=0.25, first_bucket_mb=0.01, use_program=False)
    b = DataParallel(copy.deepcopy(net), backend="nccl", compute_dtype=torch.float32,
                     bucket_cap_mb=0.25, first_bucket_mb=0.01, use_program=False)
    assert len(a.layout.buckets) >= 3
    for step in range(8):   # 2 eager steps, then the captured whole-step graph is replayed
        torch.manual_seed(31 * rank + step)
        x = torch.randn(32, 784, device=dev); y = torch.randint(0, 10, (32,), device=dev)
        la = a.train_step(x, y); lb = b.train_step(x, y)
    torch.cuda.synchronize()
    a.consolidate_state()
    assert len(a._step_graphs) == 1 and a.optimizer.step_count == 8
    assert abs(la.item() - lb.item()) < 1e-4
    assert torch.allclose(a.master, b.master, rtol=1e-4, atol=1e-6), rel(a.master, b.master)
    log(f"autograd path, {len(a.layout.buckets)} buckets overlapped with backward, eager then CUDA-graph replay: "
        f"b200 == nccl")


def main():
    rank, world, local = setup_distributed("nccl")
    dev = torch.device("cuda", local)
    assert nat.available()
    arena = SymmetricArena(dev, rank, world)
    log(f"world {world}: symmetric allocator {'own VMM (cuMemCreate + fd exchange)' if arena.vmm else 'cudaMalloc + CUDA IPC'}; "
        f"NVLS multicast memory {'available (own cuMulticast object)' if arena.nvls else 'unavailable: ' + getattr(arena, '_nvls_error', 'disabled')}")
    check_plain_allreduce(arena, rank, world, dev)
    for kind in ("adam", "adamw", "sgd"):
        for dtype in (torch.float32, torch.bfloat16):
            for two_shot in (False, True):
                check_fused_optimizer(rank, world, dev, dtype, two_shot, kind)
    for sharded in (False, True):
        check_program(rank, world, dev, sharded, fused=True)
    check_program(rank, world, dev, True, fused=False)
    check_program_pool(rank, world, dev)
    check_host_run(rank, world, dev)
    check_buffer_broadcast(rank, world, dev)
    check_autograd_overlap(rank, world, dev)
    dist.barrier()
    log("MULTI_GPU_OK")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
