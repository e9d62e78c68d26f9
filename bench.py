#!/usr/bin/env python3
"""Headline benchmark: train samples/sec of the reference's model (SimpleNet, batch
64 per rank, Adam) under data parallelism on N B200s — BASELINE.json's metric.

    python bench.py --gpus N --steps K --warmup W            # this framework
    python bench.py --impl reference --gpus N ...            # unmodified reference (baseline/_ref)

For N > 1 launch with ``python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N``.
Rank 0 prints ONE JSON line.  `value` is device-timed (CUDA events, max over
ranks) over exactly K steps after W warm-up steps, inputs rotating through an
on-device pool larger than L2.  `e2e` is the same metric through the public API
with every step's batch copied from pinned host memory and every step's loss
read back to the host.
"""
from __future__ import annotations

import argparse
import importlib.util
import json
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# rank 0 prints exactly one line: keep NCCL's own banner ("NCCL version ...") off stdout
if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
    os.environ["NCCL_DEBUG"] = "WARN"

L2_BYTES = 126 * 1024 * 1024
IN_FEATURES, CLASSES = 784, 10


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=2000)
    p.add_argument("--warmup", type=int, default=200)
    p.add_argument("--impl", choices=["b200ddp", "reference"], default="b200ddp")
    p.add_argument("--batch-size", type=int, default=64, help="per rank (reference default)")
    p.add_argument("--backend", default="b200", choices=["b200", "nccl"])
    p.add_argument("--ref-backend", default="nccl", choices=["nccl", "gloo"],
                   help="reference arm: nccl = one-word backend delta (SURVEY §6); gloo = as shipped")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--dtype", choices=["bf16", "fp32"], default="bf16",
                   help="this framework's arm only: fp32 = the reference arm's precision (torch autograd with fp32 "
                        "cuBLAS GEMMs, TF32 off, under the engine's whole-step graph and fused all-reduce + Adam kernel); "
                        "bf16 (BASELINE.json's config) = the fused tcgen05 step kernel")
    return p.parse_args()


def init_dist(args, backend: str):
    if "RANK" not in os.environ:
        if args.gpus != 1:
            raise SystemExit("--gpus > 1 must be launched with torch.distributed.run")
        os.environ.update(RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if backend == "nccl":
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        dist.init_process_group(backend)
    return dist.get_rank(), dist.get_world_size(), torch.device("cuda", local)


def make_pool(bsz: int, device, pinned: bool, seed: int):
    """Distinct synthetic batches totalling more than L2 (so inputs never come from L2)."""
    per = bsz * IN_FEATURES * 4 + bsz * 8
    n = L2_BYTES // per + 64
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, bsz, IN_FEATURES, generator=g)
    y = torch.randint(0, CLASSES, (n, bsz), generator=g)
    if pinned:
        return x.pin_memory(), y.pin_memory(), n
    return x.to(device), y.to(device), n


def max_over_ranks(v: float, device) -> float:
    t = torch.tensor([v], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _step_path(engine) -> str:
    prog = engine.program
    if prog is None:
        return ("torch autograd forward / backward (ATen + cuBLAS in the compute dtype, TF32 off) inside the engine's "
                "whole-step CUDA graph, fused all-reduce + Adam kernel of this framework")
    if prog.fused is None:
        return "GEMM-per-layer CUDA graph"
    k = prog.fused.update_clusters
    if k:
        return (f"one kernel: 8-CTA compute cluster (fwd+CE+bwd) + {k} update clusters (all-reduce + Adam) in the "
                "same launch, up to 256 steps per launch with counter hand-overs (engine.train_steps; the end-to-end "
                "arm feeds the same launches from pinned host batches, host_stepper.run_pool)")
    return "fused cluster kernel (fwd+CE+bwd) + fused all-reduce/Adam kernel, launched from C (engine.train_steps)"


def timed(fn_run, steps: int, warmup: int, device, sample_clocks: bool, fn_end=None):
    """fn_run(first, count) runs steps [first, first + count); fn_end() (optional) drains whatever
    the steps left pending, still inside the timed region.  Returns (ms total max-over-ranks,
    clocks dict).  Clocks: NVML is polled every 2 ms from the start of the warm-up (the GPU is
    already under load) and read once more the moment the timed steps have been submitted, so
    even a 20-step region (< 1 ms) has a reading taken while (or right after) it runs."""
    from b200ddp.utils.profiling import ClockSampler
    sampler = ClockSampler(index=device.index or 0, period_ms=2).start() if sample_clocks else None
    fn_run(0, warmup)
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    if sampler:
        sampler.rows.clear()          # keep only readings from the timed region on
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn_run(warmup, steps)
    if fn_end is not None:
        fn_end()
    e1.record()               # enqueued before any NVML call: a reading costs the host milliseconds
    if sampler:
        sampler.sample_now()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None
    dist.barrier()
    torch.cuda.synchronize()
    return max_over_ranks(ms, device), clocks


# ------------------------------------------------------------------ this framework
def run_ours(args):
    rank, world, device = init_dist(args, "nccl")
    from b200ddp import DataParallel, SimpleNet
    from b200ddp import _native as nat
    from b200ddp._native import bindings
    assert nat.available()
    torch.manual_seed(1234)
    backend = args.backend if world > 1 else "b200"
    fp32 = args.dtype == "fp32"
    engine = DataParallel(SimpleNet().to(device), optimizer="adam", lr=1e-3, backend=backend,
                          compute_dtype=torch.float32 if fp32 else torch.bfloat16)
    bsz = args.batch_size
    xs, ys, n = make_pool(bsz, device, pinned=False, seed=100 + rank)

    # The public multi-step call: K optimisation steps over a device-resident pool of batches.
    # With the fused MLP session it is one native call that issues two launches per step, so the
    # device-timed number is not a measurement of eight Python interpreters.
    def run(first, count):
        engine.train_steps(xs, ys, first, count)

    launches_before = [0]

    def run_counted(first, count):
        if first == args.warmup:
            launches_before[0] = bindings.launch_count()
        run(first, count)

    ms, clocks = timed(run_counted, args.steps, args.warmup, device, sample_clocks=(rank == 0))
    gpu_launches = bindings.launch_count() - launches_before[0]   # counted in C at every launch site
    launch_unit = "kernel launches of this framework (counted in C at every launch site)"
    if fp32 and gpu_launches == 0:
        # the fp32 arm replays the engine's captured step: one graph launch per step, holding ATen /
        # cuBLAS kernels and this framework's fused Adam kernel (captured launches are not re-counted)
        gpu_launches = args.steps
        launch_unit = "CUDA graph launches (one captured step each: ATen / cuBLAS kernels + this framework's fused Adam kernel)"
    per_step = gpu_launches / max(args.steps, 1)
    value = world * bsz * args.steps / (ms / 1e3)

    e2e = None
    if not args.no_e2e:
        del xs, ys
        hx, hy, n = make_pool(bsz, device, pinned=True, seed=200 + rank)
        # The public end-to-end API: pinned host batch in, Python float loss out.  Every step
        # copies its inputs host->device and its loss device->host (read one step late, so a
        # step is always queued behind the running one); the per-step host work is one native
        # call (csrc/runner.cu).
        stepper = engine.host_stepper(slots=4)
        losses = []
        use_run = hasattr(stepper, "run_pool") and os.environ.get("B200DDP_BENCH_E2E_STEP") is None

        def e2e_step(first, count):
            if use_run:     # one native call for `count` steps; every step still has its own H2D and its own loss
                losses.extend(stepper.run_pool(hx, hy, first, count))
                return
            for i in range(first, first + count):
                prev = stepper.step(hx[i % n], hy[i % n])
                if prev is not None:
                    losses.append(prev)

        def e2e_end():
            if not use_run:
                losses.append(stepper.finish())     # the last step's loss, inside the timed region

        torch.cuda.synchronize()
        t_ms, _ = timed(e2e_step, args.steps, args.warmup, device, sample_clocks=False, fn_end=e2e_end)
        assert len(losses) == args.steps + args.warmup, (len(losses), args.steps + args.warmup)
        host_loss = losses
        e2e = {"value": world * bsz * args.steps / (t_ms / 1e3), "unit": "samples/s",
               "ms_per_step": t_ms / args.steps,
               "h2d_bytes_per_step": bsz * IN_FEATURES * 4 + bsz * 8,
               "d2h_bytes_per_step": 8,   # the loss and its sequence number, stored to pinned host memory
               "host_loss_is_finite": all(l == l for l in host_loss),
               "final_loss": float(host_loss[-1]),
               "api": ("DataParallel.host_stepper().run_pool(pinned_x_pool, pinned_y_pool, first, count): per step one "
                       "H2D of that step's batch and one loss to the host, up to 256 steps per kernel launch")
               if use_run and not fp32 else "DataParallel.host_stepper().step(pinned_x, pinned_y) (one native call per step)"}

    if rank == 0:
        out = {
            "metric": "train samples/sec (device-timed, max over ranks)", "value": value,
            "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": args.dtype, "data": "synthetic (random-init weights)",
            "impl": "b200ddp",
            "config": {"model": "SimpleNet 784-256-256-10 (reference train.py model)",
                       "global_batch": world * bsz, "per_rank_batch": bsz, "seq_len": None,
                       "parallelism": f"dp{world}", "optimizer": "adam", "backend": backend,
                       "grad_wire_dtype": "fp32", "master_weights": "fp32",
                       "step_path": _step_path(engine),
                       "l2": f"inputs rotate through a {n}-batch pool "
                             f"({n * (bsz * IN_FEATURES * 4 + bsz * 8) / 2**20:.0f} MiB) larger than L2"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": gpu_launches,
            "launches_per_step": per_step,
            # one launch of the fused step kernel runs many whole steps (forward, loss, backward,
            # all-reduce, Adam each); kernels_per_step counts the kernels a step's work runs in
            "steps_per_launch": (args.steps / gpu_launches) if gpu_launches else None,
            "launch_unit": launch_unit,
        }
        print(json.dumps(out), flush=True)
    dist.destroy_process_group()


# ------------------------------------------------------------------ reference arm
def load_reference():
    path = os.path.join(ROOT, "baseline", "_ref", "train.py")
    if not os.path.exists(path):
        return None, f"{path} missing (run __graft_entry__.build() where /root/reference is mounted)"
    spec = importlib.util.spec_from_file_location("reference_train", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod, None


def run_reference(args):
    ref, why = load_reference()
    if ref is None:
        print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)
        return
    import torch.nn as nn
    import torch.optim as optim
    from torch.nn.parallel import DistributedDataParallel as DDP
    if args.ref_backend == "gloo":
        if "RANK" not in os.environ:
            os.environ.update(RANK="0", WORLD_SIZE="1", LOCAL_RANK="0", MASTER_ADDR="127.0.0.1",
                              MASTER_PORT="29512")
        rank, world, local = ref.setup_distributed()          # stock: gloo, even on GPU
        device = ref.get_device(local)
    else:
        rank, world, device = init_dist(args, "nccl")          # the one-word delta of SURVEY §6
    torch.manual_seed(1234)
    model = ref.SimpleNet().to(device)                          # train.py:232
    model = DDP(model)                                          # train.py:233
    optimizer = optim.Adam(model.parameters(), lr=1e-3)         # train.py:249
    criterion = nn.CrossEntropyLoss()                           # train.py:250
    bsz = args.batch_size

    # --- device-timed: the reference's own train_epoch over pre-staged device batches
    xs, ys, n = make_pool(bsz, device, pinned=False, seed=100 + rank)

    class Batches:  # stands in for the DataLoader: same iteration protocol, data already on device
        def __init__(self, lo, count):
            self.lo, self.count = lo, count

        def __len__(self):
            return self.count

        def __iter__(self):
            for i in range(self.lo, self.lo + self.count):
                yield xs[i % n], ys[i % n]

    import logging
    logging.getLogger("reference_train").setLevel(logging.WARNING)
    from b200ddp.utils.profiling import ClockSampler
    ref.train_epoch(model, Batches(0, args.warmup), optimizer, criterion, device, 0, rank)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    sampler = ClockSampler(index=device.index or 0, period_ms=2).start() if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ref.train_epoch(model, Batches(args.warmup, args.steps), optimizer, criterion, device, 1, rank)
    e1.record()
    if sampler:
        sampler.sample_now()
    torch.cuda.synchronize()
    ms = max_over_ranks(e0.elapsed_time(e1), device)
    clocks = sampler.stop() if sampler else None
    value = world * bsz * args.steps / (ms / 1e3)

    # --- end to end: stock create_data_loader (DataLoader, 2 workers, pinned) + train_epoch
    e2e = None
    if not args.no_e2e:
        del xs, ys
        warm_ds = ref.SyntheticDataset(num_samples=args.warmup * bsz * world)
        ds = ref.SyntheticDataset(num_samples=args.steps * bsz * world)
        warm_loader, _ = ref.create_data_loader(warm_ds, bsz, rank, world)
        loader, _ = ref.create_data_loader(ds, bsz, rank, world)
        assert len(loader) == args.steps
        ref.train_epoch(model, warm_loader, optimizer, criterion, device, 2, rank)
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ref.train_epoch(model, loader, optimizer, criterion, device, 3, rank)
        e1.record(); torch.cuda.synchronize()
        t_ms = max_over_ranks(e0.elapsed_time(e1), device)
        e2e = {"value": world * bsz * args.steps / (t_ms / 1e3), "unit": "samples/s",
               "ms_per_step": t_ms / args.steps,
               "h2d_bytes_per_step": bsz * IN_FEATURES * 4 + bsz * 8, "d2h_bytes_per_step": 4}
    if rank == 0:
        print(json.dumps({
            "impl": "reference", "metric": "train samples/sec (device-timed, max over ranks)",
            "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": {"model": "SimpleNet 784-256-256-10 (reference train.py model)",
                       "global_batch": world * bsz, "per_rank_batch": bsz, "seq_len": None,
                       "parallelism": f"dp{world}", "optimizer": "adam",
                       "backend": args.ref_backend,
                       "path": "baseline/_ref/train.py unmodified: SimpleNet, DDP(model), "
                               "optim.Adam, train_epoch()"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": 0}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
