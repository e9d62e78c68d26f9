#!/usr/bin/env python3
"""Extension configs of BASELINE.json: GPT-2 small (bf16, seq 1024) and ResNet-50 (224x224),
data-parallel on N GPUs, both arms.  Same timing rules as /bench.py (CUDA events, max over
ranks, W warm-up steps, inputs rotating through an on-device pool).

    python bench/bench_models.py --model gpt2 --gpus 1 --batch-size 8 --steps 10 --warmup 3
    torchrun --nproc-per-node 8 ... bench/bench_models.py --model gpt2 --gpus 8 [--impl reference]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--model", choices=["gpt2", "resnet50"], default="gpt2")
    p.add_argument("--impl", choices=["b200ddp", "reference"], default="b200ddp")
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--batch-size", type=int, default=8, help="per rank")
    p.add_argument("--seq-len", type=int, default=1024)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--backend", default="b200", choices=["b200", "nccl"])
    p.add_argument("--pool", type=int, default=8, help="distinct batches in the on-device pool")
    return p.parse_args()


def init(args):
    if "RANK" not in os.environ:
        os.environ.update(RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29521")
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return dist.get_rank(), dist.get_world_size(), torch.device("cuda", local)


def pool(args, rank, dev):
    g = torch.Generator().manual_seed(50 + rank)
    if args.model == "gpt2":
        t = torch.randint(0, 50257, (args.pool, args.batch_size, args.seq_len + 1), generator=g)
        return t[..., :-1].contiguous().to(dev), t[..., 1:].contiguous().to(dev)
    x = torch.randn(args.pool, args.batch_size, 3, 224, 224, generator=g).to(dev)
    y = torch.randint(0, 1000, (args.pool, args.batch_size), generator=g).to(dev)
    return x, y


def timed(step, args, dev):
    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(args.warmup + i)
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    args = parse()
    rank, world, dev = init(args)
    xs, ys = pool(args, rank, dev)
    torch.manual_seed(7)
    losses = []
    if args.impl == "b200ddp":
        from b200ddp import DataParallel
        if args.model == "gpt2":
            from b200ddp.models.gpt2 import GPT2, GPT2Config
            model = GPT2(GPT2Config(n_positions=args.seq_len)).to(dev)
        else:
            from b200ddp.models.resnet50 import resnet50
            model = resnet50().to(dev)
        backend = args.backend if world > 1 else "b200"
        eng = DataParallel(model, optimizer="adamw", lr=3e-4, weight_decay=0.01, backend=backend,
                           compute_dtype=torch.bfloat16)

        def step(i):
            losses.append(eng.train_step(xs[i % args.pool], ys[i % args.pool]))
        extra = {"buckets": len(eng.layout.buckets), "backend": backend,
                 "wire_dtype": str(eng.grad.dtype).split(".")[-1]}
    else:
        from torch.nn.parallel import DistributedDataParallel as DDP
        if args.model == "gpt2":
            from transformers import GPT2Config as HFConfig, GPT2LMHeadModel
            cfg = HFConfig(n_positions=args.seq_len, resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0)
            model = GPT2LMHeadModel(cfg).to(dev)
        else:
            import torchvision
            model = torchvision.models.resnet50().to(dev)
        model = DDP(model)                     # the reference's wrapper call (train.py:233)
        opt = torch.optim.AdamW(model.parameters(), lr=3e-4, weight_decay=0.01)
        crit = torch.nn.CrossEntropyLoss()

        def step(i):                           # the reference's loop body (train.py:135-141) + autocast
            x, y = xs[i % args.pool], ys[i % args.pool]
            opt.zero_grad()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                if args.model == "gpt2":
                    logits = model(x).logits
                    loss = crit(logits.view(-1, logits.size(-1)).float(), y.view(-1))
                else:
                    loss = crit(model(x).float(), y)
            loss.backward()
            opt.step()
            losses.append(loss.detach())
        extra = {"backend": "nccl", "amp": "autocast bf16, fp32 params/grads"}
    ms = timed(step, args, dev)
    unit_per_sample = args.seq_len if args.model == "gpt2" else 1
    samples = world * args.batch_size * args.steps / (ms / 1e3)
    if rank == 0:
        print(json.dumps({
            "impl": args.impl, "model": args.model, "n_gpus": world, "per_rank_batch": args.batch_size,
            "seq_len": args.seq_len if args.model == "gpt2" else None,
            "metric": "train samples/sec (device-timed, max over ranks)", "value": samples,
            "tokens_per_s": samples * unit_per_sample if args.model == "gpt2" else None,
            "ms_per_step": ms / args.steps, "steps": args.steps, "warmup": args.warmup,
            "first_loss": float(losses[0]), "last_loss": float(losses[-1]), "dtype": "bf16",
            "data": "synthetic", **extra}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
