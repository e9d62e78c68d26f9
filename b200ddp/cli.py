"""Command line — superset of the reference's (reference ``train.py:213-221``).

The six reference flags keep their names, types and defaults.  The rest are new
and default to the product configuration on a B200 box.
"""
from __future__ import annotations

import argparse


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(prog="b200ddp.train",
                                description="Blackwell-native data-parallel trainer")
    # --- reference flags (train.py:213-221) ---
    p.add_argument("--epochs", type=int, default=10)
    p.add_argument("--batch-size", type=int, default=64)
    p.add_argument("--lr", type=float, default=0.001)
    p.add_argument("--num-samples", type=int, default=10000)
    p.add_argument("--checkpoint-dir", type=str, default="./checkpoints")
    p.add_argument("--resume", type=str, default=None)
    # --- new ---
    p.add_argument("--backend", choices=["auto", "b200", "nccl", "gloo"], default="auto",
                   help="gradient reduction path: b200 = fused peer-memory kernels, "
                        "nccl = NCCL baseline, gloo = CPU plumbing")
    p.add_argument("--model", choices=["simplenet", "gpt2", "resnet50"], default="simplenet")
    p.add_argument("--dtype", choices=["fp32", "bf16"], default=None,
                   help="compute dtype (default: bf16 on CUDA, fp32 on CPU)")
    p.add_argument("--wire-dtype", choices=["fp32", "bf16"], default=None,
                   help="gradient dtype on the wire (default: the compute dtype)")
    p.add_argument("--optimizer", choices=["adam", "adamw", "sgd"], default="adam")
    p.add_argument("--weight-decay", type=float, default=None)
    p.add_argument("--momentum", type=float, default=0.9)
    p.add_argument("--bucket-cap-mb", type=float, default=25.0)
    p.add_argument("--seq-len", type=int, default=1024, help="gpt2 only")
    p.add_argument("--synthetic-on-device", action="store_true",
                   help="keep the synthetic data set in HBM; batches are device gathers")
    p.add_argument("--seed", type=int, default=None,
                   help="seed model init and data (default: unseeded like the reference)")
    p.add_argument("--log-interval", type=int, default=10)
    p.add_argument("--resume-next-epoch", action="store_true",
                   help="resume at stored epoch + 1 instead of re-running the stored epoch")
    p.add_argument("--no-program", action="store_true",
                   help="disable the fused CUDA-graph step program (debug)")
    p.add_argument("--profile", action="store_true", help="emit NVTX ranges per step phase")
    p.add_argument("--inject-fault", type=str, default=None, metavar="RANK:STEP",
                   help="test hook: the given rank exits non-zero at the given global step")
    return p


def parse_args(argv=None) -> argparse.Namespace:
    return build_parser().parse_args(argv)
