#!/usr/bin/env python3
"""Gradient all-reduce bandwidth sweep, 1 KB - 1 GB, ours vs NCCL (BASELINE.json config 5).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/allreduce_sweep.py

For every size: NCCL `dist.all_reduce` (the reference's collective once its backend is
nccl), our one-shot and two-shot peer-memory kernels (plain all-reduce),
and the fused two-shot all-reduce + 1/W + Adam + bf16 parameter push.  Device-timed
with CUDA events, max over ranks, buffers rotated so consecutive iterations touch
different memory (a set larger than L2 for the small sizes).

Reported per row: time, bus bandwidth S/t * 2(W-1)/W, and the fraction of the NVLink roofline,
each algorithm against ITS OWN per-direction link bytes:
    two-shot (P2P)   2 S (W-1)/W
    NVLS             S (1 + 1/W)          (one switch-reduced pull of the shard, one multicast push)
Every fraction is quoted twice: against the measured peer-copy rate (770 GB/s per direction,
B200_PROFILING.md) and against the nominal NVLink 5 rate (900 GB/s, BASELINE.json).  For the
fused kernel the roofline is max(link time, HBM time): link bytes per direction are
max(out, in) with, for P2P two-shot, S_grad (W-1)/W pulled + S_param (W-1)/W pushed, and for NVLS
out = S_grad (W-1)/W + S_param / W, in = S_grad / W + S_param (W-1)/W; HBM is 28 B/param / W
(sharded optimizer) over the measured copy bandwidth.
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

LINK_GBS = 770.0        # measured peer copy, per direction
LINK_NOMINAL_GBS = 900.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dtype", choices=["fp32", "bf16"], default="bf16")
    ap.add_argument("--max-bytes", type=int, default=1 << 30)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--out", type=str, default="")
    args = ap.parse_args()

    from b200ddp import _native as nat
    from b200ddp._native import bindings
    from b200ddp.parallel.process_group import setup_distributed
    from b200ddp.parallel.symm import SymmetricArena
    rank, world, local = setup_distributed("nccl")
    dev = torch.device("cuda", local)
    assert nat.available()
    try:
        hbm = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        hbm = 6650.0
    dtype = torch.bfloat16 if args.dtype == "bf16" else torch.float32
    esz = 2 if dtype == torch.bfloat16 else 4
    arena = SymmetricArena(dev, rank, world)

    sizes = []
    s = 1024
    while s <= args.max_bytes:
        sizes.append(s)
        s *= 4
    if sizes[-1] != args.max_bytes:
        sizes.append(args.max_bytes)

    def time_it(fn, iters):
        """`iters` back-to-back calls captured in one CUDA graph (no host launch overhead in
        the number, for NCCL and for us alike); replayed once warm, then timed."""
        fn(0)
        torch.cuda.synchronize(); dist.barrier()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for i in range(iters):
                fn(i)
        g.replay()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        del g
        return float(t.item())

    rows = []
    for nbytes in sizes:
        n = max(64, nbytes // esz // 64 * 64)
        nbytes = n * esz
        # rotate through `slots` disjoint regions so small sizes do not sit in L2
        slots = max(2, min(64, (192 << 20) // nbytes)) if nbytes < (96 << 20) else 2
        buf = arena.alloc(n * slots, dtype)
        buf.copy_(torch.randn(n * slots, device=dev).to(dtype))
        out = torch.empty(n, dtype=dtype, device=dev)
        nccl_bufs = [torch.randn(n, device=dev).to(dtype) for _ in range(slots)]
        iters = args.iters if nbytes <= (64 << 20) else max(5, args.iters // 4)

        t_nccl = time_it(lambda i: dist.all_reduce(nccl_bufs[i % slots]), iters)
        t_one = None
        if nbytes <= (32 << 20):
            t_one = time_it(lambda i: arena.allreduce_(buf, (i % slots) * n, n, algo=0, scale=1.0, out=out), iters)
        t_two = time_it(lambda i: arena.allreduce_(buf, (i % slots) * n, n, algo=1, scale=1.0 / world), iters)
        t_nvls = None
        if arena.nvls:
            t_nvls = time_it(lambda i: arena.allreduce_(buf, (i % slots) * n, n, algo=2, scale=1.0 / world), iters)

        # fused: gradient bucket of n elements -> Adam on the owner's shard -> bf16 parameter push
        t_fused = None
        if nbytes <= (512 << 20):
            master = torch.zeros(n, dtype=torch.float32, device=dev)
            m = torch.zeros_like(master); v = torch.zeros_like(master)
            shadow = arena.alloc(n, torch.bfloat16)
            gbuf = arena.alloc(n, dtype)
            gbuf.copy_(torch.randn(n, device=dev).to(dtype) * 1e-3)
            c = arena.comm(gbuf, shadow)
            hp = bindings.OptHParams(0, 1e-3, 0.9, 0.999, 1e-8, 0.0, 0.0, 1.0 / world)
            import ctypes as C

            def fused(i):
                nat.check(arena.L.b200_allreduce_optimizer(
                    C.byref(c), 2 if (arena.nvls and world > 2) else 1, C.byref(hp), 0, n, int(dtype == torch.bfloat16), master.data_ptr(),
                    m.data_ptr(), v.data_ptr(), shadow.data_ptr(), 1, None, 10 + i, 0,
                    nat.stream_ptr(dev)), "fused")
            t_fused = time_it(fused, iters)
            del master, m, v

        f = (world - 1) / world
        nvls_fused = arena.nvls and world > 2
        bytes_two = 2 * nbytes * f                       # per direction, P2P two-shot
        bytes_nvls = nbytes * (1 + 1 / world)            # per direction, NVLS
        sp = n * 2                                       # bf16 parameters pushed by the fused kernel
        if nvls_fused:
            bytes_fused = max(nbytes * f + sp / world, nbytes / world + sp * f)
        else:
            bytes_fused = nbytes * f + sp * f
        hbm_ms = (n / world) * 28 / (hbm * 1e9) * 1e3

        def frac(link_bytes, t_ms, gbs, floor_ms=0.0):
            return max(link_bytes / (gbs * 1e9) * 1e3, floor_ms) / t_ms if t_ms else None

        row = {"bytes": nbytes, "dtype": args.dtype, "world": world,
               "nccl_ms": t_nccl, "oneshot_ms": t_one, "twoshot_ms": t_two, "nvls_ms": t_nvls,
               "nvls_busbw_gbs": (nbytes * 2 * f / t_nvls / 1e6) if t_nvls else None,
               "fused_adam_ms": t_fused, "fused_algo": "nvls" if nvls_fused else "two-shot",
               "nccl_busbw_gbs": nbytes * 2 * f / t_nccl / 1e6,
               "twoshot_busbw_gbs": nbytes * 2 * f / t_two / 1e6,
               "oneshot_busbw_gbs": (nbytes * 2 * f / t_one / 1e6) if t_one else None,
               "twoshot_roofline_frac": frac(bytes_two, t_two, LINK_GBS),
               "twoshot_roofline_frac_nominal": frac(bytes_two, t_two, LINK_NOMINAL_GBS),
               "nvls_roofline_frac": frac(bytes_nvls, t_nvls, LINK_GBS),
               "nvls_roofline_frac_nominal": frac(bytes_nvls, t_nvls, LINK_NOMINAL_GBS),
               "nccl_roofline_frac": frac(bytes_two, t_nccl, LINK_GBS),
               "fused_roofline_frac": frac(bytes_fused, t_fused, LINK_GBS, hbm_ms),
               "fused_roofline_frac_nominal": frac(bytes_fused, t_fused, LINK_NOMINAL_GBS, hbm_ms)}
        rows.append(row)
        if rank == 0:
            one = f"{t_one * 1e3:9.1f}" if t_one else "        -"
            fu = (f"{t_fused * 1e3:9.1f} ({row['fused_roofline_frac']:.2f} / {row['fused_roofline_frac_nominal']:.2f})"
                  if t_fused else "        -")
            nv = (f"{t_nvls * 1e3:9.1f} ({row['nvls_roofline_frac']:.2f} / {row['nvls_roofline_frac_nominal']:.2f} of its roofline)"
                  if t_nvls else "        -")
            print(f"{nbytes:>11d} B  nccl {t_nccl * 1e3:9.1f} us  one-shot {one} us  two-shot "
                  f"{t_two * 1e3:9.1f} us ({row['twoshot_busbw_gbs']:6.1f} GB/s bus, "
                  f"{row['twoshot_roofline_frac']:.2f} / {row['twoshot_roofline_frac_nominal']:.2f} of link roofline @770/@900)  nvls {nv} us  "
                  f"fused+adam[{row['fused_algo']}] {fu} us", flush=True)
        del buf, out, nccl_bufs
        torch.cuda.empty_cache()
    if rank == 0 and args.out:
        json.dump(rows, open(args.out, "w"), indent=1)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
