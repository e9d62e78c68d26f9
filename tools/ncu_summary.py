#!/usr/bin/env python3
"""Summarises gpurun_out/ncu_<name>.ncu-rep (ncu --set full, one launch) into profiles/ncu/<name>.md:
duration, DRAM bytes and bandwidth against the measured copy bandwidth, tensor-pipe activity,
occupancy limits and the top stall reasons.  Runs here (no GPU):  python tools/ncu_summary.py"""
import csv
import glob
import io
import json
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles", "ncu")
PEAKS = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0}
KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor_op_hmma.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__grid_size", "launch__block_size", "launch__cluster_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "launch__waves_per_multiprocessor", "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
        "smsp__inst_executed.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum"]


def to_bytes(v, unit):
    f = float(v)
    return f * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def to_us(v, unit):
    return float(v) * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6, "usecond": 1, "nsecond": 1e-3, "msecond": 1e3}.get(unit, 1)


def main():
    os.makedirs(OUT, exist_ok=True)
    index = []
    for rep in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "ncu_*.ncu-rep"))):
        name = os.path.basename(rep)[4:-8]
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        if len(rows) < 3:
            continue
        hdr, units, val = rows[0], rows[1], rows[2]
        d = {h: (v, u) for h, u, v in zip(hdr, units, val)}
        kernel = d.get("Kernel Name", ("?", ""))[0]
        us = to_us(*d["gpu__time_duration.sum"])
        rd, wr = to_bytes(*d["dram__bytes_read.sum"]), to_bytes(*d["dram__bytes_write.sum"])
        gbs = (rd + wr) / (us * 1e-6) / 1e9
        stalls = sorted(((float(v[0]), h) for h, v in d.items()
                         if h.startswith("smsp__average_warp_latency_issue_stalled_") or h.startswith("smsp__average_warps_issue_stalled_")
                         if v[0] not in ("", "n/a")), reverse=True)[:5]
        tensor = d.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", ("", ""))[0]
        lines = [f"# `{kernel}` — ncu `--set full --clock-control none`, one launch inside `tools/prof_targets.py` (round 2)", "",
                 f"Duration {us:.1f} us (cold caches: ncu flushes L2 and serialises launches; compare shares, not absolutes). "
                 f"DRAM {rd / 1e6:.2f} MB read + {wr / 1e6:.2f} MB written = {gbs:.0f} GB/s = "
                 f"{gbs / PEAKS['hbm_gbs']:.2f} of the measured copy bandwidth ({PEAKS['hbm_gbs']:.0f} GB/s)."
                 + (f"  Tensor pipe active {float(tensor):.1f} % of active cycles." if tensor not in ("", "n/a") and float(tensor) > 0 else ""), "",
                 "| metric | unit | value |", "|---|---|---|"]
        for k in KEEP:
            if k in d and d[k][0] not in ("", "n/a"):
                lines.append(f"| {k} | {d[k][1]} | {d[k][0]} |")
        if stalls:
            lines += ["", "Top issue-stall reasons (warp-cycles per issued instruction):", ""]
            lines += [f"- {h.split('stalled_')[-1]}: {v:.2f}" for v, h in stalls]
        open(os.path.join(OUT, name + ".md"), "w").write("\n".join(lines) + "\n")
        index.append((name, kernel, us, gbs, tensor))
    with open(os.path.join(OUT, "INDEX.md"), "w") as f:
        f.write("# ncu captures, one launch per kernel family (round 2)\n\nWritten by `tools/ncu_summary.py` from the reports "
                "`tools/gpu_r2_ncu_families.sh` brings back.  Durations are cold-cache, serialised launches.\n\n"
                "Not here: the multi-GPU collectives (`allreduce_optimizer_kernel`, `allreduce_kernel`, `broadcast_kernel`, the "
                "update CTAs of `mlp_step_kernel` at W > 1).  ncu on this pool is for one GPU only (a multi-rank command under it "
                "replays every kernel ~40 times against peers that have moved on), so NVLink byte counters cannot be captured; "
                "their evidence is the CUDA-event sweeps with per-algorithm rooflines (`../r2_allreduce_sweep_bf16_4gpu.txt`, "
                "`../r1_call9_allreduce_sweep_bf16_8gpu_nvls.txt`), the SASS listings (`../sass/allreduce_optimizer_kernel.sass`: "
                "`multimem.ld_reduce` / `multimem.st`, fence counts) and the in-kernel wall-clock trace of the 8-GPU step "
                "(`../mlp_fused_timeline.md`).  `mlp_step` below is the single-GPU launch of the same kernel (24 steps).\n\n"
                "| capture | kernel | us | DRAM GB/s | of copy bandwidth | tensor pipe % |\n|---|---|---|---|---|---|\n")
        for name, kernel, us, gbs, tensor in index:
            t = f"{float(tensor):.1f}" if tensor not in ("", "n/a") and float(tensor) > 0 else ""
            f.write(f"| [{name}]({name}.md) | `{kernel[:60]}` | {us:.1f} | {gbs:.0f} | {gbs / PEAKS['hbm_gbs']:.2f} | {t} |\n")
    print(len(index), "summaries")


if __name__ == "__main__":
    main()
