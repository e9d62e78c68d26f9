#!/usr/bin/env python3
"""Writes one SASS listing per kernel family of libb200ddp.so into profiles/sass/ and an index
with the mnemonics that identify the Blackwell datapath (B200_PROFILING.md "What proves a
Blackwell-native kernel").  Runs without a GPU:  python tools/dump_sass.py"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "b200ddp", "_native", "libb200ddp.so")
OUT = os.path.join(ROOT, "profiles", "sass")
# family -> preferred instantiation (substring of the demangled name); first match wins
PREFER = {
    "gemm_kernel": "gemm_kernel<256, true>", "gemm_kernel_2sm": "gemm_kernel_2sm<256>",
    "allreduce_optimizer_kernel": "allreduce_optimizer_kernel<false, 2, true>",
    "allreduce_kernel": "allreduce_kernel<true, 2>", "ce_kernel": "ce_kernel<true, 256>",
    "mlp_step_kernel": "mlp_step_kernel<true>", "optimizer_kernel": "optimizer_kernel<false>",
    "layernorm_fwd_kernel": "layernorm_fwd_kernel<3>", "layernorm_bwd_kernel": "layernorm_bwd_kernel<3>",
}
MARKS = ["UTCHMMA", "UTCQMMA", "UTMALDG", "UTMASTG", "UBLKCP", "LDTM", "STTM", "UTCBAR", "UCGABAR_ARV", "HMMA",
         "LDGMC", "STGMC", "MEMBAR.ALL.SYS", "MEMBAR.ALL.GPU", "CCTL.IVALL", "LDG.E.128", "STG.E.128"]


def main():
    dump = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    blocks, name, buf = {}, None, []
    for line in dump.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            if name:
                blocks[name] = buf
            name, buf = m.group(1), []
        if name:
            buf.append(line)
    if name:
        blocks[name] = buf
    dem = subprocess.run(["c++filt"], input="\n".join(blocks), capture_output=True, text=True).stdout.splitlines()
    fams = collections.OrderedDict()
    for mangled, d in zip(blocks, dem):
        short = re.sub(r"\(.*", "", d).replace("void ", "").replace("b200::", "")
        fam = re.sub(r"<.*", "", short)
        if not fam or fam.startswith("_Z"):      # anonymous-namespace kernels do not demangle to a bare name
            m = re.search(r"\d+([A-Za-z_]+_kernel)", mangled)
            fam = m.group(1) if m else mangled[:40]
            short = short or fam
        fams.setdefault(fam, []).append((short, mangled))
    os.makedirs(OUT, exist_ok=True)
    rows = []
    for fam, insts in sorted(fams.items()):
        pick = next((i for i in insts if PREFER.get(fam) and PREFER[fam] in i[0]), insts[0])
        # keep the mnemonics and operands, drop the 128-bit encodings (two thirds of the bytes)
        body = [re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/", "", l).rstrip() for l in blocks[pick[1]]]
        body = [l for l in body if l.strip()]
        text = "\n".join(body)
        counts = {k: len(re.findall(r"\b" + re.escape(k), text)) for k in MARKS}
        n_inst = sum(1 for l in body if re.match(r"\s+/\*[0-9a-f]{4}\*/", l))
        path = os.path.join(OUT, fam + ".sass")
        with open(path, "w") as f:
            f.write(f"// {pick[0]}   ({len(insts)} instantiation(s) in the library: " + ", ".join(i[0] for i in insts) + ")\n")
            f.write(text + "\n")
        rows.append((fam, pick[0], n_inst, counts))
    with open(os.path.join(OUT, "INDEX.md"), "w") as f:
        f.write("# SASS listings, one per kernel family (`cuobjdump -sass`, sm_100a)\n\n"
                "Written by `tools/dump_sass.py`.  Mnemonic counts of the listed instantiation "
                "(UTCHMMA = tcgen05.mma, UTMALDG/UTMASTG = TMA load/store, LDTM = tcgen05.ld, "
                "LDGMC/STGMC = multimem.ld_reduce / multimem.st, HMMA = legacy mma.sync: none).\n\n")
        keys = [k for k in MARKS if any(r[3][k] for r in rows)]
        f.write("| listing | instantiation | instr | " + " | ".join(keys) + " |\n|---|---|---|" + "---|" * len(keys) + "\n")
        for fam, inst, n, c in rows:
            f.write(f"| [{fam}.sass]({fam}.sass) | `{inst}` | {n} | " + " | ".join(str(c[k]) if c[k] else "" for k in keys) + " |\n")
    print(f"{len(rows)} listings in {OUT}")


if __name__ == "__main__":
    sys.exit(main())
