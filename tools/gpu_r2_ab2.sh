#!/usr/bin/env bash
# A/B of the update-CTA variants at N GPUs: bench lines only.  usage: bash tools/gpu_r2_ab2.sh N
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
one() { local name=$1; shift; env "$@" timeout 200 $TR --master-port $((29800 + RANDOM % 100)) bench.py --gpus $N --steps 2000 --warmup 200 > gpurun_out/ab_$name.log 2>&1; python - gpurun_out/ab_$name.log $name <<'PY'
import json, sys
for line in open(sys.argv[1]):
    if line.startswith('{"metric"'):
        d = json.loads(line); print(sys.argv[2], "device us/step", round(d["ms_per_step"] * 1e3, 2), "e2e us/step", round(d["e2e"]["ms_per_step"] * 1e3, 2), "launches", d["gpu_launches"])
PY
}
one regs B200DDP_MLP_COMM_REGS=1
one mem B200DDP_MLP_COMM_REGS=0
one regs2 B200DDP_MLP_COMM_REGS=1
one separate B200DDP_MLP_UPDATE_CLUSTERS=0
one single_step B200DDP_MLP_STEPS_PER_LAUNCH=1
