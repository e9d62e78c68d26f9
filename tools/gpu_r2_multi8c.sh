#!/usr/bin/env bash
# Final N-GPU confirmation: numerics worker, bench at 2000 and 20 steps (device + end to end).
set -u
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; grep -v "OMP_NUM_THREADS\|^\*\*\*\*\|^$" "gpurun_out/$name.log" | tail -n "${TAILN:-12}" | cut -c1-${CUT:-400}; }
TAILN=9 run m${N}c_worker 400 $TR --master-port 29701 tests/multi_gpu_worker.py
CUT=2200 TAILN=1 run m${N}c_bench2000 200 $TR --master-port 29703 bench.py --gpus $N --steps 2000 --warmup 200
CUT=300 TAILN=1 run m${N}c_bench20 200 $TR --master-port 29702 bench.py --gpus $N --steps 20 --warmup 5
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/m*c_bench*.log")):
    for line in open(f):
        if line.startswith('{"metric"'):
            d = json.loads(line); print(f, "device us/step", round(d["ms_per_step"] * 1e3, 2), "e2e us/step", round(d["e2e"]["ms_per_step"] * 1e3, 2), "value", round(d["value"]), "e2e value", round(d["e2e"]["value"]))
PY
