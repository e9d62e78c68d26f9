#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; tail -n "${TAILN:-15}" "gpurun_out/$name.log"; }
TAILN=25 run t_all_gpu 1200 python -m pytest tests/ -x -q -m gpu --timeout 600 -p no:cacheprovider
TAILN=3 run bench_ours_1 600 python bench.py --steps 2000 --warmup 200
B200DDP_NO_PDL=1 TAILN=3 run bench_ours_1_nopdl 600 python bench.py --steps 2000 --warmup 200 --no-e2e
TAILN=3 run bench_ours_n$N 600 $TR --master-port 29535 bench.py --gpus $N --steps 2000 --warmup 200
TAILN=14 run sweep_bf16_n$N 900 $TR --master-port 29541 bench/allreduce_sweep.py --dtype bf16 --out gpurun_out/sweep_bf16_n$N.json
TAILN=14 run sweep_fp32_n$N 900 $TR --master-port 29542 bench/allreduce_sweep.py --dtype fp32 --max-bytes 268435456 --out gpurun_out/sweep_fp32_n$N.json
