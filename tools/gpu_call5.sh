#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; tail -n "${TAILN:-15}" "gpurun_out/$name.log"; }
TAILN=30 run sweep_bf16_n$N 900 $TR --master-port 29541 bench/allreduce_sweep.py --dtype bf16 --out gpurun_out/sweep_bf16_n$N.json
TAILN=30 run sweep_fp32_n$N 900 $TR --master-port 29542 bench/allreduce_sweep.py --dtype fp32 --max-bytes 268435456 --out gpurun_out/sweep_fp32_n$N.json
TAILN=3 run gpt2_ours_n$N 900 $TR --master-port 29543 bench/bench_models.py --model gpt2 --gpus $N --batch-size 8 --steps 10 --warmup 3
TAILN=3 run gpt2_ref_n$N 900 $TR --master-port 29544 bench/bench_models.py --model gpt2 --gpus $N --batch-size 8 --steps 10 --warmup 3 --impl reference
