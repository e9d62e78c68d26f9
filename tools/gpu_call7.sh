#!/usr/bin/env bash
# 8-GPU validation: multi-rank numerics, flagship bench both arms, one-shot vs two-shot, sweep, GPT-2.
set -u
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; grep -v "OMP_NUM_THREADS\|^\*\*\*\*" "gpurun_out/$name.log" | tail -n "${TAILN:-15}"; }
TAILN=20 run multi_n$N 600 $TR --master-port 29533 tests/multi_gpu_worker.py
TAILN=2 run bench_ours_n$N 300 $TR --master-port 29535 bench.py --gpus $N --steps 2000 --warmup 200
B200DDP_ONE_SHOT_MAX_BYTES=67108864 TAILN=2 run bench_ours_n${N}_oneshot 300 $TR --master-port 29536 bench.py --gpus $N --steps 2000 --warmup 200 --no-e2e
TAILN=2 run bench_ref_n$N 600 $TR --master-port 29534 bench.py --impl reference --gpus $N --steps 500 --warmup 50
TAILN=14 run sweep_bf16_n$N 600 $TR --master-port 29541 bench/allreduce_sweep.py --dtype bf16 --out gpurun_out/sweep_bf16_n$N.json
TAILN=14 run sweep_fp32_n$N 600 $TR --master-port 29542 bench/allreduce_sweep.py --dtype fp32 --max-bytes 268435456 --out gpurun_out/sweep_fp32_n$N.json
TAILN=2 run gpt2_ours_n$N 600 $TR --master-port 29543 bench/bench_models.py --model gpt2 --gpus $N --batch-size 8 --steps 10 --warmup 3
TAILN=2 run gpt2_ref_n$N 600 $TR --master-port 29544 bench/bench_models.py --model gpt2 --gpus $N --batch-size 8 --steps 10 --warmup 3 --impl reference
