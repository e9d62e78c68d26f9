#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; grep -v "OMP_NUM_THREADS\|^\*\*\*\*" "gpurun_out/$name.log" | tail -n "${TAILN:-15}"; }
TAILN=30 run t_all_gpu 1200 python -m pytest tests/ -x -q -m gpu --timeout 600 -p no:cacheprovider
TAILN=2 run gpt2_native 600 python bench/bench_models.py --model gpt2 --batch-size 8 --steps 10 --warmup 3
B200DDP_ATTENTION=sdpa TAILN=2 run gpt2_sdpa 600 python bench/bench_models.py --model gpt2 --batch-size 8 --steps 10 --warmup 3
