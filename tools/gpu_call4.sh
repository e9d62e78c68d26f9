#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; tail -n "${TAILN:-15}" "gpurun_out/$name.log"; }
TAILN=40 run t_transformer 600 python -m pytest tests/test_gpu_kernels.py -q -m gpu -s -k "layernorm or colsum or gpt2" --timeout 300 -p no:cacheprovider
TAILN=5 run gpt2_ours 900 python bench/bench_models.py --model gpt2 --batch-size 8 --steps 10 --warmup 3
TAILN=5 run gpt2_ref 900 python bench/bench_models.py --model gpt2 --batch-size 8 --steps 10 --warmup 3 --impl reference
