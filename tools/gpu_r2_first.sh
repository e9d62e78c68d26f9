#!/usr/bin/env bash
# Round-2 first GPU pass (1 GPU): the fused cluster-kernel step, then the whole GPU tier, then both bench arms.
set -u
mkdir -p gpurun_out
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; tail -n "${TAILN:-12}" "gpurun_out/$name.log"; }
nvidia-smi -L > gpurun_out/gpus.txt 2>&1
TAILN=30 run a_fused 300 python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k "fused_mlp or mlp_program" --timeout 120 -p no:cacheprovider
TAILN=6 run a_smoke 200 python -c "import __graft_entry__ as g; g.smoke()"
TAILN=3 run a_bench20 200 python bench.py --steps 20 --warmup 5
TAILN=3 run a_bench2000 200 python bench.py --steps 2000 --warmup 200
TAILN=15 run a_tests 900 python -m pytest tests -q -m gpu -x --timeout 300 -p no:cacheprovider
TAILN=3 run a_ref20 300 python bench.py --impl reference --steps 20 --warmup 5
