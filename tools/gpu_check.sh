#!/usr/bin/env bash
# One GPU-box pass: kernel tests in isolated processes, smoke, both bench arms.
# Usage (via gpurun): bash tools/gpu_check.sh [steps] [warmup]
set -u
mkdir -p gpurun_out
STEPS=${1:-1000}; WARM=${2:-100}
nvidia-smi -L > gpurun_out/gpus.txt 2>&1
run() { # name, timeout, cmd...
  local name=$1 t=$2; shift 2
  echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; local rc=$?
  echo "rc=$rc"; tail -n 25 "gpurun_out/$name.log"; return 0
}
run t_optim 300 python -m pytest tests/test_gpu_kernels.py -q -m gpu -k "optimizer or cross_entropy_kernel" --timeout 120 -p no:cacheprovider
run t_gemm_basic 300 python -m pytest tests/test_gpu_kernels.py -q -m gpu -k "gemm_layouts" --timeout 120 -p no:cacheprovider
run t_gemm_epi 300 python -m pytest tests/test_gpu_kernels.py -q -m gpu -s -k "gemm and not layouts" --timeout 120 -p no:cacheprovider
run t_program 300 python -m pytest tests/test_gpu_kernels.py -q -m gpu -k "program or checkpoint" --timeout 120 -p no:cacheprovider
run smoke 200 python -c "import __graft_entry__ as g; g.smoke()"
run bench_ours 300 python bench.py --steps "$STEPS" --warmup "$WARM"
run bench_ref 600 python bench.py --impl reference --steps "$STEPS" --warmup "$WARM"
