#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; tail -n "${TAILN:-15}" "gpurun_out/$name.log"; }
run t_noalloc_alone 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu --timeout 120 -p no:cacheprovider
run t_all_gpu 900 python -m pytest tests/ -x -q -m gpu --timeout 300 -p no:cacheprovider
TAILN=40 run diag 300 python tools/diag_program.py
run gemm_plain 200 python tools/prof_gemm.py 4096 5
run bench_short 200 python bench.py --steps 20 --warmup 5 --no-e2e
echo "=== ncu launch list"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 20 --warmup 5 --no-e2e > gpurun_out/ncu_launches.log 2>&1; echo "rc=$?"
echo "=== ncu gemm full"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 5 -c 2 -o gpurun_out/prof_gemm python tools/prof_gemm.py 4096 5 > gpurun_out/ncu_gemm.log 2>&1; echo "rc=$?"
tail -5 gpurun_out/ncu_gemm.log
