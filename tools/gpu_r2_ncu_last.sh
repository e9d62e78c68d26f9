set -u
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
cap() { local tgt=$1 re=$2 name=$3 skip=${4:-1}
  timeout 600 ncu --set full --clock-control none --import-source on -k "regex:$re" -s $skip -c 1 -f -o gpurun_out/ncu_$name python tools/prof_targets.py $tgt > gpurun_out/ncu_$name.log 2>&1
  echo "$name rc=$? $(grep -c '==PROF== Profiling' gpurun_out/ncu_$name.log)"; }
timeout 200 python tools/prof_targets.py mlpstep > /dev/null 2>&1 || exit 1
timeout 200 python tools/prof_targets.py gpt2step > /dev/null 2>&1 || exit 1
cap mlpstep mlp_step_kernel mlp_step 1
cap gpt2step layernorm_bwd_kernel layernorm_bwd 25
cap gpt2step reduce_partials_kernel reduce_partials 25
