#!/usr/bin/env bash
# One ncu --set full capture (1 launch) per kernel family; each target first runs plainly.
# usage: bash tools/gpu_r2_ncu_families.sh            (reports land in gpurun_out/ncu_<name>.ncu-rep)
set -u
mkdir -p gpurun_out
cap() { # target, kernel regex, name, skip
  local tgt=$1 re=$2 name=$3 skip=${4:-1}
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$re" -s $skip -c 1 -f -o gpurun_out/ncu_$name \
      python tools/prof_targets.py $tgt > gpurun_out/ncu_$name.log 2>&1
  echo "$name rc=$? $(grep -c '==PROF== Profiling' gpurun_out/ncu_$name.log)"
}
for tgt in mlpstep gpt2step rn50step; do
  timeout 300 python tools/prof_targets.py $tgt > gpurun_out/plain_$tgt.log 2>&1 || { echo "plain $tgt failed"; tail -5 gpurun_out/plain_$tgt.log; exit 1; }
done
if [ "${1:-all}" = "changed" ]; then   # the kernels touched since the last full pass
cap mlpstep mlp_step_kernel mlp_step 1
cap gpt2step flash_attn_fwd2_kernel flash_fwd 12
cap gpt2step flash_attn_bwd_kernel flash_bwd 12
cap gpt2step optimizer_kernel optimizer 0
cap rn50step bn_reduce_kernel bn_reduce 60
cap rn50step bn_apply_kernel bn_apply 60
cap rn50step bn_bwd_apply_kernel bn_bwd_apply 60
cap rn50step bn_finalize_fwd_kernel bn_finalize 30
cap rn50step im2col_flat_kernel im2col_flat 0
cap rn50step gemm_kernel_2sm conv_gemm_2sm 30
cap rn50step "gemm_kernel<256" conv_wgrad_splitk 20
else
cap mlpstep mlp_step_kernel mlp_step 1
cap gpt2step optimizer_kernel optimizer 0
cap gpt2step flash_attn_fwd2_kernel flash_fwd 12
cap gpt2step flash_attn_bwd_kernel flash_bwd 12
cap gpt2step gemm_kernel_2sm gemm_2sm 60
cap gpt2step gemm_kernel_m256 gemm_m256 1
cap gpt2step layernorm_fwd_kernel layernorm_fwd 25
cap gpt2step layernorm_bwd_kernel layernorm_bwd 25
cap gpt2step ce_row_vec_kernel ce_row_vec 1
cap gpt2step embedding_scatter_kernel embedding_scatter 1
cap gpt2step colsum_stage1_kernel colsum 20
cap rn50step bn_reduce_kernel bn_reduce 60
cap rn50step bn_apply_kernel bn_apply 60
cap rn50step bn_bwd_apply_kernel bn_bwd_apply 60
cap rn50step im2col_flat_kernel im2col_flat 0
cap rn50step col2im_kernel col2im 10
cap rn50step maxpool_fwd_kernel maxpool_fwd 1
fi
ls -la gpurun_out/*.ncu-rep | awk '{s+=$5} END {print s/1e6 " MB of reports"}'
