#!/bin/bash
mkdir -p gpurun_out
python tools/one_step.py > gpurun_out/k_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/k_plain.log; exit 1; }
for k in self_attn_bwd_sp_kernel cross_attn_bwd_reg_kernel; do
  ncu --profile-from-start off --set full --import-source on --clock-control none -k regex:$k -c 1 -f -o gpurun_out/k_$k python tools/one_step.py > gpurun_out/k_ncu_$k.log 2>&1; echo "$k rc=$?"
done
