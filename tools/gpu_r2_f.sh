#!/bin/bash
# round 2, call F: ncu --set full (with source counters) of the attention backward / GroupNorm backward kernels of one eager step
mkdir -p gpurun_out
python tools/one_step.py > gpurun_out/f_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/f_plain.log; exit 1; }
for k in cross_attn_bwd_mma_kernel self_attn_bwd_mma_kernel gn_bwd_kernel cross_attn_fwd_mma_kernel self_attn_fwd_mma_kernel; do
  ncu --profile-from-start off --set full --import-source on --clock-control none -k regex:$k -c 1 -f -o gpurun_out/f_$k python tools/one_step.py > gpurun_out/f_ncu_$k.log 2>&1; echo "$k rc=$?"
done
ls -la gpurun_out/*.ncu-rep
