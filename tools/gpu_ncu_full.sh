#!/bin/bash
# `ncu --set full` captures of the final round-2 build (run only after tools/gpu_final.sh has passed on the same build):
# a 12-launch sample of the dominant GEMM family inside one training step, and every kernel of the concept-encoder trunk,
# the first-stage specific kernels, the weight pack and the optimiser inside one full step from images.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
O=gpurun_out
python tools/one_step.py train > $O/ncu_pre_train.log 2>&1 || exit 1
timeout 1200 ncu --profile-from-start off --set full --import-source on --clock-control none -k regex:sgemm2_kernel \
  --launch-skip 40 --launch-count 12 -f -o $O/r02_sgemm2_full python tools/one_step.py train > $O/ncu_sgemm2.log 2>&1; echo "sgemm2 rc=$?"
ncu -i $O/r02_sgemm2_full.ncu-rep --page raw --csv > $O/r02_sgemm2_full_raw.csv 2>/dev/null
python tools/one_step.py full > $O/ncu_pre_full.log 2>&1 || exit 1
timeout 1500 ncu --profile-from-start off --set full --clock-control none -k "regex:e4_|vq_|pack_kernel|adamw_dev|concept_mlp|wgrad2_group" \
  -c 80 -f -o $O/r02_full_step_kernels python tools/one_step.py full > $O/ncu_full_kernels.log 2>&1; echo "full-step kernels rc=$?"
ncu -i $O/r02_full_step_kernels.ncu-rep --page raw --csv > $O/r02_full_step_kernels_raw.csv 2>/dev/null
rm -f $O/r02_full_step_kernels.ncu-rep
ls -la $O/*.ncu-rep $O/*_raw.csv
