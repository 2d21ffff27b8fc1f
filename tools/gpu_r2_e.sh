#!/bin/bash
# round 2, call E: marginal in-graph cost per kernel family (ablation) + GEMM shape timings of the current library
mkdir -p gpurun_out
timeout 900 python tools/exp_ablate.py 256 20 > gpurun_out/e_ablate.log 2>&1; echo "ablate rc=$?"; cat gpurun_out/e_ablate.log | tail -30
timeout 300 python tools/exp_gemm_pacing.py > gpurun_out/e_pacing.log 2>&1; echo "pacing rc=$?"; cat gpurun_out/e_pacing.log | tail -10
timeout 600 python -m pytest tests/test_train_gpu.py -m gpu -q -x -k "loss_curve_shapes3d" -s 2>&1 | grep -E "shapes3d|passed|failed" | tail -3
