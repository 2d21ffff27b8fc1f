#!/bin/bash
# five runs of the Shapes3D loss-curve test on one build: run-to-run spread of the chaotic batch-8 trajectory
mkdir -p gpurun_out
for v in 1 2 3 4 5; do
  timeout 300 python -m pytest tests/test_train_gpu.py -m gpu -q -k loss_curve_shapes3d -s 2>&1 | grep -E "shapes3d 400|passed|failed"
done | tee gpurun_out/n_loss_curve_runs.txt
