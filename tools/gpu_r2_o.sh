#!/bin/bash
# round 2, call O: first-stage VQ encoder on the GPU
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_vq_gpu.py -m gpu -q -x -s > gpurun_out/o_pytest.log 2>&1; echo "tests rc=$?"; grep -E "rel err|passed|failed|Error|error|assert" gpurun_out/o_pytest.log | head -20
