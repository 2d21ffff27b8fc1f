#!/bin/bash
# round 2, call H: fp32 validation forward
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fp32_validation_gpu.py -m gpu -q -x -s > gpurun_out/h_pytest.log 2>&1; echo "tests rc=$?"; grep -E "rel err|passed|failed|Error|error" gpurun_out/h_pytest.log | head -20
