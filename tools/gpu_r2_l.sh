#!/bin/bash
# round 2, call L: GroupNorm backward with the dy' stash: parity + family cost
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ops_gpu.py tests/test_unet_gpu.py -m gpu -q -x > gpurun_out/l_pytest.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/l_pytest.log
python tools/exp_ablate_quick.py 2>&1 | tail -6
