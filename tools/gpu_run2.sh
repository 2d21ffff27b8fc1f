#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_train_gpu.py::test_loss_curve_1k_vs_reference > gpurun_out/run2_pytest.log 2>&1
tail -5 gpurun_out/run2_pytest.log
python bench.py --steps 10 --warmup 3 --no-ddim --no-cpu > gpurun_out/run2_bench.json 2> gpurun_out/run2_bench.err
ENCDIFF_EXP_LN_CAP=1184 python bench.py --steps 10 --warmup 3 --no-ddim --no-cpu > gpurun_out/run2_cap1184.json 2> gpurun_out/run2_cap1184.err
ENCDIFF_EXP_LN_CAP=296 python bench.py --steps 10 --warmup 3 --no-ddim --no-cpu > gpurun_out/run2_cap296.json 2> gpurun_out/run2_cap296.err
