#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_unet_gpu.py tests/test_train_gpu.py -m gpu -q -x > gpurun_out/run24_pytest.log 2>&1
tail -3 gpurun_out/run24_pytest.log
run() { env "$@" python bench.py --steps 10 --warmup 3 --no-ddim --no-cpu 2>> gpurun_out/run24.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); f=d['roofline']['families']; print('$*', round(d['ms_per_step'],3), f['sgemm']['ms'], f['wgrad']['ms'], d['loss_first_last'])" >> gpurun_out/run24.log; }
OLD=$GRAFT_REPO_ROOT/gpurun_old/libencdiff_b200_fb44ffd.so
run ENCDIFF_B200_NO_BUILD=1 ENCDIFF_B200_LIB=$OLD
run X=1
run ENCDIFF_B200_NO_BUILD=1 ENCDIFF_B200_LIB=$OLD
run X=1
