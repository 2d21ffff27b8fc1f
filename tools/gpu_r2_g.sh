#!/bin/bash
# round 2, call G: register-resident cross-attention kernels: parity + step time
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ops_gpu.py tests/test_unet_gpu.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/g_pytest.log 2>&1; echo "tests rc=$?"; tail -15 gpurun_out/g_pytest.log
timeout 300 python bench.py --steps 20 --warmup 5 --no-ddim --no-cpu --no-comparators > gpurun_out/g_bench.json 2> gpurun_out/g_bench.err; echo "bench rc=$?"
python tools/bench_summary.py gpurun_out/g_bench.json | head -12
