#!/bin/bash
# round 2, call C: warp-uniform producer / MMA issue loops: old vs new library on one box
mkdir -p gpurun_out
echo "== old"; ENCDIFF_B200_NO_BUILD=1 ENCDIFF_B200_LIB=encdiff_b200/_build/lib_old.so python tools/exp_gemm_pacing.py 2>&1 | tee gpurun_out/c_pacing_old.log
echo "== new"; python tools/exp_gemm_pacing.py 2>&1 | tee gpurun_out/c_pacing_new.log
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_unet_gpu.py -m gpu -q -x > gpurun_out/c_pytest.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/c_pytest.log
for v in old new; do
  if [ $v = old ]; then export ENCDIFF_B200_NO_BUILD=1 ENCDIFF_B200_LIB=encdiff_b200/_build/lib_old.so; else unset ENCDIFF_B200_LIB; fi
  timeout 300 python bench.py --steps 20 --warmup 5 --no-ddim --no-cpu --no-comparators > gpurun_out/c_bench_$v.json 2> gpurun_out/c_bench_$v.err; echo "bench $v rc=$?"
  python tools/bench_summary.py gpurun_out/c_bench_$v.json 2>/dev/null | head -12
done
