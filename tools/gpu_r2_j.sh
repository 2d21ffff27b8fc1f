#!/bin/bash
# round 2, call J: single-pass self-attention backward: parity + step time (A/B against the two-pass kernel)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ops_gpu.py tests/test_unet_gpu.py -m gpu -q -x > gpurun_out/j_pytest.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/j_pytest.log
for v in new staged; do
  if [ $v = staged ]; then export ENCDIFF_ATTN_STAGED=1; else unset ENCDIFF_ATTN_STAGED; fi
  timeout 300 python bench.py --steps 20 --warmup 5 --no-ddim --no-cpu --no-comparators > gpurun_out/j_bench_$v.json 2> gpurun_out/j_bench_$v.err; echo "bench $v rc=$?"
  python tools/bench_summary.py gpurun_out/j_bench_$v.json | head -3
done
