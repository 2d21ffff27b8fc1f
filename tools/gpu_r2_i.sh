#!/bin/bash
# round 2, call I: whole GPU suite + bench after the attention / fp32-validation work
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/i_pytest.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/i_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/i_bench.json 2> gpurun_out/i_bench.err; echo "bench rc=$?"
python tools/bench_summary.py gpurun_out/i_bench.json | head -12
timeout 600 python tools/exp_ablate.py 256 20 > gpurun_out/i_ablate.log 2>&1; cat gpurun_out/i_ablate.log
