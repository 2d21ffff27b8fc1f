#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/run4_pytest.log 2>&1
tail -5 gpurun_out/run4_pytest.log
python bench.py --steps 20 --warmup 5 > gpurun_out/run4_bench.json 2> gpurun_out/run4_bench.err
python bench.py --workload celeba64 --steps 10 --warmup 3 --no-ddim --no-cpu > gpurun_out/run4_celeba64.json 2> gpurun_out/run4_celeba64.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/run4_reference.json 2> gpurun_out/run4_reference.err
