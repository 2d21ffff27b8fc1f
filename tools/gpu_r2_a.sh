#!/bin/bash
# round 2, call A: new tests alone + whole GPU suite + bench
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q -s > gpurun_out/a_zz_alone.log 2>&1; echo "zz alone rc=$?"
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/a_pytest.log 2>&1; echo "suite rc=$?"
tail -5 gpurun_out/a_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err; echo "bench rc=$?"
cut -c1-600 gpurun_out/a_bench.json
