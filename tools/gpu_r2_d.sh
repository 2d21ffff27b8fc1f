#!/bin/bash
# round 2, call D: baseline of the committed build: whole GPU suite, bench (both arms), ncu launch list of one step
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/d_pytest.log 2>&1; echo "suite rc=$?"; tail -5 gpurun_out/d_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/d_bench.json 2> gpurun_out/d_bench.err; echo "bench rc=$?"
cut -c1-400 gpurun_out/d_bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/d_bench_ref.json 2> gpurun_out/d_bench_ref.err; echo "ref rc=$?"
cut -c1-400 gpurun_out/d_bench_ref.json
python tools/one_step.py > gpurun_out/d_one_step.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 3000 --csv --log-file gpurun_out/d_launches.csv python tools/one_step.py > gpurun_out/d_ncu.log 2>&1; echo "ncu rc=$?"
