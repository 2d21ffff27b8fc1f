#!/bin/bash
# Round-2 final verification of the committed state on one B200: GPU tests, smoke, both bench arms, ncu launch lists of
# one UNet training step and of one full step from images, in-graph ablation, first-stage / concept-encoder benches.
# usage (from the repo root on the GPU box):  bash tools/gpu_final.sh     -> everything lands in gpurun_out/final_*
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > $O/final_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/final_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/final_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/final_smoke.log
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $O/final_reference.json 2> $O/final_reference.err; echo "reference arm rc=$?"
timeout 1500 python bench.py --steps 20 --warmup 5 > $O/final_bench.json 2> $O/final_bench.err; echo "bench rc=$?"
python tools/bench_summary.py $O/final_bench.json 2>/dev/null | head -40
# launch lists: only after the same programs exited 0 without the profiler (above: bench; one_step is the same step, eager)
python tools/one_step.py train > $O/final_one_step.log 2>&1 && \
  timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum \
    --clock-control none -c 3000 --csv --log-file $O/final_launches.csv python tools/one_step.py train > $O/final_ncu.log 2>&1
echo "ncu train rc=$?"
python tools/one_step.py full > $O/final_one_step_full.log 2>&1 && \
  timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum \
    --clock-control none -c 4000 --csv --log-file $O/final_launches_full.csv python tools/one_step.py full > $O/final_ncu_full.log 2>&1
echo "ncu full rc=$?"
python tools/launch_summary.py $O/final_launches.csv 60 > $O/final_launches.txt 2>&1
python tools/launch_summary.py $O/final_launches_full.csv 60 > $O/final_launches_full.txt 2>&1
python tools/ncu_dram_summary.py $O/final_launches.csv $O/final_sgemm2_dram.json > $O/final_dram_summary.txt 2>&1
timeout 900 python tools/exp_ablate.py > $O/final_ablation.txt 2>&1; echo "ablation rc=$?"
timeout 600 python tools/vq_bench.py 2>/dev/null | tail -1 > $O/final_vq_bench.json; echo "vq bench rc=$?"
timeout 300 python tools/enc4_profile.py > $O/final_enc4_profile.txt 2>&1; echo "enc4 profile rc=$?"
timeout 300 python tools/exp_pipeline_breakdown.py 2>&1 | tail -6 > $O/final_pipeline_breakdown.txt; echo "breakdown rc=$?"
