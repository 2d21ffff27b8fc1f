#!/bin/bash
# final launch list of one B = 256 training step (ncu, time + DRAM bytes per launch) and a --set full capture of the top GEMM variants
set -x
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
python tools/one_step.py train 256 > gpurun_out/run8_plain.log 2>&1 || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
  --profile-from-start off -c 1200 --csv --log-file gpurun_out/run8_launches.csv python tools/one_step.py train 256 > gpurun_out/run8_ncu.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:sgemm2_kernel --launch-skip 40 -c 8 \
  -o gpurun_out/run8_sgemm2 -f python tools/one_step.py train 256 > gpurun_out/run8_ncu_sgemm2.log 2>&1
