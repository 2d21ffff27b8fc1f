#!/bin/bash
# 8-GPU box: data-parallel scaling of the final build, N = 8, 4, 2, 1 back to back on the same box
set -x
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
for n in 8 4 2; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 --no-cpu > gpurun_out/run7_dp$n.json 2> gpurun_out/run7_dp$n.err
done
python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu > gpurun_out/run7_dp1.json 2> gpurun_out/run7_dp1.err
