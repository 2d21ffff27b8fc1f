#!/bin/bash
# 2-GPU box: chunked all-reduce overlapped with AdamW vs the single all-reduce (ENCDIFF_AR_CHUNKS), + the multi-GPU tests
set -x
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py tests/test_train_gpu.py -m gpu -q -s > gpurun_out/run9_pytest.log 2>&1
tail -3 gpurun_out/run9_pytest.log
for k in 4 1 8 2; do
  ENCDIFF_AR_CHUNKS=$k python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2952$k bench.py --gpus 2 --steps 30 --warmup 5 --no-ddim > gpurun_out/run9_dp2_k$k.json 2> gpurun_out/run9_dp2_k$k.err
done
