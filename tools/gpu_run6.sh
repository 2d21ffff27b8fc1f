#!/bin/bash
# 2-GPU box: multi-GPU tests + data-parallel bench at N = 2 (and N = 1 on the same box)
set -x
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -q -s > gpurun_out/run6_pytest.log 2>&1
tail -5 gpurun_out/run6_pytest.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/run6_dp2.json 2> gpurun_out/run6_dp2.err
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/run6_smoke.log 2>&1
