#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/run3_pytest.log 2>&1
tail -15 gpurun_out/run3_pytest.log
for v in 256 128 0; do
  ENCDIFF_LN_FUSE_MAXC=$v python bench.py --steps 10 --warmup 3 --no-ddim --no-cpu > gpurun_out/run3_fuse$v.json 2> gpurun_out/run3_fuse$v.err
done
ENCDIFF_EXP_LN_CAP=148 python bench.py --steps 10 --warmup 3 --no-ddim --no-cpu > gpurun_out/run3_cap148.json 2> gpurun_out/run3_cap148.err
ENCDIFF_EXP_LN_CAP=222 python bench.py --steps 10 --warmup 3 --no-ddim --no-cpu > gpurun_out/run3_cap222.json 2> gpurun_out/run3_cap222.err
for k in gn_bwd gn_fwd ln_bwd; do
  timeout 300 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:$k -c 4 \
    -o gpurun_out/run3_$k -f python tools/one_step.py train 256 > gpurun_out/run3_ncu_$k.log 2>&1
done
