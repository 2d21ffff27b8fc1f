#!/bin/bash
# experiment batch 1: baseline bench, atomics-off / LN grid-cap variants (per-family times), ncu --set full of norm kernels
set -x
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 --no-ddim --no-cpu > gpurun_out/exp1_base.json 2> gpurun_out/exp1_base.err
ENCDIFF_EXP_NOATOM=1 python bench.py --steps 10 --warmup 3 --no-ddim --no-cpu > gpurun_out/exp1_noatom.json 2> gpurun_out/exp1_noatom.err
ENCDIFF_EXP_LN_CAP=148 python bench.py --steps 10 --warmup 3 --no-ddim --no-cpu > gpurun_out/exp1_cap148.json 2> gpurun_out/exp1_cap148.err
ENCDIFF_EXP_LN_CAP=2048 python bench.py --steps 10 --warmup 3 --no-ddim --no-cpu > gpurun_out/exp1_cap2048.json 2> gpurun_out/exp1_cap2048.err
for k in gn_bwd ln_bwd gn_fwd geglu_bwd; do
  timeout 300 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:$k -c 6 \
    -o gpurun_out/exp1_$k -f python tools/one_step.py train 256 > gpurun_out/exp1_ncu_$k.log 2>&1
done
ls -la gpurun_out
