#!/bin/bash
# round 2, call B: cluster GroupNorm: parity + timing sweep
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ops_gpu.py tests/test_unet_gpu.py -m gpu -q -x > gpurun_out/b_pytest.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/b_pytest.log
for v in off 4 2 1; do
  if [ $v = off ]; then export ENCDIFF_GN_NO_CLUSTER=1; else unset ENCDIFF_GN_NO_CLUSTER; export ENCDIFF_GN_VPT=$v; fi
  timeout 300 python bench.py --steps 20 --warmup 5 --no-ddim --no-cpu --no-comparators > gpurun_out/b_bench_$v.json 2> gpurun_out/b_bench_$v.err; echo "bench $v rc=$?"
  python - <<PY
import json
d=json.loads(open('gpurun_out/b_bench_$v.json').read().strip().splitlines()[-1])
f=d['roofline']['families']
print('$v', 'ms/step', round(d['ms_per_step'],3), {k: f[k]['ms'] for k in ('gn1','gn2','gn1.bwd','gn2.bwd','st.norm','st.norm.bwd')})
PY
done
