#!/bin/bash
# round 2, call M: graph replay in the autograd (drop-in) path: parity + comparator
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_unet_gpu.py tests/test_dropin_gpu.py tests/test_train_gpu.py -m gpu -q -x > gpurun_out/m_pytest.log 2>&1; echo "tests rc=$?"; tail -6 gpurun_out/m_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-ddim --no-cpu > gpurun_out/m_bench.json 2> gpurun_out/m_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/m_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/m_bench.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], json.dumps(d.get('comparators'))[:900])
print({k:v for k,v in d['roofline'].items() if k in ('achieved','frac','family_ms_per_step','eager_events')})
PY
