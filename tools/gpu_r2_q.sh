#!/bin/bash
# round 2, call Q: whole GPU suite + full bench (all legs)
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/q_pytest.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/q_pytest.log
timeout 1200 python bench.py --steps 20 --warmup 5 > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/q_bench.json').read().strip().splitlines()[-1])
print("ms/step", d['ms_per_step'], "value", d['value'], "e2e", d['e2e']['value'])
r=d['roofline']; print("roofline", {k:r[k] for k in ('achieved','frac','family_ms_per_step','share_of_step') if k in r}, r.get('eager_events',{}).get('frac'))
print("ddim", d.get('ddim50')); print("swap", d.get('swap_ddim200'))
print("cpu", d.get('cpu_baseline')); print("cfg1", d.get('config1_cpu_full_step'))
print("comparators", json.dumps(d.get('comparators'))[:1500])
print("pipeline", d.get('full_pipeline'), d.get('full_pipeline_error'))
PY
