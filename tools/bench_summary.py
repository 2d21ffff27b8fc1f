"""Print a compact summary of a bench.py JSON line (tools only)."""
import json
import sys

for path in sys.argv[1:]:
    lines = [x for x in open(path) if x.startswith("{")]
    if not lines:
        print(path, "no JSON line")
        print(open(path).read()[-1500:])
        continue
    d = json.loads(lines[-1])
    print(f"{path}: {d['value']:.0f} img/s  {d['ms_per_step']:.2f} ms/step  e2e {d['e2e']['value']:.0f}  "
          f"ddim50 {d.get('ddim50', {}).get('img_per_s', 0):.0f} img/s  sgemm {d['roofline']['achieved']:.0f} TF/s "
          f"clocks {d.get('clocks')}")
    fam = d["roofline"]["families"]
    print("   " + "  ".join(f"{k}:{v['ms']:.2f}" for k, v in list(fam.items())[:14]))
