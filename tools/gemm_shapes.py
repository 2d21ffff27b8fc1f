"""Join the ENCDIFF_GEMM_LOG shape log of one profiled step with the ncu launch list (same order) and print, per distinct
GEMM shape, count / avg time / achieved TFLOP/s.  usage: python tools/gemm_shapes.py launches.csv gemmlog.txt"""
import collections
import csv
import re
import sys


def main():
    rows = csv.DictReader([l for l in open(sys.argv[1]) if not l.startswith("==")])
    dur = {"sgemm2": [], "wgrad2": []}
    seen = set()
    for r in rows:
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        for k in dur:
            if k + "_kernel" in r["Kernel Name"]:
                dur[k].append(float(r["Metric Value"].replace(",", "")) / 1e3)   # us
    logs = {"sgemm2": [], "wgrad2": []}
    for l in open(sys.argv[2]):
        m = re.match(r"GEMMLOG (\w+) (.*)", l.strip())
        if m:
            logs[m.group(1)].append(dict(kv.split("=") for kv in m.group(2).split()))
    for k in dur:
        n = len(dur[k])
        lg = logs[k][-n:]                       # the profiled step is the last one logged
        assert len(lg) == n, (k, len(lg), n)
        agg = collections.OrderedDict()
        for d, t in zip(lg, dur[k]):
            key = tuple(d.items())
            a = agg.setdefault(key, [0, 0.0])
            a[0] += 1
            a[1] += t
        print(f"== {k}: {n} launches, {sum(dur[k]) / 1e3:.3f} ms")
        for key, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            d = {a: int(b) for a, b in key}
            if k == "sgemm2":
                fl = 2.0 * d["M"] * d["N"] * d["K"]
            else:
                fl = 2.0 * d["m_tiles"] * 128 * d["Cp"] * d["Cq"] * d["taps"]
            print(f"  n={c:3d} total {t:8.1f} us avg {t / c:6.1f} us  {fl / (t / c) / 1e6:7.1f} TF/s  " +
                  " ".join(f"{a}={b}" for a, b in key))


if __name__ == "__main__":
    main()
