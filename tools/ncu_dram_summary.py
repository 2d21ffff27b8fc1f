"""Summarise `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv` of ONE training step
(tools/one_step.py) per kernel: launches, total / average time, DRAM bytes per launch.  Writes the dominant kernel's
figures to a JSON file bench.py reads for `roofline.traffic`.
usage: python tools/ncu_dram_summary.py launches.csv [out.json] [kernel-substring, default sgemm2_kernel]"""
import collections
import csv
import json
import re
import sys


def to_bytes(value, unit):
    v = float(value.replace(",", ""))
    u = unit.strip().lower()
    return v * {"byte": 1, "bytes": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "": 1}.get(u, 1)


def to_ns(value, unit):
    v = float(value.replace(",", ""))
    return v * {"ns": 1, "nsecond": 1, "us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6, "s": 1e9, "second": 1e9}.get(unit.strip().lower(), 1)


def main():
    path = sys.argv[1]
    out = sys.argv[2] if len(sys.argv) > 2 else None
    dom = sys.argv[3] if len(sys.argv) > 3 else "sgemm2_kernel"
    rows = csv.DictReader([l for l in open(path) if not l.startswith("==")])
    K = collections.defaultdict(lambda: {"name": "", "ns": 0.0, "rd": 0.0, "wr": 0.0})
    for r in rows:
        k = K[int(r["ID"])]
        k["name"] = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "").replace("edb::", "")
        m = r["Metric Name"]
        if m == "gpu__time_duration.sum":
            k["ns"] = to_ns(r["Metric Value"], r["Metric Unit"])
        elif m == "dram__bytes_read.sum":
            k["rd"] = to_bytes(r["Metric Value"], r["Metric Unit"])
        elif m == "dram__bytes_write.sum":
            k["wr"] = to_bytes(r["Metric Value"], r["Metric Unit"])
    agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
    for k in K.values():
        a = agg[k["name"]]
        a[0] += 1
        a[1] += k["ns"]
        a[2] += k["rd"]
        a[3] += k["wr"]
    tot = sum(a[1] for a in agg.values())
    print(f"# {len(K)} launches, {tot / 1e6:.3f} ms (serialised, cold-cache ncu replay times)")
    print(f"# {'kernel':44s} {'n':>5s} {'ms':>8s} {'share':>6s} {'avg us':>8s} {'DRAM MB/launch':>15s} {'GB/s':>7s}")
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        mb = (a[2] + a[3]) / a[0] / 1e6
        print(f"  {name[:44]:44s} {a[0]:5d} {a[1] / 1e6:8.3f} {100 * a[1] / tot:5.1f}% {a[1] / a[0] / 1e3:8.1f} {mb:15.2f} "
              f"{(a[2] + a[3]) / max(a[1], 1):7.0f}")
    if out:
        hit = [(n, a) for n, a in agg.items() if dom in n]
        if hit:
            n = sum(a[0] for _, a in hit)
            d = {"kernel": dom, "launches": n, "dram_bytes_per_launch": sum(a[2] + a[3] for _, a in hit) / n,
                 "dram_read_bytes_per_launch": sum(a[2] for _, a in hit) / n,
                 "dram_write_bytes_per_launch": sum(a[3] for _, a in hit) / n,
                 "avg_ns_under_ncu": sum(a[1] for _, a in hit) / n,
                 "share_of_step_under_ncu": sum(a[1] for _, a in hit) / tot, "source": path}
            json.dump(d, open(out, "w"), indent=1)
            print("wrote", out, d)


if __name__ == "__main__":
    main()
