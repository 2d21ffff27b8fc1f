"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by (kernel, grid) (tools only)."""
import collections
import csv
import re
import sys


def main():
    path = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    flt = sys.argv[3] if len(sys.argv) > 3 else ""
    lines = [l for l in open(path) if not l.startswith("==")]
    L = {}
    for row in csv.DictReader(lines):
        if row["Metric Name"] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("void ", "").replace("edb::", "")
        L[int(row["ID"])] = (name, row["Grid Size"], float(row["Metric Value"].replace(",", "")))
    tot = sum(v[2] for v in L.values())
    print(f"# launches {len(L)} total {tot / 1e3:.1f} us")
    byk = collections.defaultdict(lambda: [0, 0.0])
    agg = collections.defaultdict(lambda: [0, 0.0, 1e18, 0.0])
    for name, grid, d in L.values():
        byk[name][0] += 1
        byk[name][1] += d
        a = agg[(name, grid)]
        a[0] += 1
        a[1] += d
        a[2] = min(a[2], d)
        a[3] = max(a[3], d)
    for k, v in sorted(byk.items(), key=lambda kv: -kv[1][1])[:25]:
        print(f"{v[1] / 1e3:9.1f} us  n={v[0]:4d}  {100 * v[1] / tot:5.1f}%  {k[:60]}")
    print("# by (kernel, grid)")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        if flt and flt not in k[0]:
            continue
        top -= 1
        if top < 0:
            break
        print(f"{v[1] / 1e3:9.1f} us  n={v[0]:3d}  avg {v[1] / v[0] / 1e3:7.1f}  min {v[2] / 1e3:6.1f} max {v[3] / 1e3:6.1f}  "
              f"{100 * v[1] / tot:4.1f}%  {k[0][:34]:34s} grid {k[1]}")


if __name__ == "__main__":
    main()
