"""Top stall locations of one kernel launch from an ncu report's SASS source page.
usage: python tools/ncu_hot.py report.ncu-rep [launch-index] [top-n]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
skip = sys.argv[2] if len(sys.argv) > 2 else "0"
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", skip, "--launch-count", "1"],
                     capture_output=True, text=True).stdout.splitlines()
rows = list(csv.reader(out))
hdr = rows[1]
i_src, i_s, i_ex = hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
data = []
for k, r in enumerate(rows[2:]):
    try:
        data.append((int(r[i_s]), k, r[i_src].strip(), int(r[i_ex])))
    except (ValueError, IndexError):
        pass
tot = sum(d[0] for d in data)
print(rows[0][1], "total samples", tot, "instructions", len(data))
for s, k, src, ex in sorted(data, reverse=True)[:top]:
    print(f"{100.0 * s / max(tot, 1):5.1f}%  #{k:5d}  exec {ex:8d}  {src}")
