"""Instruction mix (per warp) of one kernel launch from an ncu report's SASS page.
usage: python tools/ncu_mix.py report.ncu-rep [launch-index]"""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
skip = sys.argv[2] if len(sys.argv) > 2 else "0"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", skip, "--launch-count", "1"],
                     capture_output=True, text=True).stdout.splitlines()
rows = list(csv.reader(out))
hdr = rows[1]
i_src, i_ex = hdr.index("Source"), hdr.index("Instructions Executed")
data = [(r[i_src].strip(), int(r[i_ex])) for r in rows[2:] if len(r) > i_ex and r[i_ex].isdigit()]
data = data[:len(data) // 2]          # the page lists the function twice
mx = max(e for _, e in data)
tot = sum(e for _, e in data)
print(rows[0][1], "| warps", mx, "| instructions per warp", round(tot / mx, 1))
ops = collections.Counter()
for s, e in data:
    parts = s.split()
    op = parts[1] if parts[0].startswith("@") else parts[0]
    ops[op.split(".")[0]] += e
print("  ".join(f"{k} {v / mx:.0f}" for k, v in ops.most_common(24)))
