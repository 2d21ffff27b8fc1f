"""Instructions executed and stall samples per CUDA source line of one kernel launch (ncu report with --import-source on).
usage: python tools/ncu_lines.py report.ncu-rep [top-n]"""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout.splitlines()
rows = list(csv.reader(out))
cur_file = ""
agg = collections.OrderedDict()
hdr = None
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r and r[0] == "Line No":
        hdr = r
        i_ex = hdr.index("Instructions Executed")
        i_st = hdr.index("Warp Stall Sampling (All Samples)")
        continue
    if hdr is None or len(r) < len(hdr) - 2:
        continue
    if r[0].strip().isdigit():          # a CUDA source line followed by its SASS rows
        key = (cur_file, int(r[0]), r[1].strip()[:110])
        try:
            ex, st = int(r[i_ex]), int(r[i_st])
        except ValueError:
            continue
        a = agg.setdefault(key, [0, 0])
        a[0] += ex
        a[1] += st
tot_ex = sum(a[0] for a in agg.values()) or 1
tot_st = sum(a[1] for a in agg.values()) or 1
print(f"total warp-instructions {tot_ex}, stall samples {tot_st}")
for (f, ln, src), (ex, st) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{100.0 * ex / tot_ex:5.1f}% instr {100.0 * st / tot_st:5.1f}% stall  {f}:{ln:<4d} {src}")
