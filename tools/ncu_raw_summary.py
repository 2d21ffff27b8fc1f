"""One line per launch from `ncu -i report.ncu-rep --page raw --csv`: time, DRAM bytes and GB/s, L2 hit rate, throughput
percentages, tensor-pipe activity, registers, dynamic shared memory, achieved occupancy (tools only).
usage: python tools/ncu_raw_summary.py raw.csv [kernel-substring]"""
import csv
import re
import sys

COLS = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "rdMB"), ("dram__bytes_write.sum", "wrMB"),
        ("lts__t_sector_hit_rate.pct", "L2hit%"), ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
        ("launch__registers_per_thread", "regs"), ("launch__shared_mem_per_block_dynamic", "dsmemKB"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%")]
SCALE = {"byte": 1e-6, "kbyte": 1e-3, "mbyte": 1.0, "gbyte": 1e3, "ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3,
         "usecond": 1.0, "msecond": 1e3, "second": 1e6}


def main():
    rows = list(csv.reader([l for l in open(sys.argv[1]) if not l.startswith("==")]))
    flt = sys.argv[2] if len(sys.argv) > 2 else ""
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    print("# kernel | grid | " + " | ".join(n for _, n in COLS) + " | GB/s")
    for r in rows[2:]:
        name = re.sub(r"\(.*", "", r[idx["Kernel Name"]]).replace("void ", "").replace("edb::", "").replace("<unnamed>::", "")
        if flt and flt not in name:
            continue
        vals = []
        for key, short in COLS:
            if key not in idx:
                vals.append(float("nan"))
                continue
            v = float(r[idx[key]].replace(",", "") or "nan")
            u = units[idx[key]].strip().lower()
            if short in ("us", "rdMB", "wrMB") and u in SCALE:
                v *= SCALE[u]
            if short == "dsmemKB":
                v *= {"byte": 1e-3, "kbyte": 1.0, "mbyte": 1e3}.get(u.split("/")[0], 1.0)
            vals.append(v)
        gbs = (vals[1] + vals[2]) * 1e6 / (vals[0] * 1e-6) / 1e9 if vals[0] > 0 else float("nan")
        print(f"{name[:44]:44s} | {r[idx['Grid Size']]:>12s} | " + " | ".join(f"{v:8.2f}" for v in vals) + f" | {gbs:7.0f}")


if __name__ == "__main__":
    main()
