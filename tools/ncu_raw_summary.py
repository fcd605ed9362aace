"""Summarise `ncu -i X.ncu-rep --page raw --csv` per launch: duration, DRAM bytes, tensor-pipe %, occupancy."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
want = [("Kernel Name", "kernel"), ("Grid Size", "grid"), ("Block Size", "block"),
        ("gpu__time_duration.sum", "dur"), ("dram__bytes_read.sum", "rd"), ("dram__bytes_write.sum", "wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("launch__registers_per_thread", "regs")]
idx = [(hdr.index(k), n) for k, n in want if k in hdr]
print("# " + sys.argv[1] + " (ncu --set full --clock-control none; per launch)")
for r in rows[2:]:
    out = []
    for i, n in idx:
        v = r[i]
        if n == "kernel":
            v = v.split("(")[0].split("::")[-1].strip()[:28]
        elif n in ("grid", "block"):
            v = v.replace(" ", "")
        else:
            try:
                v = f"{float(v.replace(',', '')):.3f}{units[i] if n in ('dur', 'rd', 'wr') else ''}"
            except ValueError:
                pass
        out.append(f"{n}={v}")
    print("  ".join(out))
