"""profiles/rNN_traffic.json from `ncu -i X.ncu-rep --page raw --csv` of the full-size bench workload:
per kernel, DRAM bytes (read + write) of its first launch.  usage: ncu_traffic.py raw.csv streams hours sample_rate chunks_per_stream > out.json"""
import csv
import json
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
ki, ri, wi, di = (hdr.index(k) for k in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum"))
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
tscale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0}
out = {}
for r in rows[2:]:
    name = r[ki].split("(")[0].split("::")[-1].split("<")[0].replace("void ", "").strip()
    if name in out:
        continue
    rd = float(r[ri].replace(",", "")) * scale[units[ri]]
    wr = float(r[wi].replace(",", "")) * scale[units[wi]]
    out[name] = {"dram_bytes_per_launch": rd + wr, "dram_read": rd, "dram_write": wr,
                 "ncu_duration_ms": float(r[di].replace(",", "")) * tscale[units[di]],
                 "chunks_per_stream": int(sys.argv[5])}
json.dump({"source": "ncu --set full --clock-control none, first launch of each kernel in one bench step",
           "workload": {"streams": int(sys.argv[2]), "hours": float(sys.argv[3]), "sample_rate": int(sys.argv[4])},
           "kernels": out}, sys.stdout, indent=1)
