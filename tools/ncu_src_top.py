"""Top stalled SASS instructions of one launch: `ncu -i X.ncu-rep --page source --csv --print-source sass --launch-skip N --launch-count 1 > f.csv`."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
hdr = next(r for r in rows if "# Samples" in r)
si, src, ie = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = [r for r in rows if len(r) == len(hdr) and r[si].isdigit()]
print("total samples", sum(int(r[si]) for r in data), "instructions", len(data))
top = sorted(enumerate(data), key=lambda x: -int(x[1][si]))[:top_n]
for n, r in sorted(top):
    st = sorted(((int(r[i]), hdr[i][6:]) for i in stall), reverse=True)[:2]
    print(n, r[src].strip()[:72].ljust(72), r[si], r[ie], st)
