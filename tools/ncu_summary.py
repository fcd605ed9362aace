"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (count, total, share, avg)."""
import collections
import csv
import sys

rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    v = float(r[vi].replace(",", ""))
    v = v / 1e3 if r[ui] in ("ns", "nsecond") else (v * 1e3 if r[ui] in ("ms", "msecond") else v)
    a = agg.setdefault(r[ki][:70], [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
print(f"# {sys.argv[1]}: {sum(a[0] for a in agg.values())} launches, {tot/1e3:.2f} ms under ncu (cold-cache, serialised: compare shares)")
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{k:70s} n={a[0]:4d} total={a[1]/1e3:9.3f} ms share={a[1]/tot:6.1%} avg={a[1]/a[0]:10.1f} us")
