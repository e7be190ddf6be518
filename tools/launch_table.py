"""Aggregate an ncu CSV (gpu__time_duration.sum, dram__bytes_read.sum, dram__bytes_write.sum per launch) over the LAST
n launches of each kernel name: `python tools/launch_table.py file.csv [tail_fraction]` -> kernel, launches, total us,
DRAM MB, GB/s."""
import csv
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
h = rows[0]
ki, mi, vi, ui, ii = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit"), h.index("ID")
per = OrderedDict()
for r in rows[1:]:
    d = per.setdefault(int(r[ii]), {"name": r[ki].split("(")[0].replace("void ", "")[:60]})
    v = float(r[vi].replace(",", ""))
    u = r[ui]
    if r[mi].startswith("gpu__time"):
        d["us"] = v * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u, 1e-3)
    else:
        d[r[mi]] = v * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1.0)
ids = sorted(per)
frac = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
ids = ids[int(len(ids) * (1 - frac)):]
agg = OrderedDict()
for i in ids:
    d = per[i]
    a = agg.setdefault(d["name"], [0, 0.0, 0.0])
    a[0] += 1
    a[1] += d.get("us", 0.0)
    a[2] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
tot = sum(a[1] for a in agg.values())
for n, (c, us, by) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-60s %4d x %9.1f us total %5.1f %%  %8.1f MB  %7.0f GB/s" % (n, c, us, 100 * us / tot, by / 1e6, by / us / 1e3 if us else 0))
print("total %.1f us over %d launches" % (tot, len(ids)))
