"""Summarise an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv`): kernels of the LAST step
(from the last preprocess_fwd on), microseconds each.  `python tools/launch_summary.py launches.csv`"""
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
names = []
for r in rows[1:]:
    v = float(r[vi].replace(",", ""))
    v = {"ns": v / 1e3, "us": v, "ms": v * 1e3}.get(r[ui], v / 1e3)
    names.append((r[ki].split("(")[0][:48], v))
first = max(i for i, (n, v) in enumerate(names) if "preprocess_fwd" in n)
tot = sum(v for n, v in names[first:])
for n, v in names[first:]:
    print("%-50s %8.1f us %5.1f %%" % (n, v, 100 * v / tot))
print("%-50s %8.1f us" % ("total", tot))
