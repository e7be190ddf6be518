"""Summarise an ncu launch list (`ncu --metrics gpu__time_duration.sum[,dram__bytes_read.sum,dram__bytes_write.sum] --csv`):
kernels of the LAST step (from the last preprocess_fwd on), microseconds each and, when captured, the DRAM bytes of the
launch.  `python tools/launch_summary.py launches.csv [traffic.json]` (the optional second argument writes the per-kernel
DRAM bytes per launch that bench.py's roofline.traffic reads)."""
import csv
import json
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
ii, ki, mi, vi, ui = (hdr.index(k) for k in ("ID", "Kernel Name", "Metric Name", "Metric Value", "Metric Unit"))
launches = OrderedDict()
for r in rows[1:]:
    v = float(r[vi].replace(",", ""))
    if r[mi].startswith("gpu__time"):
        v = {"ns": v / 1e3, "us": v, "ms": v * 1e3, "nsecond": v / 1e3, "usecond": v, "msecond": v * 1e3}.get(r[ui], v / 1e3)
    else:
        v = {"byte": v, "Kbyte": v * 1e3, "Mbyte": v * 1e6, "Gbyte": v * 1e9}.get(r[ui], v)
    launches.setdefault(r[ii], {"name": r[ki].split("(")[0].replace("void ", "").replace("ws::", "")[:44]})[r[mi]] = v
L = list(launches.values())
first = max(i for i, l in enumerate(L) if "preprocess_fwd" in l["name"])
step = L[first:]
tot = sum(l.get("gpu__time_duration.sum", 0.0) for l in step)
print("%-46s %9s %6s %10s %10s" % ("kernel", "us", "%", "DRAM rd MB", "DRAM wr MB"))
agg = {}
for l in step:
    t = l.get("gpu__time_duration.sum", 0.0)
    rd, wr = l.get("dram__bytes_read.sum"), l.get("dram__bytes_write.sum")
    print("%-46s %9.1f %6.1f %10s %10s" % (l["name"], t, 100 * t / tot, "-" if rd is None else "%.1f" % (rd / 1e6),
                                          "-" if wr is None else "%.1f" % (wr / 1e6)))
    a = agg.setdefault(l["name"], [0, 0.0, 0.0])
    a[0] += 1; a[1] += t; a[2] += (rd or 0) + (wr or 0)
print("%-46s %9.1f us (ncu serialises launches and runs each cold: shares, not absolutes)" % ("total", tot))
if len(sys.argv) > 2:
    stage_of = {"render_bwd": "render_bwd", "render_fwd": "render_fwd", "preprocess_fwd": "preprocess", "preprocess_bwd": "preprocess_bwd",
                "emit_tiles_ordered": "emit_keys", "identify_ranges4": "ranges_pack"}
    by = {}
    for name, (n, t, b) in agg.items():
        base = name.split("<")[0]
        if base in stage_of:
            by[stage_of[base]] = b / n
    by["sort"] = sum(b for name, (n, t, b) in agg.items() if name.startswith("rs_"))
    json.dump({"note": "ncu dram__bytes_read.sum + dram__bytes_write.sum per launch, config 2 view 17 (R = 8.4 M), tools/profile_step.py, "
                       "captured with the build of the same commit (profiles/r02_launches_cfg2_step.csv)",
               "bytes_per_launch": by}, open(sys.argv[2], "w"), indent=1)
