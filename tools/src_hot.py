"""Opcode / hot-line summary of `ncu -i X.ncu-rep --page source --csv` output.
`python tools/src_hot.py src.csv [kernel index] [n lines]`"""
import csv
import sys
from collections import Counter

rows = list(csv.reader(open(sys.argv[1])))
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
nl = int(sys.argv[3]) if len(sys.argv) > 3 else 30
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]
k, end = starts[which], starts[which + 1]
print(rows[k][1][:100])
hdr = rows[k + 1]
si, ei, wi = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
body = rows[k + 2:end]
num = lambda x: int(x) if x else 0
tot = sum(num(r[ei]) for r in body)
tots = sum(num(r[wi]) for r in body)
print("warp instructions", tot, "stall samples", tots)
c, cs = Counter(), Counter()
for r in body:
    t = r[si].split()
    op = t[1] if t and t[0].startswith("@") else (t[0] if t else "?")
    c[op] += num(r[ei])
    cs[op] += num(r[wi])
for op, v in c.most_common(22):
    print("%-22s %10d %5.1f%%  samples %5.1f%%" % (op, v, 100 * v / max(tot, 1), 100 * cs[op] / max(tots, 1)))
print("--- hottest instructions by stall samples")
order = sorted(range(len(body)), key=lambda i: -num(body[i][wi]))[:nl]
for i in sorted(order):
    r = body[i]
    print("%5d %6.2f%% %9d  %s" % (i, 100 * num(r[wi]) / max(tots, 1), num(r[ei]), r[si][:90]))
