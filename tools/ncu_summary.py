"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list: the last step's launches in order and per-kernel totals.
   python tools/ncu_summary.py launches.csv [launches_per_step]"""
import collections
import csv
import re
import sys

lines = [l for l in open(sys.argv[1]) if l.startswith('"')]
r = csv.reader(lines)
hdr = next(r)
ki, vi, gi, bi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size"), hdr.index("Block Size")
rows = []
for row in r:
    name = re.sub(r"\(.*", "", row[ki]).replace("void ", "").replace("cvt::<unnamed>::", "").replace("cvt::", "")
    rows.append((name, float(row[vi].replace(",", "")) / 1e3, row[gi], row[bi]))
own = [x for x in rows if not x[0].startswith("at::")]
per = int(sys.argv[2]) if len(sys.argv) > 2 else 0
if per:
    last = own[-per:]
    print("# last step, in launch order: kernel, us, grid, block")
    for x in last:
        print("%-48s %9.1f  %-18s %s" % x)
    print("# step total us: %.1f" % sum(x[1] for x in last))
    own = last
agg = collections.defaultdict(lambda: [0, 0.0])
for n, t, _, _ in own:
    agg[n][0] += 1; agg[n][1] += t
tot = sum(v[1] for v in agg.values())
print("# per kernel: name, launches, total_us, share")
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-48s %5d %10.1f %6.2f%%" % (k, c, t, 100 * t / tot))
