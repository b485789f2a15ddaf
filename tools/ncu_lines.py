"""Per-instruction stall samples of one kernel in an ncu report: python tools/ncu_lines.py report.ncu-rep [min_samples]"""
import csv, subprocess, sys
rep = sys.argv[1]; mn = int(sys.argv[2]) if len(sys.argv) > 2 else 20
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][-1]
h = rows[hi]; body = [r for r in rows[hi + 1:] if len(r) == len(h)]
isrc, isamp, iex = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
cols = [c for c in h if c.startswith("stall_") and "Not Issued" not in c]
tot = sum(int(r[isamp] or 0) for r in body)
print("total samples", tot)
for k, r in enumerate(body):
    s = int(r[isamp] or 0)
    if s >= mn:
        st = sorted(((c[6:], int(r[h.index(c)] or 0)) for c in cols), key=lambda kv: -kv[1])[:2]
        print("%5d %5.1f%% ex=%-8s %-70s %s" % (k, 100.0 * s / tot, r[iex], r[isrc][:70], st))
