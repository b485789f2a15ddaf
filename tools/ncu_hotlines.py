"""Hot instructions (>= frac of the stall samples) of one kernel in an ncu report, with the loads around them:
python tools/ncu_hotlines.py report.ncu-rep <kernel regex> [frac]"""
import csv, subprocess, sys
rep, rx = sys.argv[1], sys.argv[2]; frac = float(sys.argv[3]) if len(sys.argv) > 3 else 0.02
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name", "regex:" + rx], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
heads = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
for n, hi in enumerate(heads):
    h = rows[hi]; end = heads[n + 1] if n + 1 < len(heads) else len(rows)
    body = [r for r in rows[hi + 1:end] if len(r) == len(h)]
    isrc, isamp, iex = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
    cols = [c for c in h if c.startswith("stall_") and "Not Issued" not in c]
    tot = sum(int(r[isamp] or 0) for r in body)
    print("== launch", n, "samples", tot, "instructions", len(body))
    for k, r in enumerate(body):
        s = int(r[isamp] or 0)
        if s >= tot * frac:
            st = sorted(((c[6:], int(r[h.index(c)] or 0)) for c in cols), key=lambda kv: -kv[1])[:2]
            print("%5d %5.1f%% ex=%-7s %-64s %s" % (k, 100.0 * s / tot, r[iex], r[isrc][:64], st))
