"""Hot spots of one kernel from an ncu report's SASS page: python tools/ncu_hot.py report.ncu-rep [launch_index] [top_n]
Groups SASS instructions into runs of 48 and prints executed-instruction and stall-sample shares."""
import csv
import subprocess
import sys

rep = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else -1
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
heads = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
start = heads[which]
pos = heads.index(start)
end = heads[pos + 1] if pos + 1 < len(heads) else len(rows)
h = rows[start]
ia, isrc, isamp, iex = h.index("Address"), h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
body = [r for r in rows[start + 1:end] if len(r) == len(h)]
tot_s = sum(int(r[isamp] or 0) for r in body)
tot_e = sum(int(r[iex] or 0) for r in body)
print("instructions in kernel:", len(body), "executed (warp-level):", tot_e, "samples:", tot_s)
B = 48
buckets = []
for i in range(0, len(body), B):
    chunk = body[i:i + B]
    s = sum(int(r[isamp] or 0) for r in chunk); e = sum(int(r[iex] or 0) for r in chunk)
    ops = {}
    for r in chunk:
        t = r[isrc].split()
        op = (t[1] if t and t[0].startswith("@") and len(t) > 1 else (t[0] if t else "?")).split(".")[0]
        ops[op] = ops.get(op, 0) + 1
    topops = ",".join("%s%d" % kv for kv in sorted(ops.items(), key=lambda kv: -kv[1])[:6])
    buckets.append((i, s, e, topops))
for i, s, e, ops in sorted(buckets, key=lambda b: -b[1])[:top]:
    print("sass[%5d..] samples %5.1f%% exec %5.1f%%  %s" % (i, 100.0 * s / max(tot_s, 1), 100.0 * e / max(tot_e, 1), ops))
stall_cols = [c for c in h if c.startswith("stall_") and "Not Issued" not in c]
agg = {c: sum(int(r[h.index(c)] or 0) for r in body) for c in stall_cols}
print({k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]})
