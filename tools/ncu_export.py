"""Key metrics of every launch in an ncu report as one JSON object per launch: python tools/ncu_export.py report.ncu-rep > summary.jsonl"""
import csv, json, subprocess, sys
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__occupancy_limit_shared_mem", "sm__cycles_elapsed.max", "smsp__inst_executed.sum"]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, u = rows[0], rows[1]
for r in rows[2:]:
    d = {"kernel": r[h.index("Kernel Name")][:90], "grid": r[h.index("Grid Size")], "block": r[h.index("Block Size")]}
    for k in KEYS:
        if k in h:
            i = h.index(k)
            try:
                d[k + " [" + u[i] + "]"] = float(r[i].replace(",", ""))
            except ValueError:
                d[k] = r[i]
    print(json.dumps(d))
