#!/usr/bin/env python
"""DRAM traffic per kernel group, per launch, from an ncu capture of ONE profiled step (profiles/ncu_traffic.json, read by bench.py's
`roofline.traffic`).

  step 1 (GPU box, under ncu):
     ncu --profile-from-start off --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
         --csv --log-file gpurun_out/traffic_train.csv python tools/ncu_traffic.py --run train --out gpurun_out/traffic_train_groups.json
  step 2 (anywhere):  python tools/ncu_traffic.py --join gpurun_out --out profiles/ncu_traffic.json --tag r02

--run issues warm-up steps, then cudaProfilerStart, ONE eager cvt_profile_step (no CUDA graph, so ncu sees the kernels in issue order),
cudaProfilerStop, and writes the (group, kernels launched) list the library recorded; --join cuts ncu's launch list by those counts."""
import argparse
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(mode, out, batch, precision):
    import torch
    import transformer_stm_b200 as T
    train = mode == "train"
    eng = T.CvtEngine(T.CvtConfig.reference(precision=precision, max_batch=batch))
    eng.init_params(1234)
    eng.set_dropout("philox" if train else "off", 1234)
    g = torch.Generator(device="cuda"); g.manual_seed(0)
    img = torch.rand((batch, 128, 128, 1), device="cuda", generator=g); proc = torch.randn((batch, 5), device="cuda", generator=g)
    y = 1000 + 100 * torch.randn((batch, 1), device="cuda", generator=g)
    for _ in range(3):
        eng.profile_step(img, proc, y, batch, train)
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    eng.profile_step(img, proc, y, batch, train)
    eng.synchronize()
    torch.cuda.profiler.stop()
    json.dump({"mode": mode, "batch": batch, "precision": precision, "groups": eng.profile_group_launches()}, open(out, "w"))
    eng.close()


def join(folder, out, tag):
    res = {"source": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum of one eager cvt_profile_step (%s)" % tag, "unit": "bytes per launch of the group"}
    for mode in ("train", "infer"):
        gp, cp = os.path.join(folder, "traffic_%s_groups.json" % mode), os.path.join(folder, "traffic_%s.csv" % mode)
        if not (os.path.exists(gp) and os.path.exists(cp)):
            continue
        groups = json.load(open(gp))["groups"]
        rows = []
        with open(cp) as f:
            lines = [l for l in f if not l.startswith("==")]
        rd = csv.DictReader(lines)
        per = {}
        order = []
        for r in rd:
            kid = r["ID"]
            if kid not in per:
                per[kid] = {"name": r["Kernel Name"], "bytes": 0.0, "ns": 0.0}
                order.append(kid)
            v = float(r["Metric Value"].replace(",", ""))
            unit = r.get("Metric Unit", "")
            if r["Metric Name"].startswith("dram__bytes"):
                per[kid]["bytes"] += v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
            elif r["Metric Name"].startswith("gpu__time_duration"):
                per[kid]["ns"] += v * {"ns": 1, "us": 1e3, "ms": 1e6}.get(unit, 1)
        ks = [per[k] for k in order]
        tot = sum(c for _, c in groups)
        if tot != len(ks):
            print("warning: %s: library counted %d launches, ncu saw %d" % (mode, tot, len(ks)), file=sys.stderr)
        i, table, detail = 0, {}, {}
        for name, cnt in groups:
            chunk = ks[i:i + cnt]; i += cnt
            table[name] = table.get(name, 0.0) + sum(k["bytes"] for k in chunk)
            detail.setdefault(name, []).extend({"kernel": k["name"].split("(")[0][:60], "bytes": k["bytes"], "us": k["ns"] / 1e3} for k in chunk)
        res[mode] = table
        res[mode + "_kernels"] = detail
    json.dump(res, open(out, "w"), indent=1)
    print("wrote", out)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--run", choices=["train", "infer"])
    ap.add_argument("--join")
    ap.add_argument("--out", required=True)
    ap.add_argument("--tag", default="")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--precision", default="tf32")
    a = ap.parse_args()
    if a.run:
        run(a.run, a.out, a.batch, a.precision)
    else:
        join(a.join, a.out, a.tag)
