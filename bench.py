#!/usr/bin/env python
"""bench.py — CvT(Par) training-step throughput (images/s) on N B200s, the metric BASELINE.json names.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one rank per GPU under torchrun)
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle port) on the host cores

One "step" = one Keras train_step of create_cvt_model's network (models/CvT(Par).py:292-354,464-476): forward
(training) + MSE + backward + Adam + BN moving update on one batch of 64 synthetic 128x128x1 images + 5 process
parameters per GPU (configs[1]; weak scaling: per-GPU batch fixed).  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "cvt_par_train_images_per_sec"
UNIT = "images/s"
BATCH_PER_GPU = 64


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=float(d["hbm_gbs"]), bf16=float(d["bf16_tflops"]), bf16_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), src="measured")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sustained=1400.0, src="fallback")


class ClockSampler(threading.Thread):
    """Samples nvidia-smi SM clocks + throttle reasons while the timed region runs."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag, self.proc = index, [], False, None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                if self.stop_flag:
                    break
                self.samples.append([x.strip() for x in line.split(",")])
        except Exception:
            pass

    def finish(self):
        self.stop_flag = True
        if self.proc:
            try:
                self.proc.terminate()
            except Exception:
                pass
        sm, mx, reasons = [], 0.0, set()
        for s in self.samples:
            try:
                sm.append(float(s[0])); mx = max(mx, float(s[1]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), s[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


# ---- algorithmic work of the kernel groups (SURVEY 8d conventions: activations read+written once, fp32) -----------
def group_work(geo, batch, mlp_ratio=4):
    """{group: (flops, bytes)} per launch-group for a forward pass over `batch` images; backward groups = 2x."""
    out = {}
    for i, g in enumerate(geo, 1):
        T, Tc, D, H = g["t"], g["tc"], g["d"], g["heads"]
        kk = g["k"] * g["k"] * g["cin"]
        B = batch
        out["s%d.conv_embed" % i] = (2.0 * B * T * kk * D, 4.0 * B * (g["hin"] * g["win"] * g["cin"] + T * D))
        out["s%d.ln_proj_bn" % i] = (B * Tc * D * (8 + 3 * 22.0), 4.0 * B * Tc * D * 5)            # read x0, write xn + 3 conv maps
        out["s%d.qkv_gemm" % i] = (2.0 * B * Tc * D * D * 6, 4.0 * B * Tc * D * 9)                 # 3 in, 3 mid (kept for bwd), 3 out
        out["s%d.attention" % i] = (4.0 * B * H * Tc * Tc * 64, 4.0 * B * Tc * D * 4)
        out["s%d.out_gemm" % i] = (2.0 * B * Tc * D * D * 2, 4.0 * B * Tc * D * 4)
        out["s%d.mlp" % i] = (2.0 * B * Tc * D * D * mlp_ratio * 2 + 8.0 * B * Tc * D, 4.0 * B * Tc * D * (3 + 2 * mlp_ratio))
    return out


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--precision", default=os.environ.get("CVT_PRECISION", "tf32"), choices=["fp32", "tf32", "bf16"])
    ap.add_argument("--batch", type=int, default=BATCH_PER_GPU)
    ap.add_argument("--mode", default="train", choices=["train", "infer"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-flush", action="store_true", help="do not flush L2 between timed steps")
    return ap.parse_args()


# ---- the reference's CPU path (oracle port; TensorFlow is unavailable on this box) -----------------
def cpu_reference_run(steps, warmup, batch, mode, budget_s=25.0):
    import torch
    from oracle import cvt_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    spec = O.ref_spec()
    params = O.init_params(spec, 1234, "keras")
    img, proc, y = O.synthetic_batch(spec, batch, 0)
    st = O.new_opt_state(params)

    def one():
        if mode == "train":
            O.train_step(spec, params, st, img, proc, y, 1e-3, None, dtype=torch.float32)
        else:
            O.forward(spec, params, img, proc, False, None, dtype=torch.float32)

    times = []
    t_all = time.time()
    for _ in range(max(warmup, 1)):
        one()
        if time.time() - t_all > budget_s:
            break
    for _ in range(max(steps, 1)):
        t0 = time.time(); one(); times.append(time.time() - t0)
        if time.time() - t_all > 2 * budget_s:
            break
    ms = 1e3 * float(np.median(times))
    return dict(value=batch / (ms / 1e3), ms_per_step=ms, cores=cores, steps=len(times),
                sample="%d %s step(s) of batch %d, fp32 torch-CPU oracle port, %d threads" % (len(times), mode, batch, cores))


def main_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = cpu_reference_run(min(a.steps, 5), min(a.warmup, 2), a.batch, a.mode)
    line = {"impl": "reference", "metric": METRIC if a.mode == "train" else "cvt_par_infer_images_per_sec", "value": r["value"], "unit": UNIT,
            "n_gpus": a.gpus, "steps": r["steps"], "warmup": min(a.warmup, 2), "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "CvT(Par) %s step, batch %d, 128x128x1 + 5 process parameters (configs[1])" % (a.mode, a.batch)},
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "TensorFlow/Keras unavailable on this box: the reference's CPU path is timed through the oracle port of its graph"}
    print(json.dumps(line))


# ---- this repo's arm ---------------------------------------------------------------------------------------------------
def main_b200(a):
    import torch
    import transformer_stm_b200 as T

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries exactly one JSON line: anything libraries print there (e.g. NCCL's version banner) goes to stderr instead
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this arm has no CPU fallback (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B, train = a.batch, a.mode == "train"
    cfg = T.CvtConfig.reference(precision=a.precision, max_batch=B, device=local)
    eng = T.CvtEngine(cfg)
    eng.init_params(1234)                                   # Keras initialisers, random-init weights
    eng.set_dropout("philox" if train else "off", seed=1234 + rank)
    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.tensor(list(eng.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        eng.comm_init(bytes(uid.cpu().tolist()), rank, world)
        # same initial weights on every rank
        flat = torch.cat([torch.tensor(v).reshape(-1) for v in eng.get_params().values()]).cuda()
        dist.broadcast(flat, 0)
        o, new = 0, {}
        for k, v in eng.get_params().items():
            new[k] = flat[o:o + v.size].cpu().numpy().reshape(v.shape); o += v.size
        eng.set_params(new)
    # synthetic inputs of SURVEY 8d, several distinct batches resident in HBM
    NB = 4
    g = torch.Generator(device="cuda"); g.manual_seed(rank)
    imgs = [torch.rand((B, 128, 128, 1), device="cuda", generator=g) for _ in range(NB)]
    procs = [torch.randn((B, 5), device="cuda", generator=g) for _ in range(NB)]
    ys = [1000.0 + 100.0 * torch.randn((B, 1), device="cuda", generator=g) for _ in range(NB)]
    h_img = [t.cpu().pin_memory() for t in imgs]; h_proc = [t.cpu().pin_memory() for t in procs]; h_y = [t.cpu().pin_memory() for t in ys]
    flush = None if a.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device="cuda")      # > 126 MB L2
    stream = torch.cuda.ExternalStream(eng.stream())
    torch.cuda.synchronize()

    def step(i):
        j = i % NB
        if train:
            eng.train_step_dev(imgs[j], procs[j], ys[j], B, 1e-3)
        else:
            eng.forward_dev(imgs[j], procs[j], B, False)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(max(a.warmup, 3)):
        step(i)
    eng.synchronize()
    print("[bench] rank %d: warm-up done" % rank, file=sys.stderr, flush=True)
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start(); time.sleep(0.3)
    barrier()
    l0 = eng.kernel_launches()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    t_wall = time.time()
    for i in range(a.steps):
        if flush is not None:
            flush.zero_()                                   # on torch's stream; ordered before the step by the sync below
            torch.cuda.current_stream().synchronize()
        ev[i][0].record(stream); step(i); ev[i][1].record(stream)
    barrier()
    t_wall = time.time() - t_wall
    launches = eng.kernel_launches() - l0
    clocks = sampler.finish() if sampler else None
    step_ms = [s.elapsed_time(e) for s, e in ev]
    total_ms = torch.tensor([float(sum(step_ms))], device="cuda")
    if dist is not None:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    total_ms = float(total_ms.item())
    value = world * B * a.steps / (total_ms / 1e3)

    # ---- end to end through the host-buffer C-ABI call (H2D of the batch + D2H of loss/predictions inside the timed region)
    print("[bench] rank %d: device-timed region done" % rank, file=sys.stderr, flush=True)
    # One call of the public host-buffer API over `steps` batches, the way model.fit / model.predict walk a dataset
    # (models/CvT(Par).py:475, CvT_test(Par).py:522): every batch is copied from pinned host memory inside the timed region and
    # its loss / predictions are read back; the library double-buffers the copies against the steps.
    reps = (a.steps + NB - 1) // NB
    e_img = torch.cat(h_img * reps)[: a.steps * B].contiguous().pin_memory()
    e_proc = torch.cat(h_proc * reps)[: a.steps * B].contiguous().pin_memory()
    e_y = torch.cat(h_y * reps)[: a.steps * B].contiguous().pin_memory()
    if train:
        eng.train_epoch(e_img.numpy()[: 6 * B], e_proc.numpy()[: 6 * B], e_y.numpy()[: 6 * B], B, 0.0)      # untimed: graphs of both staging slots
    else:
        eng.predict(e_img.numpy()[: 6 * B], e_proc.numpy()[: 6 * B])
    barrier()
    t0 = time.time()
    if train:
        e_loss, e_mae = eng.train_epoch(e_img.numpy(), e_proc.numpy(), e_y.numpy(), B, 1e-3)
    else:
        e_pred = eng.predict(e_img.numpy(), e_proc.numpy())
    barrier()
    e2e_s = torch.tensor([time.time() - t0], device="cuda")
    print("[bench] rank %d: end-to-end region done" % rank, file=sys.stderr, flush=True)
    if dist is not None:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e = world * B * a.steps / float(e2e_s.item())
    h2d = B * (128 * 128 * 4 + 5 * 4 + (4 if train else 0)); d2h = 8 if train else B * 4

    if rank != 0:
        eng.close()
        if dist is not None:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant kernel group, timed live with CUDA events on the model's stream
    peaks = load_peaks()
    prof_runs = [eng.profile_step(imgs[0], procs[0], ys[0], B, train) for _ in range(3)][-1]
    agg = {}
    for name, ms in prof_runs:
        agg[name] = agg.get(name, 0.0) + ms
    from_geo = [dict(t=1024, tc=1024, d=64, heads=1, k=7, cin=1, hin=128, win=128), dict(t=256, tc=256, d=128, heads=2, k=3, cin=64, hin=32, win=32),
                dict(t=64, tc=65, d=256, heads=4, k=3, cin=128, hin=16, win=16)]
    work = group_work(from_geo, B)
    top = max(agg, key=agg.get)
    base = top.replace("bwd_", "")
    base = {"s1.out": "s1.out_gemm", "s2.out": "s2.out_gemm", "s3.out": "s3.out_gemm", "s1.qkv": "s1.qkv_gemm", "s2.qkv": "s2.qkv_gemm", "s3.qkv": "s3.qkv_gemm",
            "s1.proj_ln": "s1.ln_proj_bn", "s2.proj_ln": "s2.ln_proj_bn", "s3.proj_ln": "s3.ln_proj_bn", "s1.conv": "s1.conv_embed", "s2.conv": "s2.conv_embed",
            "s3.conv": "s3.conv_embed"}.get(base, base)
    mult = 2.0 if "bwd_" in top else 1.0
    roof = None
    if base in work:
        fl, by = work[base][0] * mult, work[base][1] * mult
        t = agg[top] / 1e3
        tf32_peak = peaks["bf16"] / 2.0                      # TF32 dense peak = half the measured bf16 cuBLAS figure
        tensor_bound = (fl / (tf32_peak * 1e12)) > (by / (peaks["hbm"] * 1e9)) and a.precision != "fp32"
        if tensor_bound:
            roof = {"bound": "tensor", "achieved": fl / t / 1e12, "peak": tf32_peak, "unit": "TFLOP/s", "frac": fl / t / 1e12 / tf32_peak, "traffic": None}
        else:
            roof = {"bound": "hbm", "achieved": by / t / 1e9, "peak": peaks["hbm"], "unit": "GB/s", "frac": by / t / 1e9 / peaks["hbm"], "traffic": None}
        try:      # DRAM traffic of this kernel group from the committed ncu --set full capture (profiles/ncu_traffic.json), else null
            roof["traffic"] = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(top)
        except Exception:
            pass
        roof.update({"kernel_group": top, "group_ms": agg[top], "peak_source": peaks["src"], "algorithmic_flops": fl, "algorithmic_bytes": by})
    line = {"metric": METRIC if train else "cvt_par_infer_images_per_sec", "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": max(a.warmup, 3), "ms_per_step": total_ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": {"fp32": "f32", "tf32": "tf32", "bf16": "bf16"}[a.precision], "data": "synthetic",
            "config": {"workload": "CvT(Par) %s step, batch %d per GPU, 128x128x1 + 5 process parameters (configs[1]), dw_bn, cls on stage 3" % (a.mode, B),
                       "global_batch": B * world, "parallelism": "dp%d" % world, "l2": "flushed between timed steps" if flush is not None else "not flushed",
                       "weights": "random-init (Keras initialisers)", "dropout": "philox 0.1" if train else "off"},
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "wall_s_timed_region": t_wall,
            "groups_ms": {k: round(v, 4) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:12]}}
    if world == 1 and not a.no_cpu_baseline:
        r = cpu_reference_run(3, 1, B, a.mode, budget_s=15.0)
        line["cpu_baseline"] = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]}
    sys.stdout.flush()
    os.write(json_fd, (json.dumps(line) + "\n").encode())
    eng.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        main_reference(args)
    else:
        main_b200(args)
