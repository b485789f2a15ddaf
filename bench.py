#!/usr/bin/env python
"""bench.py — CvT(Par) training-step AND inference throughput (images/s) on N B200s, the metric BASELINE.json names.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one rank per GPU under torchrun)
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (TensorFlow if importable, else the oracle port)
  python bench.py --config cfg1|cfg2|cfg5 ...              # BASELINE.json configs[0] / [1] (default) / [4]-geometry in TF32
  python bench.py --sweep                                  # configs[3]: inference batch sweep 1..4096 (one JSON line per batch)

One "step" = one Keras train_step of create_cvt_model's network (models/CvT(Par).py:292-354,464-476): forward (training) + MSE +
backward + Adam + BN moving update on one batch of 64 synthetic 128x128x1 images + 5 process parameters per GPU (configs[1]; weak
scaling: per-GPU batch fixed).  The same run also times batch inference (model.predict, models/CvT_test(Par).py:522) and reports it
under "infer".  Prints ONE JSON line on rank 0 (one per batch with --sweep).
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

CONFIGS = {
    # name: (workload text, image (h, w), dims, heads, proc_dim, default batch, modes)
    "cfg1": ("CvT(Img) forward inference, batch 32, 128x128x1 (configs[0])", (128, 128), (64, 128, 256), (1, 2, 4), 0, 32, ("infer",)),
    "cfg2": ("CvT(Par) train step + batch inference, batch 64 per GPU, 128x128x1 + 5 process parameters (configs[1]), dw_bn, cls on stage 3",
             (128, 128), (64, 128, 256), (1, 2, 4), 5, 64, ("train", "infer")),
    "cfg3": ("CvT(Par)(many): the same network with a multi-target regression head (num_classes = 20; num_classes = 1 is cfg2), data-parallel training, "
             "batch 64 per GPU (configs[2]; run with --gpus 8 for the 8 x B200 line)", (128, 128), (64, 128, 256), (1, 2, 4), 5, 64, ("train", "infer")),
    "cfg5": ("widened CvT(Par) (configs[4] geometry: 340x345 input, dims 64/192/384, heads 1/3/6), TF32 (bf16 mode is not implemented)",
             (340, 345), (64, 192, 384), (1, 3, 6), 5, 8, ("train", "infer")),
}


NUM_CLASSES = {"cfg3": 20}                                # every other configuration: 1 (models/CvT(Par).py:388)


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

    def wait_first(self, timeout_s):
        t0 = time.time()
        while not self.samples and time.time() - t0 < timeout_s:
            time.sleep(0.02)

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


# ---- algorithmic work (SURVEY 8d conventions: activations read + written once at fp32, fused groups) --------------------------
def stage_geo(hw, dims, heads, cls_last=True):
    geo, (h, w), cin = [], hw, 1
    for i, (d, nh, k, s) in enumerate(zip(dims, heads, (7, 3, 3), (4, 2, 2))):
        ho, wo = -(-h // s), -(-w // s)
        t = ho * wo
        geo.append(dict(t=t, tc=t + (1 if (cls_last and i == len(dims) - 1) else 0), d=d, heads=nh, k=k, cin=cin, hin=h, win=w))
        h, w, cin = ho, wo, d
    return geo


def group_work(geo, batch, mlp_ratio=4):
    """{group: (flops, bytes)} per kernel group for a forward pass over `batch` images: ALGORITHMIC work of the perfectly fused group as
    SURVEY 8d's table counts it (activations read + written once at fp32: K1 input map + T*D, K2 4 T*D, K3 6, K4 4, K5 3, K6 2) - not the
    traffic of the kernels as they happen to be split, so that an unfused group cannot flatter its roofline fraction."""
    out = {}
    for i, g in enumerate(geo, 1):
        T, Tc, D, H = g["t"], g["tc"], g["d"], g["heads"]
        kk = g["k"] * g["k"] * g["cin"]
        B = batch
        e = 4.0 * B * Tc * D
        out["s%d.conv_embed" % i] = (2.0 * B * T * kk * D, 4.0 * B * (g["hin"] * g["win"] * g["cin"] + T * D))
        out["s%d.ln_proj_bn" % i] = (B * Tc * D * (8 + 3 * 22.0), 4 * e)
        out["s%d.qkv_gemm" % i] = (2.0 * B * Tc * D * D * 6, 6 * e)
        out["s%d.attention" % i] = (4.0 * B * H * Tc * Tc * 64, 4 * e)
        out["s%d.out_gemm" % i] = (2.0 * B * Tc * D * D * 2, 3 * e)
        out["s%d.mlp" % i] = (2.0 * B * Tc * D * D * mlp_ratio * 2 + 8.0 * B * Tc * D, 2 * e)
    return out


def step_algorithmic(geo, batch, n_params, train, mlp_ratio=4):
    """Whole-step algorithmic (flops, bytes) with perfectly FUSED groups, SURVEY 8d table: per image forward K1 reads the input map and
    writes T*D; K2 5 T*D elements; K3 6; K4 4; K5 3; K6 2.  Training = 3x forward (+ 3 x parameter bytes for the gradient write and Adam)."""
    fl = by = 0.0
    for g in geo:
        T, Tc, D, H = g["t"], g["tc"], g["d"], g["heads"]
        kk = g["k"] * g["k"] * g["cin"]
        fl += 2.0 * T * kk * D + Tc * D * (8 + 66.0) + 2.0 * Tc * D * D * 8 + 4.0 * H * Tc * Tc * 64 + 2.0 * Tc * D * D * mlp_ratio * 2
        by += 4.0 * (g["hin"] * g["win"] * g["cin"] + T * D) + 4.0 * Tc * D * (4 + 6 + 4 + 3 + 2)
    fl *= batch; by *= batch
    if train:
        fl *= 3.0; by = 3.0 * by + 3.0 * 4.0 * n_params
    return fl, by


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--precision", default=os.environ.get("CVT_PRECISION", "tf32"), choices=["fp32", "tf32", "tf32x3"])
    ap.add_argument("--config", default="cfg2", choices=sorted(CONFIGS))
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (default: the configuration's own)")
    ap.add_argument("--mode", default="both", choices=["both", "train", "infer"])
    ap.add_argument("--sweep", action="store_true", help="configs[3]: inference at batch 1, 2, 4, ..., 4096 per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-x3", action="store_true", help="skip the extra tf32x3 (error-compensated mode) measurement of the default cfg2 line")
    ap.add_argument("--no-batch128", action="store_true", help="skip the extra batch-128 measurement of the default cfg2 line")
    ap.add_argument("--no-flush", action="store_true", help="do not flush L2 between timed steps")
    return ap.parse_args()


# ---- the reference's CPU path -----------------------------------------------------------------------------------------
def cpu_reference_run(steps, warmup, batch, mode, cfgname="cfg2", budget_s=120.0):
    """Times the reference's own CPU implementation of the path: the unmodified TensorFlow model when `import tensorflow` works
    (baseline/run_tf_reference.py, kind "tf"), else the torch-CPU port of its graph (oracle/, kind "port"), on all host cores."""
    cores = os.cpu_count() or 1
    _, hw, dims, heads, proc_dim, _, _ = CONFIGS[cfgname]
    try:
        sys.path.insert(0, os.path.join(ROOT, "baseline"))
        import run_tf_reference as RT
        if RT.tensorflow_available():
            r = RT.time_reference(mode, batch, steps, warmup, proc_dim=proc_dim, image_hw=hw, dims=dims, heads=heads)
            r.update(cores=cores, kind="tf")
            return r
    except Exception as e:                                           # TensorFlow absent (expected here): the oracle port
        print("[bench] TensorFlow reference unavailable (%s): timing the oracle port" % (str(e).splitlines()[0] if str(e) else type(e).__name__), file=sys.stderr)
    import torch
    from oracle import cvt_oracle as O
    torch.set_num_threads(cores)
    spec = O.ref_spec(proc_dim=proc_dim, image_h=hw[0], image_w=hw[1], dims=dims, heads=heads, num_classes=NUM_CLASSES.get(cfgname, 1))
    params = O.init_params(spec, 1234, "keras")
    img, proc, y = O.synthetic_batch(spec, batch, 0)
    st = O.new_opt_state(params)

    def one():
        if mode == "train":
            O.train_step(spec, params, st, img, proc, y, 1e-3, None, dtype=torch.float32)
        else:
            O.forward(spec, params, img, proc, False, None, dtype=torch.float32)

    times, t_all = [], time.time()
    for _ in range(max(warmup, 1)):
        one()
    for _ in range(max(steps, 1)):
        t0 = time.time(); one(); times.append(time.time() - t0)
        if time.time() - t_all > budget_s and len(times) >= 5:
            break
    ms = 1e3 * float(np.mean(times))
    return dict(value=batch / (ms / 1e3), ms_per_step=ms, cores=cores, steps=len(times), kind="port",
                sample="%d %s step(s) of batch %d after %d warm-up(s), fp32 torch-CPU oracle port, %d threads" % (len(times), mode, batch, max(warmup, 1), cores))


def main_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl, hw, dims, heads, proc_dim, dbatch, modes = CONFIGS[a.config]
    batch = a.batch or dbatch
    mode = a.mode if a.mode != "both" else modes[0]
    r = cpu_reference_run(a.steps, a.warmup, batch, mode, a.config)
    line = {"impl": "reference", "metric": METRIC if mode == "train" else "cvt_par_infer_images_per_sec", "value": r["value"], "unit": UNIT,
            "n_gpus": a.gpus, "steps": r["steps"], "warmup": max(a.warmup, 1), "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl, "global_batch": batch, "timed": mode},
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    if r["kind"] == "port":
        line["note"] = "TensorFlow/Keras unavailable on this box: the reference's CPU path is timed through the oracle port of its graph"
    if mode == "train" and "infer" in modes and a.mode == "both":
        ri = cpu_reference_run(max(3, a.steps // 2), 2, batch, "infer", a.config)
        line["infer"] = {"value": ri["value"], "unit": UNIT, "ms_per_step": ri["ms_per_step"], "sample": ri["sample"]}
    print(json.dumps(line))


# ---- this repo's arm ---------------------------------------------------------------------------------------------------
class Bench:
    def __init__(self, a, cfgname, batch):
        import torch
        import transformer_stm_b200 as T
        from transformer_stm_b200 import dp
        self.torch, self.T, self.dp, self.a = torch, T, dp, a
        self.rank = int(os.environ.get("RANK", "0")); self.world = int(os.environ.get("WORLD_SIZE", "1")); self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.wl, self.hw, self.dims, self.heads, self.proc_dim, _, self.modes = CONFIGS[cfgname]
        self.cfgname, self.B = cfgname, batch
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            if not dist.is_initialized():
                dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist
        self.nc = NUM_CLASSES.get(cfgname, 1)
        self.cfg = T.CvtConfig.reference(precision=a.precision, max_batch=batch, device=self.local, proc_dim=self.proc_dim, num_classes=self.nc,
                                         image_h=self.hw[0], image_w=self.hw[1], dims=self.dims, heads=self.heads)
        self.eng = T.CvtEngine(self.cfg)
        self.eng.init_params(1234)                                   # Keras initialisers, random-init weights
        if self.world > 1:
            dp.init_data_parallel(self.eng, self.dist, self.rank, self.world, torch.device("cuda", self.local))
        self.n_params = sum(p["numel"] for p in self.eng.param_infos() if p["trainable"])
        self.geo = stage_geo(self.hw, self.dims, self.heads)
        NB = 4
        g = torch.Generator(device="cuda"); g.manual_seed(self.rank)
        H, W = self.hw
        self.imgs = [torch.rand((batch, H, W, 1), device="cuda", generator=g) for _ in range(NB)]
        self.procs = [torch.randn((batch, max(self.proc_dim, 1)), device="cuda", generator=g) for _ in range(NB)] if self.proc_dim else [None] * NB
        self.ys = [1000.0 + 100.0 * torch.randn((batch, self.nc), device="cuda", generator=g) for _ in range(NB)]
        self.NB = NB
        self.flush = None if a.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device="cuda")      # > 126 MB L2
        self.stream = torch.cuda.ExternalStream(self.eng.stream())
        self.peaks = load_peaks()
        torch.cuda.synchronize()

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def maxr(self, v):
        t = self.torch.tensor([float(v)], device="cuda")
        if self.dist is not None:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def timed(self, train, steps, warmup, sampler=None):
        torch, eng, B = self.torch, self.eng, self.B
        eng.set_dropout("philox" if train else "off", seed=1234 + self.rank)

        def step(i):
            j = i % self.NB
            if train:
                eng.train_step_dev(self.imgs[j], self.procs[j], self.ys[j], B, 1e-3)
            else:
                eng.forward_dev(self.imgs[j], self.procs[j], B, False)

        # untimed: every one of the NB input-buffer sets is seen twice first (the library runs a (buffers, batch) key eagerly the first time and
        # captures its CUDA graph the second time - no capture may land inside the timed region), then the W warm-up steps proper
        for i in range(2 * self.NB):
            step(i)
        eng.synchronize()
        if sampler:
            sampler.start()
            sampler.wait_first(5.0)                                  # nvidia-smi's own start-up (driver locks) stays out of the timed region
        # (No rank-dependent number of extra steps here: a data-parallel train step is collective.  A loop that kept rank 0 - the only rank with a
        #  sampler - stepping until it had three clock samples let it run ahead at 8 GPUs: its peer-memory tail kernel trapped after the bounded
        #  spin and the other ranks hung.  tests/test_bench_protocol.py pins the step counts.)
        for i in range(max(warmup, 3)):                              # the W warm-up steps run last, so the timed region starts on a busy, clocked-up GPU
            step(i)
        self.barrier()
        l0 = eng.kernel_launches()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        t_wall = time.time()
        for i in range(steps):
            if self.flush is not None:
                # the previous step has to be over before the flush starts: the host runs ahead of the step stream, and without this wait every
                # flush of the loop executed during the first timed step (measured: step 0 at 2.1-2.9 ms, the rest unflushed at 1.37 ms)
                eng.synchronize()
                self.flush.zero_()                                   # on torch's stream; ordered before the step by the sync below
                torch.cuda.current_stream().synchronize()
            ev[i][0].record(self.stream); step(i); ev[i][1].record(self.stream)
        self.barrier()
        t_wall = time.time() - t_wall
        launches = eng.kernel_launches() - l0
        per = [s.elapsed_time(e) for s, e in ev]
        total_ms = self.maxr(sum(per))
        return dict(value=self.world * B * steps / (total_ms / 1e3), ms_per_step=total_ms / steps, launches=int(launches), wall_s=t_wall,
                    per_step_ms={"min": min(per), "median": float(np.median(per)), "max": max(per), "argmax": int(np.argmax(per)), "all": [round(v, 4) for v in per]})

    def e2e(self, train, steps):
        """End to end through the Keras-facing API with ordinary numpy arrays, as the reference scripts hold them (float64 [N,H,W], pageable):
        model.fit(epochs=1, shuffle=True) / model.predict over steps*B samples.  Every batch is gathered from the host arrays, copied to the
        device and its loss / predictions read back inside the timed region.  Also timed with float32 inputs."""
        from transformer_stm_b200 import keras_api as K
        B, H, W = self.B, self.hw[0], self.hw[1]
        # a wall-clock region of a few dozen ms is at the mercy of one host hiccup: at least 64 batches per epoch, 3 timed epochs, median reported
        steps = max(2, min(max(steps, 64), (1 << 30) // (B * H * W * 8)))      # at most 1 GiB of float64 images on the host (batch sweep)
        n = steps * B
        rng = np.random.default_rng(100 + self.rank)
        x64 = rng.random((n, H, W)); p64 = rng.standard_normal((n, max(self.proc_dim, 1))); y64 = 1000.0 + 100.0 * rng.standard_normal(n if self.nc == 1 else (n, self.nc))
        model = K.Model(self.cfg, engine_factory=lambda cfg: self.eng, seed=5)
        model.compile(optimizer=K.Adam(1e-3), loss="mean_squared_error", metrics=["mae"])
        out = {}
        for tag, (x, p, y) in (("f64", (x64, p64, y64)), ("f32", (x64.astype(np.float32), p64.astype(np.float32), y64.astype(np.float32)))):
            xin = [x, p] if self.proc_dim else x
            warm = [x[: 6 * B], p[: 6 * B]] if self.proc_dim else x[: 6 * B]
            if train:
                model.fit(warm, y[: 6 * B], epochs=1, batch_size=B, verbose=0)          # untimed: graphs of both staging slots, ring allocation
            else:
                model.predict(warm)
            reps = []
            for _ in range(3):
                self.barrier()
                t0 = time.time()
                if train:
                    model.fit(xin, y, epochs=1, batch_size=B, shuffle=True, verbose=0)
                else:
                    model.predict(xin)
                self.barrier()
                reps.append(self.world * n / self.maxr(time.time() - t0))
            out[tag] = float(np.median(reps))
            out[tag + "_reps"] = [round(v, 1) for v in reps]
        elem = 8
        h2d = B * (H * W * elem + self.proc_dim * 4 + (4 * self.nc if train else 0)); d2h = 8 if train else B * 4 * self.nc
        return {"value": out["f64"], "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "api": "keras_api.Model.fit(epochs=1, shuffle=True)" if train else "keras_api.Model.predict", "inputs": "float64 numpy, pageable",
                "samples_per_epoch": n, "protocol": "3 timed calls of %d batches each, median; all three in `repeats`" % steps, "repeats": out["f64_reps"],
                "value_float32_inputs": out["f32"]}

    def roofline(self, train, ms_per_step):
        eng, B, peaks = self.eng, self.B, self.peaks
        prof = [eng.profile_step(self.imgs[0], self.procs[0], self.ys[0], B, train) for _ in range(3)][-1]
        agg = {}
        for name, ms in prof:
            agg[name] = agg.get(name, 0.0) + ms
        work = group_work(self.geo, B)
        top = max(agg, key=agg.get)
        base = top.replace("bwd_", "")
        for short, full in (("out", "out_gemm"), ("qkv", "qkv_gemm"), ("proj_ln", "ln_proj_bn"), ("conv", "conv_embed")):
            if base.endswith("." + short):
                base = base[: -len(short)] + full
        mult = 2.0 if "bwd_" in top else 1.0
        tf32_peak = peaks["bf16"] / 2.0                              # TF32 dense peak = half the measured bf16 cuBLAS figure
        roof = None
        if base in work:
            fl, by = work[base][0] * mult, work[base][1] * mult
            t = agg[top] / 1e3
            tensor_bound = (fl / (tf32_peak * 1e12)) > (by / (peaks["hbm"] * 1e9)) and self.a.precision != "fp32"
            if tensor_bound:
                roof = {"bound": "tensor", "achieved": fl / t / 1e12, "peak": tf32_peak, "unit": "TFLOP/s", "frac": fl / t / 1e12 / tf32_peak, "traffic": None}
            else:
                roof = {"bound": "hbm", "achieved": by / t / 1e9, "peak": peaks["hbm"], "unit": "GB/s", "frac": by / t / 1e9 / peaks["hbm"], "traffic": None}
            try:      # DRAM bytes of this group per launch from the committed ncu capture of the same kernels (tools/ncu_traffic.py)
                tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
                key = "train" if train else "infer"
                roof["traffic"] = tr.get(key, {}).get(top)
                roof["traffic_source"] = tr.get("source")
            except Exception:
                pass
            roof.update({"kernel_group": top, "group_ms": agg[top], "peak_source": peaks["src"], "algorithmic_flops": fl, "algorithmic_bytes": by})
        # whole step against the SURVEY 8d fused-kernel ceiling: this is the number that cannot flatter
        sfl, sby = step_algorithmic(self.geo, B, self.n_params, train)
        t = ms_per_step / 1e3
        if roof is None:
            roof = {"bound": "hbm", "achieved": None, "peak": peaks["hbm"], "unit": "GB/s", "frac": None, "traffic": None}
        roof.update({"step_frac": sby / t / 1e9 / peaks["hbm"], "step_algorithmic_bytes": sby, "step_algorithmic_flops": sfl,
                     "step_tensor_frac": sfl / t / 1e12 / tf32_peak})
        groups = {k: round(v, 4) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:12]}
        return roof, groups

    def parity(self):
        """First-step check of the benched configuration against the oracle (untimed, world 1): predictions and gradients of this batch."""
        from oracle import cvt_oracle as O          # checker only
        import torch
        eng, B = self.eng, self.B
        spec = O.ref_spec(proc_dim=self.proc_dim, image_h=self.hw[0], image_w=self.hw[1], dims=self.dims, heads=self.heads, num_classes=self.nc)
        params = {k: v.astype(np.float64) for k, v in eng.get_params().items()}
        img = self.imgs[0].cpu().numpy(); proc = self.procs[0].cpu().numpy() if self.proc_dim else None; y = self.ys[0].cpu().numpy()
        eng.set_dropout("off")
        eng.forward_dev(self.imgs[0], self.procs[0], B, True)
        eng.loss_backward_dev(self.ys[0])
        pred = eng.tap("pred")
        grads = eng.get_grads()
        r = O.loss_and_grads(spec, params, img, proc, y, None, dtype=torch.float64 if B * self.hw[0] <= 64 * 128 else torch.float32)
        gmax = max(float(np.abs(v).max()) for v in r["grads"].values())
        per = {k: float(np.abs(grads[k] - v).max()) / max(float(np.abs(v).max()), 1e-3 * gmax) for k, v in r["grads"].items()}
        worst = max(per, key=per.get)
        gerr = per[worst]
        num = sum(float(((grads[k].astype(np.float64) - v) ** 2).sum()) for k, v in r["grads"].items())
        den = sum(float((np.asarray(v, dtype=np.float64) ** 2).sum()) for v in r["grads"].values())
        perr = float(np.abs(pred.reshape(-1) - r["pred"].reshape(-1)).max() / np.abs(r["pred"]).max())
        over = sorted(per.values())
        return {"pred": perr, "grad": gerr, "grad_worst_tensor": worst, "grad_median_tensor": over[len(over) // 2], "grad_l2_all": (num / den) ** 0.5,
                "tensors_over_1e-3": int(sum(e > 1e-3 for e in over)), "tensors": len(over),
                "norm": "max|a-b| / max|b| per tensor (gradient floor 1e-3 of the largest gradient); grad_l2_all = ||a-b|| / ||b|| over all gradients",
                "contract": "north_star asks rtol 1e-3; the fp32 mode meets it, this TF32 tensor-core mode is measured here (DESIGN.md 4)",
                "oracle": "torch-CPU port; pinned to the reference's model source executed over oracle/mini_keras.py (tests/golden/refsrc_*.npz), unpinned against TensorFlow's kernels"}

    def close(self):
        self.eng.close()


def run_config(a, cfgname, batch, modes, first=True):
    import torch
    b = Bench(a, cfgname, batch)
    rank, world = b.rank, b.world
    res = {}
    sampler = ClockSampler(b.local) if rank == 0 else None
    par = None
    if world == 1 and not a.no_parity and "train" in modes and cfgname != "cfg5" and batch * b.hw[0] * b.hw[1] <= 64 * 128 * 128 * 2:
        par = b.parity()
    for mode in modes:
        train = mode == "train"
        r = b.timed(train, a.steps, a.warmup, sampler if mode == modes[0] else None)
        print("[bench] rank %d: %s device-timed region done (%.1f img/s)" % (rank, mode, r["value"]), file=sys.stderr, flush=True)
        if mode == modes[0] and sampler:
            r["clocks"] = sampler.finish()
        r["e2e"] = b.e2e(train, a.steps)
        print("[bench] rank %d: %s end-to-end region done" % (rank, mode), file=sys.stderr, flush=True)
        if rank == 0:
            r["roofline"], r["groups_ms"] = b.roofline(train, r["ms_per_step"])
        res[mode] = r
    dp_check = None
    if world > 1 and "train" in modes:
        dp_check = {"replica_divergence": b.dp.replica_divergence(b.eng, b.dist, world, torch.device("cuda", b.local)),
                    "meaning": "max |checksum(rank) - checksum(rank 0)| over 3 weight checksums after the timed steps; 0.0 = identical replicas"}
    if rank != 0:
        b.close()
        return None
    head = res[modes[0]]
    train_line = modes[0] == "train"
    line = {"metric": METRIC if train_line else "cvt_par_infer_images_per_sec", "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": max(a.warmup, 3), "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": {"fp32": "f32", "tf32": "tf32", "tf32x3": "tf32x3"}[a.precision], "data": "synthetic",
            "config": {"workload": b.wl, "name": cfgname, "per_gpu_batch": batch, "global_batch": batch * world, "parallelism": "dp%d" % world,
                       "l2": "flushed between timed steps" if b.flush is not None else "not flushed",
                       "weights": "random-init (Keras initialisers)", "dropout": "philox 0.1 (train), off (inference)"},
            "e2e": head["e2e"], "gpu_launches": head["launches"], "clocks": head.get("clocks"), "roofline": head["roofline"],
            "wall_s_timed_region": head["wall_s"], "per_step_ms": head["per_step_ms"], "groups_ms": head["groups_ms"]}
    if "infer" in res and train_line:
        i = res["infer"]
        line["infer"] = {"metric": "cvt_par_infer_images_per_sec", "value": i["value"], "unit": UNIT, "ms_per_step": i["ms_per_step"], "per_step_ms": i["per_step_ms"], "e2e": i["e2e"],
                         "gpu_launches": i["launches"], "roofline": i["roofline"], "groups_ms": i["groups_ms"]}
    if par is not None:
        line["parity"] = par
    if dp_check is not None:
        line["dp_check"] = dp_check
    if world == 1 and not a.no_cpu_baseline and first:
        r = cpu_reference_run(5, 2, batch, modes[0], cfgname, budget_s=25.0)
        line["cpu_baseline"] = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]}
    b.close()
    return line


def main_b200(a):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    # stdout carries exactly the JSON line(s): anything libraries print there (e.g. NCCL's version banner) goes to stderr instead
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this arm has no CPU fallback (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    wl, hw, dims, heads, proc_dim, dbatch, modes = CONFIGS[a.config]
    if a.mode != "both":
        modes = tuple(m for m in modes if m == a.mode) or modes
    if a.sweep:
        for i, bsz in enumerate([1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096]):
            line = run_config(a, a.config, bsz, ("infer",), first=False)
            if rank == 0 and line is not None:
                os.write(json_fd, (json.dumps(line) + "\n").encode())
    else:
        line = run_config(a, a.config, a.batch or dbatch, modes)
        if a.config == "cfg2" and not a.batch and a.mode == "both" and not a.no_batch128 and int(os.environ.get("WORLD_SIZE", "1")) == 1:
            # the reference's own training batch (models/CvT(Par).py:47: 128) beside the headline batch 64: device-timed only, same protocol
            b = Bench(a, a.config, 128)
            extra = {}
            for mode in ("train", "infer"):
                r = b.timed(mode == "train", a.steps, a.warmup)
                roof, _ = b.roofline(mode == "train", r["ms_per_step"]) if rank == 0 else (None, None)
                extra[mode] = {"value": r["value"], "unit": UNIT, "ms_per_step": r["ms_per_step"], "step_frac": roof["step_frac"] if roof else None}
            b.close()
            if line is not None:
                line["batch128"] = dict(extra, note="per-GPU batch 128 (the reference scripts' batch_size), value / ms_per_step timed like the headline")
        if (a.config == "cfg2" and not a.batch and a.mode == "both" and a.precision == "tf32" and not a.no_x3
                and int(os.environ.get("WORLD_SIZE", "1")) == 1):
            # the error-compensated tensor-core mode (CVT_PREC_TF32X3) beside the headline: the mode that MEETS north_star's rtol 1e-3, with its
            # own first-step parity check and device-timed throughput (same protocol, same batch)
            import copy
            a3 = copy.copy(a); a3.precision = "tf32x3"
            b = Bench(a3, a.config, dbatch)
            extra = {"parity": b.parity() if not a.no_parity else None}
            for mode in ("train", "infer"):
                r = b.timed(mode == "train", a.steps, a.warmup)
                extra[mode] = {"value": r["value"], "unit": UNIT, "ms_per_step": r["ms_per_step"], "gpu_launches": r["launches"]}
            b.close()
            if line is not None:
                line["tf32x3"] = dict(extra, note="precision tf32x3: every dense product as hi.hi + lo.hi + hi.lo on tcgen05 (operand tiles split in "
                                                  "shared memory), attention single-pass TF32; `bench.py --precision tf32x3` gives its full line")
        if rank == 0 and line is not None:
            os.write(json_fd, (json.dumps(line) + "\n").encode())
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        main_reference(args)
    else:
        main_b200(args)
