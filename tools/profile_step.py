"""Runs a few CvT(Par) train (or inference) steps on cuda:0 with nothing else in the process: the command ncu wraps.
   python tools/profile_step.py [--precision tf32] [--batch 64] [--steps 4] [--mode train|infer]"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import transformer_stm_b200 as T  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--precision", default="tf32")
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--steps", type=int, default=4)
ap.add_argument("--mode", default="train")
ap.add_argument("--graphs", type=int, default=0)
a = ap.parse_args()
B = a.batch
eng = T.CvtEngine(T.CvtConfig.reference(precision=a.precision, max_batch=B))
eng.init_params(1234)
eng.set_dropout("philox" if a.mode == "train" else "off", 1234)
if a.graphs:
    eng.set_graphs(True)
g = torch.Generator(device="cuda"); g.manual_seed(0)
img = torch.rand((B, 128, 128, 1), device="cuda", generator=g)
proc = torch.randn((B, 5), device="cuda", generator=g)
y = 1000.0 + 100.0 * torch.randn((B, 1), device="cuda", generator=g)
torch.cuda.synchronize()
for i in range(a.steps):
    if a.mode == "train":
        eng.train_step_dev(img, proc, y, B, 1e-3)
    else:
        eng.forward_dev(img, proc, B, False)
eng.synchronize()
print("launches", eng.kernel_launches(), "loss/mae", eng.read_metrics() if a.mode == "train" else None)
eng.close()
