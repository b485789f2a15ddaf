"""All kernel groups of one train (or inference) step with their CUDA-event times (cvt_profile_step, eager issue, warm L2):
   python tools/groups.py [--mode train|infer]"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import transformer_stm_b200 as T  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--mode", default="train")
ap.add_argument("--batch", type=int, default=64)
a = ap.parse_args()
B, train = a.batch, a.mode == "train"
eng = T.CvtEngine(T.CvtConfig.reference(precision="tf32", max_batch=B))
eng.init_params(1234)
eng.set_dropout("philox" if train else "off", 1234)
g = torch.Generator(device="cuda"); g.manual_seed(0)
img = torch.rand((B, 128, 128, 1), device="cuda", generator=g)
proc = torch.randn((B, 5), device="cuda", generator=g)
y = 1000.0 + 100.0 * torch.randn((B, 1), device="cuda", generator=g)
runs = [eng.profile_step(img, proc, y, B, train) for _ in range(5)][-1]
tot = sum(ms for _, ms in runs)
for name, ms in runs:
    print("%-22s %8.1f us %5.1f%%" % (name, ms * 1e3, 100 * ms / tot))
print("total %.1f us" % (tot * 1e3))
eng.close()
