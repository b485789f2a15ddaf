#!/usr/bin/env python
"""Per-tensor error of the CUDA path against the fp64 oracle for one train step (no dropout), fp32 and TF32 modes side by side: where the
TF32 mode's error lives.  python tests/scripts/parity_report.py [--batch 8] [--init parity|keras] [--hw 128]"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--init", default="parity")
    ap.add_argument("--hw", type=int, default=128)
    a = ap.parse_args()
    import torch
    import transformer_stm_b200 as T
    from oracle import cvt_oracle as O          # checker only
    from tests.parity_utils import TAPS
    spec = O.ref_spec(image_h=a.hw, image_w=a.hw); spec.drop_rate = 0.0
    params = O.init_params(spec, 1234, a.init)
    img, proc, y = O.synthetic_batch(spec, a.batch, 7)
    r = O.loss_and_grads(spec, params, img, proc, y, None)
    gmax = max(float(np.abs(v).max()) for v in r["grads"].values())
    for prec in ("fp32", "tf32"):
        eng = T.CvtEngine(T.CvtConfig.reference(image_h=a.hw, image_w=a.hw, precision=prec, max_batch=a.batch, drop_rate=0.0))
        eng.set_params(params); eng.set_dropout("off")
        d = [torch.tensor(np.ascontiguousarray(x, dtype=np.float32), device="cuda") for x in (img, proc, y)]
        torch.cuda.synchronize()
        eng.forward_dev(d[0], d[1], a.batch, True)
        eng.loss_backward_dev(d[2])
        g = eng.get_grads()
        print("== %s, batch %d, %s init" % (prec, a.batch, a.init))
        for i in (1, 2, 3):
            row = []
            for t in TAPS:
                k = "s%d.%s" % (i, t)
                if k not in r["taps"]:
                    continue
                try:
                    got = eng.tap(k)
                except Exception:
                    continue
                if t == "hpre" and prec == "tf32":
                    continue
                ref = np.asarray(r["taps"][k], dtype=np.float64).reshape(-1)
                row.append("%s %.1e" % (t, np.abs(got - ref).max() / np.abs(ref).max()))
            print("  s%d: " % i + "  ".join(row))
        pred = eng.tap("pred"); ref = np.asarray(r["taps"]["pred"], dtype=np.float64).reshape(-1)
        per = {k: float(np.abs(g[k] - v).max()) / max(float(np.abs(v).max()), 1e-3 * gmax) for k, v in r["grads"].items()}
        num = sum(float(((g[k].astype(np.float64) - v) ** 2).sum()) for k, v in r["grads"].items()); den = sum(float((np.asarray(v) ** 2).sum()) for v in r["grads"].values())
        vals = sorted(per.values())
        print("  pred %.2e   grad worst %.2e  median %.2e  L2 %.2e  (>1e-3: %d/%d)" % (np.abs(pred - ref).max() / np.abs(ref).max(), vals[-1], vals[len(vals) // 2], (num / den) ** 0.5, sum(v > 1e-3 for v in vals), len(vals)))
        for k in sorted(per, key=per.get, reverse=True)[:12]:
            print("     %-62s %.2e" % (k, per[k]))
        eng.close()


if __name__ == "__main__":
    main()
