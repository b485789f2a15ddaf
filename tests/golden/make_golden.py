"""Generates tests/golden/cvt_golden_*.npz from the CPU oracle (fp64).

PARITY UNPINNED: TensorFlow/Keras (the library that holds the reference's arithmetic) cannot be imported in the build
container, so these vectors come from oracle/cvt_oracle.py, not from the reference itself.  They pin (a) the oracle against
accidental edits and (b) the CUDA path against a fixed, committed set of numbers.  Re-generate with
    python tests/golden/make_golden.py
Inputs and weights are NOT stored: they are re-created from seeds (oracle.init_params(seed, 'parity'), oracle.synthetic_batch).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import cvt_oracle as O  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

# name -> (projection, cls, proc_dim, num_classes, (H, W), batch)
CASES = {
    "par_dwbn_cls_32": ("dw_bn", True, 5, 1, (32, 32), 3),
    "img_dwbn_cls_32": ("dw_bn", True, 0, 1, (32, 32), 2),
    "par_avg_nocls_48x40": ("avg", False, 5, 2, (48, 40), 2),
    "par_linear_cls_32": ("linear", True, 5, 1, (32, 32), 2),
}
SEED, DATA_SEED = 1234, 7
KEEP_GRADS = ("stage1_ConvEmbed/conv2d/kernel", "stage3_transformer/cls_token", "stage2_transformer/layer_normalization/gamma",
              "Final_Dense/kernel")


def build(case):
    method, cls, pd, nc, hw, batch = CASES[case]
    spec = O.ref_spec(method, cls, pd, nc, image_h=hw[0], image_w=hw[1])
    spec.drop_rate = 0.0
    params = O.init_params(spec, SEED, "parity")
    img, proc, y = O.synthetic_batch(spec, batch, DATA_SEED)
    return spec, params, img, (proc if pd > 0 else None), y


def compute(case):
    spec, params, img, proc, y = build(case)
    pred_inf, taps = O.forward(spec, params, img, proc, training=False)
    r = O.loss_and_grads(spec, params, img, proc, y, None)
    out = {"pred_inference": pred_inf, "pred_training": r["pred"], "loss": np.float64(r["loss"]), "mae": np.float64(r["mae"]),
           "s3_out_inference": taps["s3.out"]}
    for k, g in r["grads"].items():
        out["gnorm/" + k] = np.float64(np.sqrt((g.astype(np.float64) ** 2).sum()))
        out["gsum/" + k] = np.float64(g.astype(np.float64).sum())
    for k in KEEP_GRADS:
        if k in r["grads"]:
            out["grad/" + k] = r["grads"][k]
    return out


if __name__ == "__main__":
    for case in CASES:
        out = compute(case)
        path = os.path.join(HERE, "cvt_golden_%s.npz" % case)
        np.savez_compressed(path, **out)
        print(case, "->", path, os.path.getsize(path), "bytes; loss", float(out["loss"]))
