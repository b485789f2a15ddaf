"""Generates tests/golden/refsrc_*.npz: outputs of THE REFERENCE'S OWN MODEL SOURCE (models/CvT(Par).py, CvT(Img).py, tools/grad_cam_CvT.py,
tools/model_plot.py under /root/reference, compiled unmodified) executed over oracle/mini_keras.py in float64 - predictions in inference mode,
and loss + gradients of one Keras train step (batch-statistics BatchNorm, dropout with the committed keep masks, MSE).  The reference tree does
not travel to the GPU box; these vectors do.  tests/test_refsrc_golden.py checks oracle/cvt_oracle.py against them.
    python tests/golden/make_refsrc_golden.py          (in the build container, where /root/reference exists)"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import cvt_oracle as O          # noqa: E402  (parameter names / shapes, seeded weights and inputs, dropout keep masks only)
from oracle import mini_keras as MK         # noqa: E402

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))

# name: (script, spec builder, projection_method, cls_token_switch, image (h, w), proc_dim, num_classes, batch)
CASES = {
    "par_dwbn_cls": ("models/CvT(Par).py", "ref", "dw_bn", True, (32, 32), 5, 1, 4),
    "par_dwbn_nocls_ragged": ("models/CvT(Par).py", "ref", "dw_bn", False, (36, 44), 5, 2, 3),
    "par_avg_cls": ("models/CvT(Par).py", "ref", "avg", True, (32, 32), 5, 1, 3),
    "par_linear_nocls": ("models/CvT(Par).py", "ref", "linear", False, (32, 32), 5, 1, 3),
    "img_dwbn_cls": ("models/CvT(Img).py", "ref", "dw_bn", True, (32, 32), 0, 1, 3),
    "gradcam_dwbn_nocls": ("tools/grad_cam_CvT.py", "gradcam", "dw_bn", False, (32, 32), 5, 1, 3),
    "gradcam_linear_cls": ("tools/grad_cam_CvT.py", "gradcam", "linear", True, (32, 32), 5, 1, 3),
    "modelplot_linear_cls": ("tools/model_plot.py", "model_plot", "linear", True, (32, 32), 5, 1, 3),
}
SEED = 2024


def build_spec(kind, method, cls, hw, proc_dim, nc):
    if kind == "gradcam":
        return O.gradcam_spec(cls, method, proc_dim=proc_dim, num_classes=nc, image_h=hw[0], image_w=hw[1])
    spec = O.ref_spec(method, cls, proc_dim=proc_dim, num_classes=nc, image_h=hw[0], image_w=hw[1])
    if kind == "model_plot":
        spec.head = "cls" if cls else "mean"
    return spec


def run_source(script, spec, params, img, proc, method, cls, training, masks):
    inputs = {"Image_inputs": torch.tensor(img, dtype=torch.float64)}
    args = [spec.image_h, spec.image_w, spec.in_chans]
    if spec.proc_dim > 0:
        inputs["Proc_inputs"] = torch.tensor(proc, dtype=torch.float64)
        args.append(spec.proc_dim)
    args.append(spec.num_classes)
    w = []
    for name, _, trainable in O.param_specs(spec):
        t = torch.tensor(np.asarray(params[name]), dtype=torch.float64)
        w.append((name, t.requires_grad_(True) if (training and trainable) else t))
    q = None
    if masks is not None:
        q = [torch.tensor(masks["s%d.%s" % (i, nm)], dtype=torch.float64) for i in range(1, len(spec.stages) + 1)
             for nm in ("attn", "proj", "mlp1", "mlp2") if "s%d.%s" % (i, nm) in masks]
    out, _ = MK.run_reference_model(os.path.join(REF, script), dict(projection_method=method, cls_token_switch=cls), args, w, inputs,
                                    training=training, dropout_masks=q)
    return out, dict(w)


def main():
    for name, (script, kind, method, cls, hw, proc_dim, nc, batch) in CASES.items():
        spec = build_spec(kind, method, cls, hw, proc_dim, nc)
        params = O.init_params(spec, SEED, "parity")
        img, proc, y = O.synthetic_batch(spec, batch, SEED % 97)
        masks = O.make_dropout_masks(spec, batch, seed=SEED, step=0)
        pred, _ = run_source(script, spec, params, img, proc, method, cls, False, None)
        out, w = run_source(script, spec, params, img, proc, method, cls, True, masks)
        loss = ((out - torch.tensor(y, dtype=torch.float64).reshape(out.shape)) ** 2).mean()
        loss.backward()
        data = {"pred": pred.numpy(), "train_pred": out.detach().numpy(), "loss": np.float64(loss.detach())}
        for pname, t in w.items():                           # every gradient tensor as (sum, sum of squares, first 16 entries): small files, every tensor pinned
            if t.requires_grad:
                g = (t.grad if t.grad is not None else torch.zeros_like(t)).numpy().ravel()
                data["g/" + pname] = np.concatenate([[g.sum(), (g * g).sum()], g[:16]])
        data["meta"] = np.array([script, kind, method, str(cls), str(hw[0]), str(hw[1]), str(proc_dim), str(nc), str(batch), str(SEED)])
        path = os.path.join(HERE, "refsrc_%s.npz" % name)
        np.savez_compressed(path, **data)
        print("%-28s pred %s loss %.6f  %d gradient tensors  %d bytes" % (name, data["pred"].ravel()[:2], data["loss"], sum(k.startswith("g/") for k in data), os.path.getsize(path)))


if __name__ == "__main__":
    main()
