"""Helpers shared by the GPU parity tests: build an engine + the oracle on the same weights and inputs and compare
every named activation and every gradient, reporting all mismatches at once."""
import numpy as np

import transformer_stm_b200 as T
from oracle import cvt_oracle as O

TAPS = ["embed", "x0", "xn", "conv_qkv", "qkv1", "qkv2", "attn_o", "a1", "x1", "y", "hpre", "hact", "x2", "out"]


def make_pair(method="dw_bn", cls=True, proc_dim=5, nc=1, hw=(32, 32), dims=(64, 128, 256), heads=(1, 2, 4),
              batch=3, precision="fp32", seed=1234, drop_rate=0.1, max_batch=None):
    spec = O.ref_spec(method, cls, proc_dim, nc, image_h=hw[0], image_w=hw[1], dims=dims, heads=heads)
    spec.drop_rate = drop_rate
    cfg = T.CvtConfig.reference(method, cls, proc_dim, nc, image_h=hw[0], image_w=hw[1], dims=dims, heads=heads,
                                precision=precision, max_batch=max_batch or batch, drop_rate=drop_rate)
    eng = T.CvtEngine(cfg)
    params = O.init_params(spec, seed, "parity")
    eng.set_params(params)
    img, proc, y = O.synthetic_batch(spec, batch, seed=7)
    return spec, eng, params, img, proc, y


def rel_err(a, b, floor=1e-30):
    a = np.asarray(a, dtype=np.float64).reshape(-1)
    b = np.asarray(b, dtype=np.float64).reshape(-1)
    assert a.shape == b.shape, (a.shape, b.shape)
    den = max(float(np.abs(b).max()), floor)
    return float(np.abs(a - b).max()) / den


def compare_taps(eng, taps, n_stages=3, tol=1e-4, names=TAPS):
    bad, report = [], []
    for i in range(1, n_stages + 1):
        for nm in names:
            key = "s%d.%s" % (i, nm)
            if key not in taps:
                continue
            e = rel_err(eng.tap(key), taps[key])
            report.append((key, e))
            if not (e <= tol):
                bad.append((key, e))
    for key in ("head.feat", "pred"):
        e = rel_err(eng.tap(key), taps[key])
        report.append((key, e))
        if not (e <= tol):
            bad.append((key, e))
    return bad, report


def compare_grads(eng, grads, tol=1e-3, floor_frac=1e-4):
    """Per tensor max|a-b| / max(max|b|, floor).  The floor (1e-4 of the largest gradient in the model) keeps
    analytically-zero gradients (key biases: softmax is shift-invariant) from dividing noise by noise; the TF32 tests raise
    it to 1e-3, the size of one TF32 rounding, because there the noise on such a tensor is rounding noise of the whole dK sum."""
    got = eng.get_grads()
    bad, report = [], []
    floor = floor_frac * max(float(np.abs(g).max()) for g in grads.values())
    for k, g in grads.items():
        e = rel_err(got[k], g, floor)
        report.append((k, e))
        if not (e <= tol):
            bad.append((k, e))
    return bad, report


def fmt(report):
    return "\n".join("  %-60s %.3e" % r for r in report)
