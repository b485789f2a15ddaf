#!/usr/bin/env python
"""Runs the UNMODIFIED reference model on TensorFlow's CPU path — the only thing that can un-pin parity and the real
CPU baseline.  Inert in this image: TensorFlow / Keras 2 are not installed and cannot be (no network); `bench.py --impl reference`
and `tests/` call `tensorflow_available()` first and fall back to the oracle port (kind "port") when it is False.

What it does when TensorFlow 2.x with Keras 2 is importable and the reference tree is present (CVT_REFERENCE_ROOT, default
/root/reference):
  * exec()s lines 63-354 of models/CvT(Par).py (or models/CvT(Img).py:63-341 for proc_dim = 0) exactly as they are — the module-level
    `spec`, the four layer classes and create_cvt_model — into a namespace that provides the switches the script defines above them;
  * copies a canonical-name weight dict (the oracle's / the CUDA path's) into the Keras model BY POSITION, in the Keras layout
    keras_api.Model.keras_layout() documents (trainable first, BatchNorm moving statistics last);
  * dumps predictions, loss and every gradient for the oracle cross-check (`--dump out.npz`), and times model.predict(batch 32 by
    default) / model.train_on_batch with CUDA_VISIBLE_DEVICES="" and TensorFlow's default thread pools.
"""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("CVT_REFERENCE_ROOT", "/root/reference")


def tensorflow_available() -> bool:
    if not os.path.exists(os.path.join(REF, "models", "CvT(Par).py")):
        return False
    try:
        os.environ.setdefault("CUDA_VISIBLE_DEVICES", "")
        import tensorflow as tf  # noqa: F401
        return True
    except Exception:
        return False


def build_reference_model(proc_dim=5, num_classes=1, image_hw=(128, 128), projection_method="dw_bn", cls_token_switch=True,
                          dims=(64, 128, 256), heads=(1, 2, 4)):
    """exec the reference's own model-definition section; returns (keras model, namespace)."""
    import tensorflow as tf
    from tensorflow import keras
    from tensorflow.keras import layers
    script = "CvT(Par).py" if proc_dim > 0 else "CvT(Img).py"
    last = 354 if proc_dim > 0 else 341
    src = open(os.path.join(REF, "models", script), encoding="utf8").read().splitlines()
    body = "\n".join(src[62:last])                                 # lines 63..last, unmodified
    ns = {"tf": tf, "keras": keras, "layers": layers, "np": np, "projection_method": projection_method, "cls_token_switch": cls_token_switch,
          "__name__": "reference_model_section"}
    exec(compile(body, os.path.join(REF, "models", script), "exec"), ns)
    for st, d, h in zip(ns["spec"]["stages"], dims, heads):          # the widened configuration edits `spec` the way a user would
        st["embed_dim"], st["num_heads"] = d, h
    if proc_dim > 0:
        model = ns["create_cvt_model"](image_hw[0], image_hw[1], 1, proc_dim, num_classes)
    else:
        model = ns["create_cvt_model"](image_hw[0], image_hw[1], 1, num_classes)
    return model, ns


def set_weights_by_layout(model, layout, weights):
    """layout: keras_api.Model.keras_layout(); weights: canonical name -> array.  Positional within each layer, shapes checked."""
    by_name = {l.name: l for l in model.layers}
    for lname, ws in layout:
        if not ws:
            continue
        layer = by_name.get(lname)
        if layer is None and lname.startswith("layer_normalization"):
            cands = [l for l in model.layers if l.name.startswith("layer_normalization")]
            layer = cands[-1]
        vals = [np.asarray(weights[c], dtype=np.float32) for _, c in ws]
        have = layer.trainable_weights + layer.non_trainable_weights
        assert len(have) == len(vals), (lname, len(have), len(vals))
        for v, w in zip(vals, have):
            assert tuple(w.shape) == tuple(v.shape), (lname, w.name, w.shape, v.shape)
            w.assign(v)


def canonical_gradients(model, layout, grads_by_var):
    out = {}
    by_name = {l.name: l for l in model.layers}
    for lname, ws in layout:
        if not ws:
            continue
        layer = by_name.get(lname) or [l for l in model.layers if l.name.startswith("layer_normalization")][-1]
        for (kn, cn), var in zip(ws, layer.trainable_weights + layer.non_trainable_weights):
            g = grads_by_var.get(var.ref())
            if g is not None:
                out[cn] = np.asarray(g)
    return out


def run_and_dump(weights, layout, img, proc, y, path=None, **model_kw):
    """predictions (inference mode), loss / mae / gradients of one training-mode pass with dropout rate forced to 0 by the caller's spec."""
    import tensorflow as tf
    model, ns = build_reference_model(**model_kw)
    set_weights_by_layout(model, layout, weights)
    xs = [img, proc] if proc is not None else img
    pred = model(xs, training=False).numpy()
    with tf.GradientTape() as tape:
        out = model(xs, training=True)
        loss = tf.reduce_mean(tf.square(out - y.reshape(out.shape)))
    grads = tape.gradient(loss, model.trainable_variables)
    gv = {v.ref(): g for v, g in zip(model.trainable_variables, grads)}
    res = dict(pred=pred, loss=float(loss), **{"grad/" + k: v for k, v in canonical_gradients(model, layout, gv).items()})
    if path:
        np.savez_compressed(path, **res)
    return res


def time_reference(mode, batch, steps, warmup, proc_dim=5, image_hw=(128, 128), dims=(64, 128, 256), heads=(1, 2, 4)):
    import tensorflow as tf
    from tensorflow import keras
    model, ns = build_reference_model(proc_dim=proc_dim, image_hw=image_hw, dims=dims, heads=heads)
    model.compile(optimizer=keras.optimizers.Adam(learning_rate=1e-3), loss="mean_squared_error", metrics=["mae"])   # models/CvT(Par).py:464-466
    rng = np.random.default_rng(0)
    x = rng.random((batch,) + tuple(image_hw)).astype(np.float64)
    p = np.random.default_rng(1).standard_normal((batch, proc_dim)) if proc_dim > 0 else None
    y = 1000.0 + 100.0 * np.random.default_rng(2).standard_normal(batch)
    xs = [x, p] if p is not None else x

    def one():
        if mode == "train":
            model.train_on_batch(xs, y)
        else:
            model.predict(xs, batch_size=min(32, batch), verbose=0)                                                      # Keras default batch 32
    for _ in range(max(warmup, 1)):
        one()
    times = []
    for _ in range(max(steps, 1)):
        t0 = time.time(); one(); times.append(time.time() - t0)
    ms = 1e3 * float(np.mean(times))
    return dict(value=batch / (ms / 1e3), ms_per_step=ms, steps=len(times),
                sample="%d %s step(s) of batch %d after %d warm-up(s), unmodified reference model on TensorFlow %s CPU" % (len(times), mode, batch, max(warmup, 1), tf.__version__))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dump", help="write predictions / loss / gradients on the oracle's parity inputs to this .npz")
    ap.add_argument("--time", choices=["train", "infer"])
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=5)
    a = ap.parse_args()
    if not tensorflow_available():
        print("TensorFlow (or %s) unavailable on this box: nothing to run" % REF)
        return 0
    sys.path.insert(0, ROOT)
    if a.dump:
        from oracle import cvt_oracle as O
        from tests.fake_engine import OracleEngine
        import transformer_stm_b200 as T
        from transformer_stm_b200 import keras_api as K
        spec = O.ref_spec()
        spec.drop_rate = 0.0
        params = O.init_params(spec, 1234, "parity")
        img, proc, y = O.synthetic_batch(spec, 4, 7)
        layout = K.Model(T.CvtConfig.reference(), engine_factory=OracleEngine).keras_layout()
        res = run_and_dump(params, layout, img, proc, y, a.dump)
        ref = O.loss_and_grads(spec, params, img, proc, y, None)
        print("TF vs oracle: loss %.6g vs %.6g" % (res["loss"], ref["loss"]))
        for k, g in ref["grads"].items():
            e = float(np.abs(res["grad/" + k] - g).max() / (np.abs(g).max() + 1e-30))
            print("  %-60s %.3e" % (k, e))
    if a.time:
        print(time_reference(a.time, a.batch, a.steps, 2))
    return 0


if __name__ == "__main__":
    sys.exit(main())
