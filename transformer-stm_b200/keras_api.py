"""Keras-facing call surface of the reference's CvT scripts, backed by libcvt_b200 (no TensorFlow, no CPU fallback).

What the reference scripts call, and what stands in for it here (paths relative to the reference repo):

  create_cvt_model(image_height, image_width, num_channels, proc_dim, num_classes)   models/CvT(Par).py:292
  create_cvt_model(image_height, image_width, num_channels, num_classes)              models/CvT(Img).py:291
  model.compile(optimizer=Adam(learning_rate=1e-3), loss='mean_squared_error', metrics=['mae'])   :464-466
  model.fit([x, proc], y, epochs=, batch_size=, validation_data=([xv, pv], yv), callbacks=[...])  :475-476
  model.history.history  -> {'loss','mae','val_loss','val_mae','lr'}                              :492
  model.save_weights(path) / model.load_weights(path)                 :489, models/CvT_test(Par).py:513
  model.predict([x, proc]) -> float32 [N, num_classes] (default batch 32)   models/CvT_test(Par).py:522
  keras.callbacks.LearningRateScheduler(lr_scheduler), TensorBoard(...)     :469-476
  make_gradcam_heatmap(img_array, proc_array, model, last_conv_layer_name)  tools/grad_cam_CvT.py:422-481

The module-level switches the scripts edit (`spec`, `projection_method`, `cls_token_switch`; models/CvT(Par).py:25-72)
are module attributes here too; `configure()` sets them in one call.
"""
from __future__ import annotations

import math
import os
import time
from typing import Callable, Dict, List, Optional, Sequence

import numpy as np

from . import h5lite
from .engine import CvtConfig, CvtEngine

# ---- the switches the reference keeps as module-level constants (models/CvT(Par).py:25-28,66-72) ----------------
projection_method = "dw_bn"
cls_token_switch = True
# which script's create_cvt_model is mirrored: "reference" = models/CvT(Par).py:292-354 / CvT(Img).py; "gradcam" = tools/grad_cam_CvT.py:315-349
# (cls token on every stage, proj_q/k/v never applied, attention-output dropout, 'linear' = Dense, GlobalAveragePooling2D head);
# "model_plot" = tools/model_plot.py:254-300 (the reference network with the un-normalised head of :289-292)
model_variant = "reference"


def _make_spec():
    return {"stages": [
        {"embed_dim": 64, "patch_size": 7, "stride": 4, "num_heads": 1, "kernel_size": 3, "strides": 1,
         "qkv_method": projection_method, "with_cls_token": False},
        {"embed_dim": 128, "patch_size": 3, "stride": 2, "num_heads": 2, "kernel_size": 3, "strides": 1,
         "qkv_method": projection_method, "with_cls_token": False},
        {"embed_dim": 256, "patch_size": 3, "stride": 2, "num_heads": 4, "kernel_size": 3, "strides": 1,
         "qkv_method": projection_method, "with_cls_token": cls_token_switch},
    ]}


spec = _make_spec()


def configure(projection: Optional[str] = None, cls_token: Optional[bool] = None, new_spec: Optional[dict] = None,
              variant: Optional[str] = None):
    """Sets the module-level switches the reference scripts edit by hand, then rebuilds `spec` from them."""
    global projection_method, cls_token_switch, spec, model_variant
    if variant is not None:
        if variant not in ("reference", "gradcam", "model_plot"):
            raise ValueError("unknown model variant %r" % (variant,))
        model_variant = variant
    if projection is not None:
        projection_method = projection
    if cls_token is not None:
        cls_token_switch = bool(cls_token)
    spec = new_spec if new_spec is not None else _make_spec()
    return spec


# ---- keras.optimizers / keras.callbacks stand-ins -----------------------------------------------------------------
class Adam:
    """keras.optimizers.Adam(learning_rate=1e-3): beta 0.9/0.999, epsilon 1e-7 (the Keras default, not torch's 1e-8)."""

    def __init__(self, learning_rate: float = 1e-3, beta_1: float = 0.9, beta_2: float = 0.999, epsilon: float = 1e-7):
        if (beta_1, beta_2, epsilon) != (0.9, 0.999, 1e-7):
            raise ValueError("the CUDA Adam kernel implements the Keras defaults beta_1=0.9, beta_2=0.999, epsilon=1e-7 only")
        self.learning_rate = float(learning_rate)

    # keras spells it both ways
    @property
    def lr(self):
        return self.learning_rate

    @lr.setter
    def lr(self, v):
        self.learning_rate = float(v)


class Callback:
    model = None

    def set_model(self, model):
        self.model = model

    def on_train_begin(self, logs=None): pass
    def on_train_end(self, logs=None): pass
    def on_epoch_begin(self, epoch, logs=None): pass
    def on_epoch_end(self, epoch, logs=None): pass


class LearningRateScheduler(Callback):
    """keras.callbacks.LearningRateScheduler(schedule): lr <- schedule(epoch, lr) at the start of every epoch and
    logs['lr'] at its end (models/CvT(Par).py:357-360,469)."""

    def __init__(self, schedule: Callable[[int, float], float], verbose: int = 0):
        self.schedule, self.verbose = schedule, verbose

    def on_epoch_begin(self, epoch, logs=None):
        lr = float(self.model.optimizer.learning_rate)
        new = self.schedule(epoch, lr)
        if not isinstance(new, (float, np.floating, int)):
            raise ValueError('The output of the "schedule" function should be float.')
        self.model.optimizer.learning_rate = float(new)

    def on_epoch_end(self, epoch, logs=None):
        if logs is not None:
            logs["lr"] = float(self.model.optimizer.learning_rate)


class TensorBoard(Callback):
    """Accepted and ignored: the reference only uses it for logging (models/CvT(Par).py:472); see DESIGN.md, out of scope."""

    def __init__(self, log_dir="logs", histogram_freq=0, **kw):
        self.log_dir, self.histogram_freq = log_dir, histogram_freq


class History(Callback):
    def __init__(self):
        self.history: Dict[str, List[float]] = {}
        self.epoch: List[int] = []

    def on_epoch_end(self, epoch, logs=None):
        self.epoch.append(epoch)
        for k, v in (logs or {}).items():
            self.history.setdefault(k, []).append(v)


def lr_scheduler(epoch, lr):
    """models/CvT(Par).py:357-360."""
    if epoch > 0 and epoch % 50 == 0:
        return lr * 0.8
    return lr


# ---- Keras initialisers so that an untrained model starts where the reference's would -----------------
def _glorot_fans(shape):
    if len(shape) == 2:
        return shape[0], shape[1]
    rec = int(np.prod(shape[:-2]))
    return shape[-2] * rec, shape[-1] * rec


def keras_initial_weights(param_infos, seed: Optional[int] = None) -> Dict[str, np.ndarray]:
    rng = np.random.default_rng(seed)
    out = {}
    for p in param_infos:
        leaf = p["name"].rsplit("/", 1)[-1]
        shape = tuple(p["shape"])
        if leaf in ("kernel", "depthwise_kernel"):
            fi, fo = _glorot_fans(shape)
            lim = math.sqrt(6.0 / (fi + fo))
            v = rng.uniform(-lim, lim, size=shape)
        elif leaf in ("gamma", "moving_variance"):
            v = np.ones(shape)
        else:
            v = np.zeros(shape)
        out[p["name"]] = v.astype(np.float32)
    return out


class _Layer:
    """What `model.get_layer(name)` returns: enough for tools/grad_cam_CvT.py:430-433 (`.output` names the tap)."""

    def __init__(self, model, name, tap):
        self.model, self.name, self.output = model, name, tap

    def get_weights(self):
        w = self.model.get_weights_dict()
        return [v for k, v in w.items() if k.startswith(self.name + "/")]


def _as_inputs(x, has_proc, cast=True):
    """Keras input conformance: [x, proc] list (or bare x for the Img model); images [N,H,W] get a channel axis; fp32.
    cast=False keeps float64 arrays as they are (the library converts them while it stages them: no host-side copy of the data set)."""
    if isinstance(x, (list, tuple)):
        if has_proc:
            if len(x) != 2:
                raise ValueError("Layer expects 2 input(s), but it received %d input tensors" % len(x))
            img, proc = x
        else:
            img, proc = x[0], None
    else:
        if has_proc:
            raise ValueError("Layer expects 2 input(s), but it received 1 input tensors")
        img, proc = x, None

    def conform(a):
        a = np.asarray(a)
        if cast or a.dtype not in (np.float32, np.float64) or not a.flags["C_CONTIGUOUS"]:
            a = np.ascontiguousarray(a, dtype=np.float32)
        return a

    img = conform(img)
    if img.ndim == 3:
        img = img[..., None]
    if proc is not None:
        proc = conform(proc)
        if proc.shape[0] != img.shape[0]:
            raise ValueError("Data cardinality is ambiguous: image and process-parameter sample counts differ")
    return img, proc


def _as_targets(y, n, nc, cast=True):
    a = np.asarray(y)
    if cast or a.dtype not in (np.float32, np.float64) or not a.flags["C_CONTIGUOUS"]:
        a = np.ascontiguousarray(a, dtype=np.float32)
    a = a.reshape(n, -1)
    if a.shape[1] != nc:
        raise ValueError("targets have %d columns, the model has num_classes=%d" % (a.shape[1], nc))
    return a


class Model:
    """The object `create_cvt_model` returns.  Weights live on the GPU inside one cvt_model (libcvt_b200)."""

    def __init__(self, cfg: CvtConfig, engine_factory=CvtEngine, seed: Optional[int] = None):
        self.cfg = cfg
        self._factory = engine_factory
        self._eng = None
        self._seed = seed
        self._pending_weights: Optional[Dict[str, np.ndarray]] = None
        self.optimizer: Optional[Adam] = None
        self.loss = None
        self.metrics_names: List[str] = []
        self.history = History()
        self.stop_training = False
        self.inputs = ["Image_inputs"] + (["Proc_inputs"] if cfg.proc_dim > 0 else [])
        self.output = "pred"
        # seed=None behaves like Keras: fresh dropout masks and a fresh shuffle order for every model (so the 8 "(many)" repeats differ)
        self.dropout_seed = int.from_bytes(os.urandom(6), "little") if seed is None else int(seed)
        self._rng = np.random.default_rng(self.dropout_seed + 977)       # ONE shuffle stream per model, carried across fit() calls
        self.resident_limit_bytes = 64 << 30                              # data sets up to this size are kept in HBM during fit()

    # -- engine lifetime ---------------------------------------------------------------------------------------
    @property
    def engine(self):
        if self._eng is None:
            self._eng = self._factory(self.cfg)
            w = self._pending_weights
            if w is None:
                w = keras_initial_weights(self._eng.param_infos(), self._seed)
            self._eng.set_params(w)
            self._pending_weights = None
        return self._eng

    def _ensure_batch(self, batch):
        """The workspace is sized for cfg.max_batch; a larger training batch needs a bigger one (weights carried over)."""
        if batch <= self.cfg.max_batch:
            return
        if self._eng is not None:
            if self._eng.optimizer_steps() > 0:
                raise ValueError("batch_size %d exceeds the model's max_batch %d after training started; create the model with "
                                 "max_batch >= batch_size" % (batch, self.cfg.max_batch))
            self._pending_weights = self._eng.get_params()
            self._eng.close()
            self._eng = None
        self.cfg.max_batch = int(batch)

    def close(self):
        if self._eng is not None:
            self._eng.close()
            self._eng = None

    # -- keras.Model surface ----------------------------------------------------------------------------------------
    def compile(self, optimizer="adam", loss=None, metrics=None, **kw):
        if isinstance(optimizer, str):
            if optimizer.lower() != "adam":
                raise ValueError("Unknown optimizer: %s (the CUDA path implements Adam)" % optimizer)
            optimizer = Adam()
        if not isinstance(optimizer, Adam):
            raise ValueError("optimizer must be keras_api.Adam")
        if loss not in ("mean_squared_error", "mse", "MSE"):
            raise ValueError("Unknown loss function: %s (the fused head kernel implements mean_squared_error)" % (loss,))
        metrics = list(metrics or [])
        for m in metrics:
            if m not in ("mae", "mean_absolute_error"):
                raise ValueError("Unknown metric function: %s (the fused head kernel reports mae)" % (m,))
        self.optimizer, self.loss = optimizer, "mean_squared_error"
        self.metrics_names = ["loss"] + (["mae"] if metrics else [])
        if self._eng is not None:
            self._eng.optimizer_reset()

    def count_params(self):
        return int(sum(p["numel"] for p in self.engine.param_infos()))

    def summary(self, print_fn=print):
        tot = tr = 0
        for p in self.engine.param_infos():
            print_fn("%-70s %-18s %9d%s" % (p["name"], p["shape"], p["numel"], "" if p["trainable"] else "  (non-trainable)"))
            tot += p["numel"]; tr += p["numel"] if p["trainable"] else 0
        print_fn("Total params: %d\nTrainable params: %d\nNon-trainable params: %d" % (tot, tr, tot - tr))

    def get_layer(self, name=None, index=None):
        n = len(self.cfg.stages)
        for i in range(1, n + 1):
            if name == "stage%d_transformer" % i:
                return _Layer(self, name, "s%d.out" % i)
            if name == "stage%d_ConvEmbed" % i:
                return _Layer(self, name, "s%d.embed" % i)
        if name in ("Proc_Dense_1", "Proc_Dense_2", "Final_Dense"):
            return _Layer(self, name, {"Proc_Dense_1": "head.proc1", "Proc_Dense_2": "head.feat", "Final_Dense": "pred"}[name])
        raise ValueError("No such layer: %s." % name)

    def fit(self, x=None, y=None, batch_size=None, epochs=1, verbose=1, callbacks=None, validation_data=None, shuffle=True,
            initial_epoch=0, **kw):
        """model.fit of models/CvT(Par).py:475-476.  The arrays stay where the caller has them (float64 or float32, pageable):
        the first epoch streams its batches through the library's pinned ring in shuffled order and, as a side effect, leaves the
        set resident in HBM; every later epoch is a permutation gathered on the device.  The validation set is uploaded once."""
        if self.optimizer is None:
            raise RuntimeError("You must compile your model before training/testing. Use `model.compile(optimizer, loss)`.")
        has_proc = self.cfg.proc_dim > 0
        img, proc = _as_inputs(x, has_proc, cast=False)
        nc = self.cfg.num_classes
        n = img.shape[0]
        yy = _as_targets(y, n, nc, cast=False)
        bs = int(batch_size or 32)
        self._ensure_batch(min(bs, n))
        eng = self.engine
        TRAIN, VAL = 0, 1
        has_val = False
        if validation_data is not None:
            vimg, vproc = _as_inputs(validation_data[0], has_proc, cast=False)
            vy = _as_targets(validation_data[1], vimg.shape[0], nc, cast=False)
            eng.dataset_alloc(VAL, vimg.shape[0])
            eng.dataset_write(VAL, 0, vimg, vproc, vy)
            has_val = True
        cbs = list(callbacks or [])
        self.history = History()
        for cb in cbs + [self.history]:
            cb.set_model(self)
        eng.set_dropout("philox" if self.cfg.drop_rate > 0 else "off", self.dropout_seed)      # same seed: the step counter keeps running
        per_sample = 4 * (self.cfg.image_h * self.cfg.image_w * self.cfg.in_chans + self.cfg.proc_dim + nc)
        resident = n * per_sample <= self.resident_limit_bytes and epochs - initial_epoch > 1
        filled = False
        if resident:
            try:
                eng.dataset_alloc(TRAIN, n)
            except Exception:
                resident = False
        self.stop_training = False
        try:
            for cb in cbs:
                cb.on_train_begin()
            for epoch in range(initial_epoch, epochs):
                for cb in cbs:
                    cb.on_epoch_begin(epoch)
                t0 = time.time()
                order = self._rng.permutation(n).astype(np.int32) if shuffle else None
                lr = float(self.optimizer.learning_rate)
                if resident and filled:
                    losses, maes = eng.train_epoch_resident(TRAIN, order, bs, lr)
                else:
                    losses, maes = eng.train_epoch_stream(img, proc, yy, order, bs, lr, TRAIN if resident else -1)
                    filled = resident
                counts = np.minimum(bs, n - bs * np.arange(len(losses)))
                lsum = float(np.dot(losses.astype(np.float64), counts)); asum = float(np.dot(maes.astype(np.float64), counts))
                logs = {"loss": lsum / n}
                if "mae" in self.metrics_names:
                    logs["mae"] = asum / n
                if has_val:
                    vl, va = eng.evaluate_resident(VAL)
                    logs["val_loss"] = vl
                    if "mae" in self.metrics_names:
                        logs["val_mae"] = va
                for cb in cbs:
                    cb.on_epoch_end(epoch, logs)
                self.history.on_epoch_end(epoch, logs)
                if verbose:
                    print("Epoch %d/%d - %.1fs - %s" % (epoch + 1, epochs, time.time() - t0,
                                                        " - ".join("%s: %.4f" % kv for kv in logs.items())))
                if self.stop_training:
                    break
            for cb in cbs:
                cb.on_train_end()
        finally:
            if resident:
                eng.dataset_free(TRAIN)
            if has_val:
                eng.dataset_free(VAL)
        return self.history

    def predict(self, x, batch_size=None, verbose=0, **kw):
        img, proc = _as_inputs(x, self.cfg.proc_dim > 0, cast=False)
        # chunking by max_batch inside the library gives the same numbers as Keras' default batch 32 (inference has
        # no cross-sample coupling: BatchNorm uses its moving statistics)
        return self.engine.predict(img, proc)

    def evaluate(self, x=None, y=None, batch_size=None, verbose=0, return_dict=False, **kw):
        img, proc = _as_inputs(x, self.cfg.proc_dim > 0, cast=False)
        yy = _as_targets(y, img.shape[0], self.cfg.num_classes, cast=False)
        loss, mae = self.engine.evaluate(img, proc, yy)
        if return_dict:
            return {"loss": loss, "mae": mae}
        return [loss, mae] if "mae" in self.metrics_names else loss

    def train_on_batch(self, x, y, **kw):
        if self.optimizer is None:
            raise RuntimeError("You must compile your model before training/testing. Use `model.compile(optimizer, loss)`.")
        img, proc = _as_inputs(x, self.cfg.proc_dim > 0)
        self._ensure_batch(img.shape[0])
        yy = np.ascontiguousarray(np.asarray(y), dtype=np.float32).reshape(img.shape[0], -1)
        loss, mae = self.engine.train_step(img, proc, yy, float(self.optimizer.learning_rate))
        return [loss, mae] if "mae" in self.metrics_names else loss

    __call__ = predict

    # -- weights ----------------------------------------------------------------------------------------------------------
    def get_weights_dict(self) -> Dict[str, np.ndarray]:
        return self.engine.get_params()

    def get_weights(self):
        return list(self.get_weights_dict().values())

    def set_weights_dict(self, w: Dict[str, np.ndarray]):
        self.engine.set_params(w)

    def keras_layout(self):
        """[(keras layer name, [(keras weight name, canonical name), ...])] in the order Keras-2 `save_weights` writes an .h5 file for
        the model create_cvt_model builds: one entry per layer of model.layers (weight-less layers with an empty list), and inside a
        layer `trainable_weights + non_trainable_weights`, i.e. sub-layers in the order their attributes are assigned in __init__
        (models/CvT(Par).py:126-142,244-259) with the BatchNorm moving statistics LAST.  Keras' auto-numbered names
        (dense_3, layer_normalization_6) depend on a process-global counter: readers match by position and shape, as Keras itself does."""
        cfg = self.cfg
        out = [("Image_inputs", [])]
        dn = 0
        for i, st in enumerate(cfg.stages, 1):
            ce, t = "stage%d_ConvEmbed" % i, "stage%d_transformer" % i
            out.append((ce, [(ce + "/conv2d/kernel:0", ce + "/kernel"), (ce + "/conv2d/bias:0", ce + "/bias")]))
            tr, nt = [], []
            if st.with_cls_token:
                tr.append((t + "/cls_token:0", t + "/cls_token"))
            tr += [(t + "/layer_normalization/gamma:0", t + "/norm1/gamma"), (t + "/layer_normalization/beta:0", t + "/norm1/beta")]
            ca = t + "/conv_attention/"
            if st.qkv_method == "dw_bn":
                for b in "qkv":
                    tr += [(ca + b + "_proj/depthwise_conv2d/depthwise_kernel:0", t + "/attn/%s_proj/depthwise_kernel" % b),
                           (ca + b + "_proj/batch_normalization/gamma:0", t + "/attn/%s_proj/bn/gamma" % b),
                           (ca + b + "_proj/batch_normalization/beta:0", t + "/attn/%s_proj/bn/beta" % b)]
                    nt += [(ca + b + "_proj/batch_normalization/moving_mean:0", t + "/attn/%s_proj/bn/moving_mean" % b),
                           (ca + b + "_proj/batch_normalization/moving_variance:0", t + "/attn/%s_proj/bn/moving_variance" % b)]
            if st.qkv_method == "linear_dense":              # tools/grad_cam_CvT.py:103-104: the 'linear' Projection there is a Dense(dim)
                for b in "qkv":
                    d = "dense%s" % ("_%d" % dn if dn else ""); dn += 1
                    tr += [(ca + b + "_proj/" + d + "/kernel:0", t + "/attn/%s_proj/dense/kernel" % b),
                           (ca + b + "_proj/" + d + "/bias:0", t + "/attn/%s_proj/dense/bias" % b)]
            for b in "qkv":                                  # (built but never called in the tools' variants: Keras then never creates their weights -
                d = "dense%s" % ("_%d" % dn if dn else ""); dn += 1      #  filtered out below - but the name counter has moved on)
                tr += [(ca + d + "/kernel:0", t + "/attn/proj_%s/kernel" % b), (ca + d + "/bias:0", t + "/attn/proj_%s/bias" % b)]
            for b in ("query", "key", "value", "attention_output"):
                tr += [(ca + "multi_head_attention/%s/kernel:0" % b, t + "/attn/mha/%s/kernel" % b),
                       (ca + "multi_head_attention/%s/bias:0" % b, t + "/attn/mha/%s/bias" % b)]
            d = "dense_%d" % dn; dn += 1
            tr += [(ca + d + "/kernel:0", t + "/attn/proj/kernel"), (ca + d + "/bias:0", t + "/attn/proj/bias")]
            for fc in ("fc1", "fc2"):
                d = "dense_%d" % dn; dn += 1
                tr += [(t + "/sequential/" + d + "/kernel:0", t + "/mlp/%s/kernel" % fc), (t + "/sequential/" + d + "/bias:0", t + "/mlp/%s/bias" % fc)]
            out.append((t, tr + nt))
        ln = "layer_normalization_%d" % len(cfg.stages)
        if cfg.proc_dim > 0:
            out.append(("Proc_inputs", []))
        out.append((ln, [(ln + "/gamma:0", "head_norm/gamma"), (ln + "/beta:0", "head_norm/beta")]))
        if cfg.proc_dim > 0:
            for nm in ("Proc_Dense_1", "Proc_Dense_2"):
                out.append((nm, [(nm + "/kernel:0", nm + "/kernel"), (nm + "/bias:0", nm + "/bias")]))
            out.append(("concatenate", []))
        out.append(("Final_Dense", [("Final_Dense/kernel:0", "Final_Dense/kernel"), ("Final_Dense/bias:0", "Final_Dense/bias")]))
        # the model variants of the tools (configure(variant=...)) lack some of these weights: proj_q/k/v never applied, heads without a
        # LayerNormalization.  What the engine does not hold is not in the file; everything it holds has to be.
        have = [p["name"] for p in self.engine.param_infos()]
        out = [(ln, [(kn, cn) for kn, cn in ws if cn in set(have)]) for ln, ws in out]
        listed = [cn for _, ws in out for _, cn in ws]
        if sorted(listed) != sorted(have):
            missing = sorted(set(have) - set(listed))
            raise RuntimeError("keras_layout: no Keras-2 position known for %s" % (missing or "duplicated entries"))
        return out

    def save_weights(self, filepath, overwrite=True, save_format=None, **kw):
        """Keras-2 `model.save_weights(path)` (models/CvT(Par).py:489).  `.h5` / `.hdf5` / `.keras` paths (or save_format='h5') get the
        legacy Keras-2 HDF5 layout, written by h5lite (no h5py needed); any other path a flat name->array .npz archive."""
        if os.path.exists(filepath) and not overwrite:
            raise IOError("%s exists" % filepath)
        os.makedirs(os.path.dirname(os.path.abspath(filepath)), exist_ok=True)
        w = self.get_weights_dict()
        if save_format in ("h5", "hdf5", "keras") or str(filepath).lower().endswith((".h5", ".hdf5", ".keras")):
            h5lite.save_keras_weights(filepath, [(ln, [(kn, w[cn]) for kn, cn in ws]) for ln, ws in self.keras_layout()])
            return
        with open(filepath, "wb") as f:
            np.savez(f, __format__=np.array("cvt_b200_weights_v1"), **{k.replace("/", "|"): v for k, v in w.items()})

    def _weights_from_h5(self, filepath) -> Dict[str, np.ndarray]:
        """Matches a Keras-2 weight file to this model the way Keras' load_weights_from_hdf5_group does: layer by layer (by layer name;
        the anonymous head LayerNormalization by its position among the weighted layers), weights inside a layer by POSITION, every shape checked."""
        infos = {p["name"]: p for p in self.engine.param_infos()}
        saved = h5lite.load_keras_weights(filepath)
        layout = [(ln, ws) for ln, ws in self.keras_layout() if ws]
        filled = [(ln, ws) for ln, ws in saved.items() if ws]
        if len(filled) != len(layout):
            raise ValueError("Layer count mismatch when loading weights from file. Model expected %d layers, found %d saved layers."
                             % (len(layout), len(filled)))
        by_name = dict(filled)
        out = {}
        for k, (ln, ws) in enumerate(layout):
            if ln in by_name:
                fw = by_name[ln]
            elif ln.startswith("layer_normalization"):
                cands = [(n, v) for n, v in filled if n.startswith("layer_normalization")]
                if len(cands) != 1:
                    raise ValueError("cannot identify the head LayerNormalization among saved layers %s" % [n for n, _ in cands])
                fw = cands[0][1]
            else:
                raise ValueError("saved weights have no layer named %s" % ln)
            if len(fw) != len(ws):
                raise ValueError("Weight count mismatch for layer #%d (named %s). Layer expects %d weight(s). Received %d saved weight(s)"
                                 % (k, ln, len(ws), len(fw)))
            for (kn, cn), (fn, arr) in zip(ws, fw):
                want = tuple(infos[cn]["shape"])
                if tuple(arr.shape) != want:
                    raise ValueError("Shape mismatch in layer #%d (named %s) for weight %s. Weight expects shape %s. Received saved weight "
                                     "%s with shape %s" % (k, ln, kn, want, fn, tuple(arr.shape)))
                out[cn] = np.asarray(arr, dtype=np.float32)
        return out

    def load_weights(self, filepath, **kw):
        """Keras-2 `model.load_weights(path)` (models/CvT_test(Par).py:513): HDF5 files (ours or Keras' own) and the .npz archive."""
        if not os.path.exists(filepath):
            raise FileNotFoundError("Unable to open file (unable to open file: name = '%s')" % filepath)
        with open(filepath, "rb") as f:
            magic = f.read(8)
        if magic == h5lite.SIGNATURE:
            self.engine.set_params(self._weights_from_h5(filepath))
            return
        z = np.load(filepath, allow_pickle=False)
        w = {k.replace("|", "/"): z[k] for k in z.files if k != "__format__"}
        infos = {p["name"]: p for p in self.engine.param_infos()}
        missing = sorted(set(infos) - set(w)); extra = sorted(set(w) - set(infos))
        if missing or extra:
            raise ValueError("weight file does not match the model: missing %s, unexpected %s" % (missing[:3], extra[:3]))
        for k, v in w.items():
            if tuple(v.shape) != tuple(infos[k]["shape"]):
                raise ValueError("Shape mismatch in %s: file %s, model %s" % (k, v.shape, infos[k]["shape"]))
        self.engine.set_params(w)

    # -- Grad-CAM (tools/grad_cam_CvT.py:430-457): Model(inputs, [layer.output, output]) under a GradientTape -----------------------
    def feature_and_gradient(self, x, layer_name: str):
        """Returns (feature map [B,h,w,D], predictions [B,nc], d predictions[:,0].sum() / d feature map)."""
        img, proc = _as_inputs(x, self.cfg.proc_dim > 0)
        layer = self.get_layer(layer_name)
        if not layer.output.endswith(".out"):
            raise ValueError("gradients are available w.r.t. a stage{i}_transformer output")
        return self.engine.gradcam(img, proc, int(layer.output[1]) - 1)


def create_cvt_model(image_height, image_width, num_channels, proc_dim=None, num_classes=None, *, precision="tf32", max_batch=128,
                     device=0, seed=None, engine_factory=CvtEngine) -> Model:
    """create_cvt_model of models/CvT(Par).py:292 (five positional arguments) and models/CvT(Img).py:291 (four: no proc_dim).
    Reads the module-level `spec` like the reference does."""
    if num_classes is None:              # the Img signature: (image_height, image_width, num_channels, num_classes)
        num_classes, proc_dim = proc_dim, 0
    if num_classes is None:
        raise TypeError("create_cvt_model() missing 1 required positional argument: 'num_classes'")
    cfg = CvtConfig.from_spec_dict(spec, int(image_height), int(image_width), int(num_channels), int(proc_dim or 0), int(num_classes),
                                   precision=precision, max_batch=int(max_batch), device=int(device))
    if model_variant == "gradcam":       # tools/grad_cam_CvT.py:315-349
        if any(st.qkv_method == "avg" for st in cfg.stages):
            # found by executing the tool's own source (tests/test_oracle_vs_reference_source.py): its ConvAttention hands q the 'linear'
            # Projection when the method is 'avg' (:141), and in this script 'linear' is a Dense(dim) (:103) - q through a Dense, k / v pooled
            raise NotImplementedError("tools/grad_cam_CvT.py with projection_method='avg' mixes a Dense q-projection with average-pooled k / v: "
                                      "that combination is not implemented (dw_bn and linear are)")
        for st in cfg.stages:
            st.with_cls_token = bool(cls_token_switch)
            if st.qkv_method == "linear":
                st.qkv_method = "linear_dense"
        cfg.head, cfg.apply_proj_qkv, cfg.attn_drop_rate = "mean", False, 0.1
    elif model_variant == "model_plot":  # tools/model_plot.py:289-292
        cfg.head = "cls" if cfg.stages[-1].with_cls_token else "mean"
    return Model(cfg, engine_factory=engine_factory, seed=seed)


def make_gradcam_heatmap(img_array, proc_array, model: Model, last_conv_layer_name, pred_index=None, batch_size=1):
    """tools/grad_cam_CvT.py:422-481 on the CUDA path: per batch, grads = d pred[:,0] / d feature map, channel weights = mean of
    the gradient over (batch, h, w), heat-map = relu(sum_c w_c A[0]) / max.  Returns [n_batches, h, w]."""
    img_array = np.asarray(img_array)
    heatmaps = []
    for o in range(0, img_array.shape[0], batch_size):
        xs = img_array[o:o + batch_size]
        ps = None if proc_array is None else np.asarray(proc_array)[o:o + batch_size]
        feat, pred, grads = model.feature_and_gradient([xs, ps] if ps is not None else xs, last_conv_layer_name)
        if not np.any(grads):
            # the reference fails here too: tape.gradient returns None when the output does not depend on the layer (the cls-token head of
            # models/CvT(Par).py:326-329 reads the cls row only, never the spatial map `get_layer(...).output` names)
            raise ValueError("No gradients: the prediction does not depend on the output map of %r (cls-token head); Grad-CAM needs the "
                             "pooled head of tools/grad_cam_CvT.py (configure(variant='gradcam')) or cls_token=False" % (last_conv_layer_name,))
        pooled = grads.mean(axis=(0, 1, 2))
        heat = (pooled * feat[0]).sum(axis=-1)
        heat = np.maximum(heat, 0) / heat.max()
        heatmaps.append(heat)
    return np.stack(heatmaps, axis=0)


def save_predictions_to_excel(predictions, actual, train_num, test_num, r2, mse, mae, path):
    """The sheet tools/Compare_r.py:21-24 and tools/Compare_plot.py:13-26 read (writer: models/CvT_test(Par).py:440-477): columns
    Predictions, Actual, Errors(%), then Train mounts, Test mounts, R2 Score, MSE, MAE filled in the first row only.  Written as a
    real .xlsx by xlsxlite (pandas' own writer needs openpyxl)."""
    from . import xlsxlite
    predictions = np.asarray(predictions, dtype=np.float64).flatten(); actual = np.asarray(actual, dtype=np.float64).flatten()
    cols = {"Predictions": predictions.tolist(), "Actual": actual.tolist(), "Errors(%)": (np.abs(predictions - actual) / actual * 100).tolist(),
            "Train mounts": [train_num], "Test mounts": [test_num], "R2 Score": [float(r2)], "MSE": [float(mse)], "MAE": [float(mae)]}
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    xlsxlite.write_xlsx(path, cols)
    return path


def save_history_to_excel(history, path):
    """models/CvT(Par).py:492-494: `pd.DataFrame(model.history.history)` with an `epoch` column (1-based) in front, index=False."""
    from . import xlsxlite
    h = history.history if hasattr(history, "history") else dict(history)
    n = max([len(v) for v in h.values()] + [0])
    cols = {"epoch": list(range(1, n + 1))}
    cols.update({k: [float(x) for x in v] for k, v in h.items()})
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    xlsxlite.write_xlsx(path, cols)
    return path


def regression_metrics(y_true, y_pred):
    """sklearn's r2_score / mean_squared_error / mean_absolute_error as models/CvT_test(Par).py:525-527 calls them (single output)."""
    t = np.asarray(y_true, dtype=np.float64).flatten(); p = np.asarray(y_pred, dtype=np.float64).flatten()
    ss_res = float(((t - p) ** 2).sum()); ss_tot = float(((t - t.mean()) ** 2).sum())
    r2 = 1.0 - ss_res / ss_tot if ss_tot > 0 else (1.0 if ss_res == 0 else 0.0)
    return r2, float(((t - p) ** 2).mean()), float(np.abs(t - p).mean())
