"""TEST INFRASTRUCTURE: a stand-in for transformer_stm_b200.CvtEngine that runs the CPU oracle, so that the host-side logic of
keras_api (fit loop, shuffling, partial batches, LR schedule, history, weight files, Grad-CAM plumbing) can be tested without a GPU,
and so that the GPU test of the same call sequence has an independent trajectory to compare with.  Never imported by the product."""
import numpy as np

from oracle import cvt_oracle as O


class OracleEngine:
    def __init__(self, cfg):
        self.cfg = cfg
        st = [O.StageSpec(s.embed_dim, s.patch_size, s.stride, s.num_heads, s.kernel_size, s.strides, s.qkv_method, s.with_cls_token)
              for s in cfg.stages]
        self.spec = O.CvtSpec(cfg.image_h, cfg.image_w, cfg.in_chans, st, cfg.proc_dim, cfg.num_classes, cfg.mlp_ratio, cfg.drop_rate,
                              cfg.ln_eps, cfg.bn_eps, cfg.bn_momentum, cfg.proc_hidden)
        # the tools' model variants (CvtConfig.head / apply_proj_qkv / attn_drop_rate, include/cvt_b200.h ABI 3)
        self.spec.head = {"gap": "mean"}.get(getattr(cfg, "head", "auto"), getattr(cfg, "head", "auto"))
        self.spec.apply_proj_qkv = bool(getattr(cfg, "apply_proj_qkv", True))
        self.spec.attn_drop_rate = float(getattr(cfg, "attn_drop_rate", 0.0))
        self.params = {n: np.zeros(s, np.float64) for n, s, _ in O.param_specs(self.spec)}
        self.opt = O.new_opt_state(self.params)
        self.mode, self.seed, self.step = "off", 0, 0
        self.ds = {}
        self.calls = []

    # ---- weights ----
    def param_infos(self):
        return [dict(index=i, name=n, shape=tuple(s), trainable=t, numel=int(np.prod(s))) for i, (n, s, t) in enumerate(O.param_specs(self.spec))]

    def set_params(self, w):
        for k, v in w.items():
            if k not in self.params:
                raise KeyError(k)
            self.params[k] = np.asarray(v, dtype=np.float32).astype(np.float64).reshape(self.params[k].shape)

    def get_params(self):
        return {k: v.astype(np.float32) for k, v in self.params.items()}

    def optimizer_steps(self):
        return self.opt["t"]

    def optimizer_reset(self):
        self.opt = O.new_opt_state(self.params)

    def set_dropout(self, mode, seed=0):
        if (mode, seed) != (self.mode, self.seed):
            self.mode, self.seed, self.step = mode, seed, 0

    def close(self):
        pass

    # ---- steps ----
    def _img(self, a):
        return np.asarray(a, dtype=np.float32).reshape(-1, self.cfg.image_h, self.cfg.image_w, self.cfg.in_chans)

    def _train_batch(self, img, proc, y, lr):
        masks = None
        if self.mode == "philox" and self.cfg.drop_rate > 0:
            masks = O.make_dropout_masks(self.spec, img.shape[0], self.seed, self.step)
            self.step += 1
        r = O.train_step(self.spec, self.params, self.opt, img, proc, np.asarray(y, np.float32).reshape(img.shape[0], -1), lr, masks)
        return r["loss"], r["mae"]

    def train_step(self, img, proc, y, lr):
        return self._train_batch(self._img(img), None if proc is None else np.asarray(proc, np.float32), y, lr)

    def _epoch(self, img, proc, y, order, batch, lr):
        n = len(order)
        L, A = [], []
        for o in range(0, n, batch):
            idx = order[o:o + batch]
            l, a = self._train_batch(self._img(img[idx]), None if proc is None else np.asarray(proc[idx], np.float32), np.asarray(y)[idx], lr)
            L.append(l); A.append(a)
        return np.asarray(L, np.float32), np.asarray(A, np.float32)

    def train_epoch_stream(self, img, proc, y, order, batch, lr, fill_slot=-1):
        self.calls.append(("stream", fill_slot))
        order = np.arange(len(img)) if order is None else np.asarray(order)
        if fill_slot >= 0:
            self.ds[fill_slot] = (np.asarray(img), None if proc is None else np.asarray(proc), np.asarray(y))
        return self._epoch(np.asarray(img), None if proc is None else np.asarray(proc), y, order, batch, lr)

    def train_epoch_resident(self, slot, order, batch, lr):
        self.calls.append(("resident", slot))
        img, proc, y = self.ds[slot]
        order = np.arange(len(img)) if order is None else np.asarray(order)
        return self._epoch(img, proc, y, order, batch, lr)

    def train_epoch(self, img, proc, y, batch, lr):
        return self.train_epoch_stream(img, proc, y, None, batch, lr, -1)

    def dataset_alloc(self, slot, n):
        self.ds[slot] = None

    def dataset_free(self, slot):
        self.ds.pop(slot, None)

    def dataset_write(self, slot, offset, img=None, proc=None, y=None):
        assert offset == 0
        self.ds[slot] = (np.asarray(img), None if proc is None else np.asarray(proc), np.asarray(y))

    def predict(self, img, proc=None):
        pred, _ = O.forward(self.spec, self.params, self._img(img), None if proc is None else np.asarray(proc, np.float32))
        return pred.astype(np.float32)

    def evaluate(self, img, proc, y):
        p = self.predict(img, proc).astype(np.float64)
        d = p - np.asarray(y, np.float64).reshape(p.shape)
        return float((d ** 2).mean()), float(np.abs(d).mean())

    def evaluate_resident(self, slot):
        img, proc, y = self.ds[slot]
        return self.evaluate(img, proc, y)

    def gradcam(self, img, proc, stage):
        heat, feat, g = O.gradcam(self.spec, self.params, self._img(img), None if proc is None else np.asarray(proc, np.float32), "s%d.out" % (stage + 1))
        return feat.astype(np.float32), self.predict(img, proc), g.astype(np.float32)
