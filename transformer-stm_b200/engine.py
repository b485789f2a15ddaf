"""Thin object wrapper over the C-ABI: one CvtEngine == one cvt_model on one GPU."""
import ctypes as C
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import numpy as np

from . import _lib as L


@dataclass
class StageSpec:
    """One entry of the reference's spec['stages'] (models/CvT(Par).py:66-72)."""
    embed_dim: int
    patch_size: int
    stride: int
    num_heads: int
    kernel_size: int = 3
    strides: int = 1
    qkv_method: str = "dw_bn"
    with_cls_token: bool = False


@dataclass
class CvtConfig:
    image_h: int = 128
    image_w: int = 128
    in_chans: int = 1
    stages: List[StageSpec] = field(default_factory=list)
    proc_dim: int = 5
    proc_hidden: int = 256
    num_classes: int = 1
    mlp_ratio: int = 4
    drop_rate: float = 0.1
    ln_eps: float = 1e-6
    bn_eps: float = 1e-3
    bn_momentum: float = 0.99
    precision: str = "fp32"
    max_batch: int = 64
    device: int = 0
    head: str = "auto"                  # auto | cls_ln | ln_mean | cls | mean (alias gap): include/cvt_b200.h CVT_HEAD_*
    apply_proj_qkv: bool = True         # False: tools/grad_cam_CvT.py never calls proj_q/k/v
    attn_drop_rate: float = 0.0         # tools/grad_cam_CvT.py:200

    @staticmethod
    def gradcam(cls_token_switch=False, projection_method="dw_bn", **kw) -> "CvtConfig":
        """The model of tools/grad_cam_CvT.py:315-349: cls token (if switched on) on every stage, proj_q/k/v not applied, dropout on the
        attention output, `linear` = Dense(dim), GlobalAveragePooling2D head without LayerNorm."""
        c = CvtConfig.reference(projection_method, cls_token_switch, **kw)
        for st in c.stages:
            st.with_cls_token = bool(cls_token_switch)
            if projection_method == "linear":
                st.qkv_method = "linear_dense"
        c.head, c.apply_proj_qkv, c.attn_drop_rate = "mean", False, 0.1
        return c

    @staticmethod
    def reference(projection_method="dw_bn", cls_token_switch=True, proc_dim=5, num_classes=1, image_h=128, image_w=128,
                  dims=(64, 128, 256), heads=(1, 2, 4), **kw) -> "CvtConfig":
        st = [StageSpec(dims[0], 7, 4, heads[0], 3, 1, projection_method, False),
              StageSpec(dims[1], 3, 2, heads[1], 3, 1, projection_method, False),
              StageSpec(dims[2], 3, 2, heads[2], 3, 1, projection_method, bool(cls_token_switch))]
        return CvtConfig(image_h, image_w, 1, st, proc_dim, num_classes=num_classes, **kw)

    @staticmethod
    def from_spec_dict(spec: dict, image_h, image_w, in_chans, proc_dim, num_classes, **kw) -> "CvtConfig":
        """Builds a config from the reference's own `spec` dict layout."""
        st = [StageSpec(s["embed_dim"], s["patch_size"], s["stride"], s["num_heads"], s.get("kernel_size", 3),
                        s.get("strides", 1), s.get("qkv_method", "dw_bn"), bool(s.get("with_cls_token", False)))
              for s in spec["stages"]]
        return CvtConfig(image_h, image_w, in_chans, st, proc_dim, num_classes=num_classes, **kw)

    def to_c(self) -> L.Config:
        c = L.Config()
        c.image_h, c.image_w, c.in_chans, c.n_stages = self.image_h, self.image_w, self.in_chans, len(self.stages)
        if len(self.stages) > L.MAX_STAGES:
            raise ValueError("at most %d stages" % L.MAX_STAGES)
        for i, s in enumerate(self.stages):
            if s.kernel_size != 3 or s.strides != 1:
                raise ValueError("Projection kernel_size must be 3 with stride 1 (every reference spec)")
            if s.qkv_method not in L.PROJ:
                raise ValueError("Unknown method: %s" % s.qkv_method)        # models/CvT(Par).py:100
            c.stage[i] = L.StageConfig(s.embed_dim, s.patch_size, s.stride, s.num_heads, L.PROJ[s.qkv_method], int(s.with_cls_token))
        c.proc_dim, c.proc_hidden, c.num_classes, c.mlp_ratio = self.proc_dim, self.proc_hidden, self.num_classes, self.mlp_ratio
        c.drop_rate, c.ln_eps, c.bn_eps, c.bn_momentum = self.drop_rate, self.ln_eps, self.bn_eps, self.bn_momentum
        c.precision, c.max_batch, c.device = L.PREC[self.precision], self.max_batch, self.device
        if self.head not in L.HEAD:
            raise ValueError("Unknown head: %s" % self.head)
        c.head_kind, c.skip_proj_qkv, c.attn_drop_rate = L.HEAD[self.head], 0 if self.apply_proj_qkv else 1, float(self.attn_drop_rate)
        return c


def _f32(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float32)
    if shape is not None:
        a = a.reshape(shape)
    return a


def _ptr(a):
    """device pointer from an int, a torch tensor or anything with data_ptr()."""
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    if hasattr(a, "data_ptr"):
        return C.c_void_p(a.data_ptr())
    raise TypeError("expected a device pointer (int) or a tensor with data_ptr()")


class CvtEngine:
    def __init__(self, cfg: CvtConfig):
        self.lib = L.load()
        self.cfg = cfg
        self._c = cfg.to_c()
        h = C.c_void_p()
        L.check(self.lib.cvt_create(C.byref(self._c), C.byref(h)), "cvt_create")
        self.h = h
        self._params = None

    def close(self):
        if getattr(self, "h", None):
            self.lib.cvt_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- weights ----
    def param_infos(self):
        if self._params is None:
            out = []
            n = L.check(self.lib.cvt_param_count(self.h))
            for i in range(n):
                name, rank, tr, numel = C.c_char_p(), C.c_int32(), C.c_int32(), C.c_int64()
                shape = (C.c_int64 * 4)()
                L.check(self.lib.cvt_param_info(self.h, i, C.byref(name), C.byref(rank), shape, C.byref(tr), C.byref(numel)))
                out.append(dict(index=i, name=name.value.decode(), shape=tuple(shape[:rank.value]), trainable=bool(tr.value), numel=numel.value))
            self._params = out
        return self._params

    def param_names(self):
        return [p["name"] for p in self.param_infos()]

    def set_params(self, params: Dict[str, np.ndarray]):
        infos = {p["name"]: p for p in self.param_infos()}
        for k, v in params.items():
            if k not in infos:
                raise KeyError("unknown parameter %s" % k)
            p = infos[k]
            a = _f32(v)
            if a.size != p["numel"]:
                raise ValueError("%s: expected %s, got %s" % (k, p["shape"], a.shape))
            L.check(self.lib.cvt_set_param(self.h, p["index"], a.ctypes.data_as(C.c_void_p)), "cvt_set_param")

    def get_params(self) -> Dict[str, np.ndarray]:
        out = {}
        for p in self.param_infos():
            a = np.empty(p["shape"], dtype=np.float32)
            L.check(self.lib.cvt_get_param(self.h, p["index"], a.ctypes.data_as(C.c_void_p)), "cvt_get_param")
            out[p["name"]] = a
        return out

    def get_grads(self) -> Dict[str, np.ndarray]:
        out = {}
        for p in self.param_infos():
            if not p["trainable"]:
                continue
            a = np.empty(p["shape"], dtype=np.float32)
            L.check(self.lib.cvt_get_grad(self.h, p["index"], a.ctypes.data_as(C.c_void_p)), "cvt_get_grad")
            out[p["name"]] = a
        return out

    def init_params(self, seed=0):
        L.check(self.lib.cvt_init_params(self.h, seed), "cvt_init_params")

    # ---- host-buffer entry points ----
    def _img(self, img, n=None):
        c = self.cfg
        a = _f32(img)
        return a.reshape((-1, c.image_h, c.image_w, c.in_chans))

    def predict(self, img, proc=None) -> np.ndarray:
        """model.predict on the caller's arrays as they are (float32 / float64, pageable): chunks of max_batch, copies double-buffered."""
        if self.cfg.proc_dim > 0 and proc is None:
            raise L.CvtError("cvt_predict failed: this model needs process parameters")
        ha, keep = self._host_arrays(img, proc, None)
        out = np.empty((ha.n, self.cfg.num_classes), dtype=np.float32)
        L.check(self.lib.cvt_predict_arrays(self.h, C.byref(ha), out.ctypes.data_as(C.c_void_p)), "cvt_predict_arrays")
        return out

    def train_step(self, img, proc, y, lr: float):
        a = self._img(img)
        n = a.shape[0]
        p = _f32(proc, (n, self.cfg.proc_dim)) if (self.cfg.proc_dim > 0 and proc is not None) else None
        t = _f32(y, (n, self.cfg.num_classes))
        loss, mae = C.c_float(), C.c_float()
        L.check(self.lib.cvt_train_step_host(self.h, a.ctypes.data_as(C.c_void_p), p.ctypes.data_as(C.c_void_p) if p is not None else None,
                                             t.ctypes.data_as(C.c_void_p), n, lr, C.byref(loss), C.byref(mae)), "cvt_train_step_host")
        return loss.value, mae.value

    def train_epoch(self, img, proc, y, batch: int, lr: float):
        """One pass over (img, proc, y) in the given order, batches of `batch` (last partial batch kept, models/CvT(Par).py:475);
        returns per-batch (loss, mae) arrays.  Copies of batch i+1 overlap the step on batch i."""
        return self.train_epoch_stream(img, proc, y, None, batch, lr, -1)

    def evaluate(self, img, proc, y):
        ha, keep = self._host_arrays(img, proc, y)
        loss, mae = C.c_float(), C.c_float()
        L.check(self.lib.cvt_evaluate_arrays(self.h, C.byref(ha), C.byref(loss), C.byref(mae)), "cvt_evaluate_arrays")
        return loss.value, mae.value

    # ---- device-pointer entry points (torch tensors or raw ints) ----
    def forward_dev(self, img_dev, proc_dev, batch: int, training: bool, out_dev=None):
        L.check(self.lib.cvt_forward(self.h, _ptr(img_dev), _ptr(proc_dev), batch, int(training), _ptr(out_dev)), "cvt_forward")

    def loss_backward_dev(self, y_dev, read=True):
        loss, mae = C.c_float(), C.c_float()
        L.check(self.lib.cvt_loss_backward(self.h, _ptr(y_dev), C.byref(loss) if read else None, C.byref(mae) if read else None), "cvt_loss_backward")
        return loss.value, mae.value

    def adam_step(self, lr: float):
        L.check(self.lib.cvt_adam_step(self.h, lr), "cvt_adam_step")

    def train_step_dev(self, img_dev, proc_dev, y_dev, batch: int, lr: float):
        L.check(self.lib.cvt_train_step(self.h, _ptr(img_dev), _ptr(proc_dev), _ptr(y_dev), batch, lr), "cvt_train_step")

    def read_metrics(self):
        loss, mae = C.c_float(), C.c_float()
        L.check(self.lib.cvt_read_metrics(self.h, C.byref(loss), C.byref(mae)), "cvt_read_metrics")
        return loss.value, mae.value

    def optimizer_reset(self):
        L.check(self.lib.cvt_optimizer_reset(self.h))

    def optimizer_steps(self) -> int:
        return int(self.lib.cvt_optimizer_step_count(self.h))

    # ---- resident datasets / streamed epochs (include/cvt_b200.h, "input pipeline") ----
    def _host_arrays(self, img, proc, y):
        """cvt_host_arrays over the caller's numpy arrays without copying them when they are C-contiguous float32 / float64
        (the reference hands model.fit float64 [N,H,W], [N,5], [N]); returns (struct, keep-alive list)."""
        c = self.cfg
        keep = []

        def conform(a, tail):
            if a is None:
                return None, 0
            a = np.asarray(a)
            if a.dtype not in (np.float32, np.float64) or not a.flags["C_CONTIGUOUS"]:
                a = np.ascontiguousarray(a, dtype=np.float32)
            if int(np.prod(a.shape[1:], dtype=np.int64)) != tail:
                raise ValueError("expected rows of %d values, got array of shape %s" % (tail, a.shape))
            keep.append(a)
            return a.ctypes.data, int(a.dtype == np.float64)

        n = len(img) if img is not None else (len(y) if y is not None else len(proc))
        ha = L.HostArrays()
        ha.img, ha.img_f64 = conform(img, c.image_h * c.image_w * c.in_chans)
        ha.proc, ha.proc_f64 = conform(proc if c.proc_dim > 0 else None, c.proc_dim)
        ha.y, ha.y_f64 = conform(y, c.num_classes)
        for a in keep:
            if a.shape[0] != n:
                raise ValueError("Data cardinality is ambiguous: arrays of %d and %d samples" % (n, a.shape[0]))
        ha.n = n
        return ha, keep

    def dataset_alloc(self, slot: int, n: int):
        L.check(self.lib.cvt_dataset_alloc(self.h, slot, n), "cvt_dataset_alloc")

    def dataset_free(self, slot: int):
        L.check(self.lib.cvt_dataset_free(self.h, slot), "cvt_dataset_free")

    def dataset_count(self, slot: int) -> int:
        return int(self.lib.cvt_dataset_count(self.h, slot))

    def dataset_write(self, slot: int, offset: int, img=None, proc=None, y=None):
        ha, keep = self._host_arrays(img, proc, y)
        L.check(self.lib.cvt_dataset_write(self.h, slot, offset, C.byref(ha)), "cvt_dataset_write")

    def dataset_write_u8(self, slot: int, offset: int, pixels: np.ndarray):
        """pixels: uint8 [count, src_h, src_w, 3] (cv2.imread layout, BGR) or [count, src_h, src_w] / [.., 1] (already grey)."""
        a = np.ascontiguousarray(pixels, dtype=np.uint8)
        if a.ndim == 3:
            a = a[..., None]
        cnt, sh, sw, sc = a.shape
        L.check(self.lib.cvt_dataset_write_u8(self.h, slot, offset, cnt, a.ctypes.data_as(C.c_void_p), sh, sw, sc), "cvt_dataset_write_u8")

    def dataset_read(self, slot: int, offset: int, count: int):
        c = self.cfg
        img = np.empty((count, c.image_h, c.image_w, c.in_chans), np.float32)
        proc = np.empty((count, c.proc_dim), np.float32) if c.proc_dim > 0 else None
        y = np.empty((count, c.num_classes), np.float32)
        L.check(self.lib.cvt_dataset_read(self.h, slot, offset, count, img.ctypes.data_as(C.c_void_p),
                                          proc.ctypes.data_as(C.c_void_p) if proc is not None else None, y.ctypes.data_as(C.c_void_p)), "cvt_dataset_read")
        return img, proc, y

    @staticmethod
    def _order(order):
        if order is None:
            return None, 0, None
        o = np.ascontiguousarray(order, dtype=np.int32)
        return o.ctypes.data_as(C.c_void_p), o.size, o

    def train_epoch_stream(self, img, proc, y, order, batch: int, lr: float, fill_slot: int = -1):
        """One fit epoch over host arrays in `order` (None = as stored): rows are gathered into the library's pinned ring, copied while
        the previous step runs and, with fill_slot >= 0, also stored into that resident dataset."""
        ha, keep = self._host_arrays(img, proc, y)
        op, on, okeep = self._order(order)
        n = on if order is not None else ha.n
        nb = (n + batch - 1) // batch
        loss = np.empty(nb, np.float32); mae = np.empty(nb, np.float32)
        L.check(self.lib.cvt_train_epoch_stream(self.h, C.byref(ha), op, on, int(batch), lr, int(fill_slot), loss.ctypes.data_as(C.c_void_p),
                                                mae.ctypes.data_as(C.c_void_p)), "cvt_train_epoch_stream")
        return loss, mae

    def train_epoch_resident(self, slot: int, order, batch: int, lr: float):
        op, on, okeep = self._order(order)
        n = on if order is not None else self.dataset_count(slot)
        nb = (n + batch - 1) // batch
        loss = np.empty(nb, np.float32); mae = np.empty(nb, np.float32)
        L.check(self.lib.cvt_train_epoch_resident(self.h, slot, op, on, int(batch), lr, loss.ctypes.data_as(C.c_void_p),
                                                  mae.ctypes.data_as(C.c_void_p)), "cvt_train_epoch_resident")
        return loss, mae

    def evaluate_resident(self, slot: int):
        loss, mae = C.c_float(), C.c_float()
        L.check(self.lib.cvt_evaluate_resident(self.h, slot, C.byref(loss), C.byref(mae)), "cvt_evaluate_resident")
        return loss.value, mae.value

    def predict_resident(self, slot: int) -> np.ndarray:
        out = np.empty((self.dataset_count(slot), self.cfg.num_classes), np.float32)
        L.check(self.lib.cvt_predict_resident(self.h, slot, out.ctypes.data_as(C.c_void_p)), "cvt_predict_resident")
        return out

    # ---- dropout ----
    def set_dropout(self, mode: str, seed: int = 0):
        L.check(self.lib.cvt_set_dropout(self.h, {"off": 0, "philox": 1, "masks": 2}[mode], seed), "cvt_set_dropout")

    def set_dropout_masks(self, masks: Dict[str, np.ndarray]):
        """masks: {"s{i}.proj"|"s{i}.mlp1"|"s{i}.mlp2"|"s{i}.attn": uint8 [B,Tc,width]} (oracle.make_dropout_masks layout)."""
        sub = {"proj": 0, "mlp1": 1, "mlp2": 2, "attn": 3}
        for k, v in masks.items():
            st, nm = k.split(".")
            site = (int(st[1:]) - 1) * 4 + sub[nm]
            a = np.ascontiguousarray(v, dtype=np.uint8)
            L.check(self.lib.cvt_set_dropout_mask(self.h, site, a.ctypes.data_as(C.c_void_p), a.size), "cvt_set_dropout_mask")

    # ---- taps ----
    def tap(self, name: str) -> np.ndarray:
        n = L.check(self.lib.cvt_get_tap(self.h, name.encode(), None, 0), "cvt_get_tap")
        a = np.empty(n, dtype=np.float32)
        L.check(self.lib.cvt_get_tap(self.h, name.encode(), a.ctypes.data_as(C.c_void_p), n), "cvt_get_tap")
        return a

    def backward_to_tap(self, dout, name: str, count: int) -> np.ndarray:
        d = _f32(dout)
        a = np.empty(count, dtype=np.float32)
        L.check(self.lib.cvt_backward_to_tap(self.h, d.ctypes.data_as(C.c_void_p), name.encode(), a.ctypes.data_as(C.c_void_p), count), "cvt_backward_to_tap")
        return a

    def gradcam(self, img, proc, stage: int):
        """tools/grad_cam_CvT.py:430-457: forward (inference mode), then d sum_b pred[b, 0] / d output of block `stage` (0-based).
        Returns (feature map [B,h,w,D], predictions [B,nc], gradient [B,h,w,D]); the cls row (not part of the block's output map) is dropped."""
        pred = self.predict(img, proc)
        n = pred.shape[0]
        if n > self.cfg.max_batch:
            raise ValueError("gradcam: %d images exceed max_batch %d (the tool calls it with batch_size 1)" % (n, self.cfg.max_batch))
        st = self.cfg.stages[stage]
        D = st.embed_dim
        feat = self.tap("s%d.out" % (stage + 1))
        T = feat.size // (n * D)
        Tc = T + int(st.with_cls_token)
        dout = np.zeros((n, self.cfg.num_classes), np.float32)
        dout[:, 0] = 1.0
        g = self.backward_to_tap(dout, "s%d.x2" % (stage + 1), n * Tc * D).reshape(n, Tc, D)
        if st.with_cls_token:
            g = g[:, 1:]
        # spatial shape of the block's map
        h, w = self.stage_hw(stage)
        return feat.reshape(n, h, w, D), pred, np.ascontiguousarray(g).reshape(n, h, w, D)

    def stage_hw(self, stage: int):
        """Output map size of stage `stage` (0-based): Conv2D 'same' gives ceil(in / stride) per stage."""
        h, w = self.cfg.image_h, self.cfg.image_w
        for s in self.cfg.stages[:stage + 1]:
            h, w = -(-h // s.stride), -(-w // s.stride)
        return h, w

    # ---- data parallel ----
    def comm_unique_id(self) -> bytes:
        buf = (C.c_uint8 * 128)()
        L.check(self.lib.cvt_comm_unique_id(buf), "cvt_comm_unique_id")
        return bytes(buf)

    def comm_init(self, unique_id: bytes, rank: int, world: int):
        buf = (C.c_uint8 * 128).from_buffer_copy(unique_id)
        L.check(self.lib.cvt_comm_init(self.h, buf, rank, world), "cvt_comm_init")

    def comm_peer_export(self) -> bytes:
        """64-byte cudaIpc handle of this rank's workspace (cvt_comm_peer_export)."""
        buf = (C.c_uint8 * 64)()
        L.check(self.lib.cvt_comm_peer_export(self.h, buf), "cvt_comm_peer_export")
        return bytes(buf)

    def comm_peer_open(self, handles: bytes, sync_grads: bool = False):
        """handles: world x 64 bytes in rank order; afterwards the train step ends in the peer-memory reduce + Adam + all-gather kernel."""
        buf = (C.c_uint8 * len(handles)).from_buffer_copy(handles)
        L.check(self.lib.cvt_comm_peer_open(self.h, buf, int(sync_grads)), "cvt_comm_peer_open")

    def gemm_dispatch_counts(self, reset: bool = False):
        """(gemm_tma, gemm_tc, gemm_simt) launches issued by this thread since the last reset (cvt_gemm_dispatch_counts)."""
        out = (C.c_int64 * 3)()
        L.check(self.lib.cvt_gemm_dispatch_counts(out, int(reset)))
        return tuple(int(v) for v in out)

    def comm_peer_close(self):
        L.check(self.lib.cvt_comm_peer_close(self.h), "cvt_comm_peer_close")

    # ---- misc ----
    def synchronize(self):
        L.check(self.lib.cvt_synchronize(self.h))

    def stream(self) -> int:
        return self.lib.cvt_stream(self.h) or 0

    def set_graphs(self, enable: bool):
        L.check(self.lib.cvt_set_graphs(self.h, int(enable)))

    def kernel_launches(self) -> int:
        return self.lib.cvt_kernel_launches(self.h)

    def workspace_bytes(self) -> int:
        return self.lib.cvt_workspace_bytes(self.h)

    def profile_step(self, img_dev, proc_dev, y_dev, batch, training=True, max_entries=256):
        names = (C.c_char_p * max_entries)()
        ms = (C.c_float * max_entries)()
        n = C.c_int()
        L.check(self.lib.cvt_profile_step(self.h, _ptr(img_dev), _ptr(proc_dev), _ptr(y_dev), batch, int(training), max_entries, names, ms, C.byref(n)),
                "cvt_profile_step")
        return [(names[i].decode(), ms[i]) for i in range(n.value)]

    def profile_group_launches(self, max_entries=256):
        names = (C.c_char_p * max_entries)()
        cnt = (C.c_int * max_entries)()
        n = C.c_int()
        L.check(self.lib.cvt_profile_group_launches(self.h, max_entries, names, cnt, C.byref(n)), "cvt_profile_group_launches")
        return [(names[i].decode(), cnt[i]) for i in range(n.value)]
