"""CPU oracle for the CvT hot path of POCHUN-CHEN/Transformer-STM.

TEST INFRASTRUCTURE ONLY.  Nothing under ``transformer-stm_b200/`` (the product)
may import this file; only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` do.

PARITY: PINNED TO THE REFERENCE'S MODEL SOURCE, UNPINNED AGAINST TENSORFLOW'S KERNELS.
The reference's arithmetic lives in TensorFlow 2.x / Keras 2.x, an un-vendored and
un-pinned dependency (`models/CvT(Par).py:1-3`; the README points at a requirements.txt
that does not exist).  TensorFlow cannot be imported or installed in the build container
or on the GPU box and the reference ships neither tests nor golden vectors nor weights
for this path.  Two anchors instead:
  (1) the reference's own call sites, EXECUTED: the model classes and create_cvt_model of
      models/CvT(Par).py, CvT(Img).py, CvT(Par)(many).py, CvT_test(Par)*.py, tools/grad_cam_CvT.py
      and tools/model_plot.py are compiled from /root/reference unmodified and run over
      oracle/mini_keras.py (a float64 define-by-run stand-in for the Keras layers they use);
      predictions and the gradients of one Keras train step (batch-statistics BatchNorm,
      dropout with given keep masks, MSE) agree with this file to 1e-10
      (tests/test_oracle_vs_reference_source.py; vectors committed as tests/golden/refsrc_*.npz
      by tests/golden/make_refsrc_golden.py and checked wherever the tests run,
      tests/test_refsrc_golden.py).  That pins the WIRING - which layer feeds which, variable
      creation order, the dead ConvEmbed LayerNorm, norm1 used twice, attention(q, v, k), the
      cls-token bypass, every head - to what the reference source does when run.
  (2) the arithmetic INSIDE each Keras layer follows the *published* Keras-2 layer
      definitions (restated here and, independently, in oracle/mini_keras.py) and is
      self-checked by (tests/test_oracle.py): fp64 finite-difference gradients, independent
      re-derivations of every layer with torch.nn.functional library ops, and the shape
      goldens of `model_plot_Cls.png` / `model_plot_NoCls.png`.  No number produced by
      TensorFlow itself exists here: at that level parity stays unpinned
      (baseline/run_tf_reference.py closes it on a box that has TensorFlow).

What each function follows (paths relative to /root/reference):
  conv_same ............ ConvEmbed.call        models/CvT(Par).py:194-217  (LN branch dead: :209 vs :304)
  projection ........... Projection.call       models/CvT(Par).py:83-112   (dw_bn | avg | linear)
  conv_attention ....... ConvAttention.call    models/CvT(Par).py:144-191  + keras.layers.MultiHeadAttention
  block ................ ConvTransformerBlock  models/CvT(Par).py:261-289  (norm1 used twice, :272 and :278)
  forward .............. create_cvt_model      models/CvT(Par).py:292-354, models/CvT(Img).py:291-341
  loss / adam .......... model.compile         models/CvT(Par).py:464-466  (Keras Adam, eps 1e-7)
  lr_scheduler ......... lr_scheduler          models/CvT(Par).py:357-360

Tensors are NHWC / [B, T, D] exactly as in the reference.  All math is torch on CPU
in the dtype given (float64 master, float32 for the timed CPU baseline).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

HEAD_DIM_NOTE = "key_dim = embed_dim // num_heads (64 everywhere in the reference spec)"


# --------------------------------------------------------------------------------------
# Spec (mirrors the reference's module-level `spec` dict + switches, models/CvT(Par).py:25-72)
# --------------------------------------------------------------------------------------
@dataclass
class StageSpec:
    embed_dim: int
    patch_size: int
    stride: int
    num_heads: int
    kernel_size: int = 3
    strides: int = 1
    qkv_method: str = "dw_bn"          # dw_bn | avg | linear (identity, models/CvT(Par).py:101) | linear_dense (Dense(dim), tools/grad_cam_CvT.py:103)
    with_cls_token: bool = False


@dataclass
class CvtSpec:
    image_h: int = 128
    image_w: int = 128
    in_chans: int = 1
    stages: List[StageSpec] = field(default_factory=list)
    proc_dim: int = 5                   # 0 -> CvT(Img) variant (no process-parameter branch)
    num_classes: int = 1
    mlp_ratio: int = 4
    drop_rate: float = 0.1
    ln_eps: float = 1e-6
    bn_eps: float = 1e-3
    bn_momentum: float = 0.99
    proc_hidden: int = 256
    # variants of the tools/ scripts; the defaults are models/CvT(Par).py
    head: str = "auto"                  # auto | cls_ln | ln_mean (models/CvT(Par).py:325-337) | cls | mean (tools/model_plot.py:289-292;
                                        #   "mean" is also GlobalAveragePooling2D, tools/grad_cam_CvT.py:336)
    apply_proj_qkv: bool = True         # False: proj_q/k/v are built but never called (tools/grad_cam_CvT.py:146-149 vs :168-201)
    attn_drop_rate: float = 0.0         # > 0: Dropout on the MultiHeadAttention output before self.proj (tools/grad_cam_CvT.py:200)

    @property
    def head_kind(self) -> str:
        if self.head != "auto":
            return self.head
        return "cls_ln" if self.stages[-1].with_cls_token else "ln_mean"


def ref_spec(projection_method: str = "dw_bn", cls_token_switch: bool = True, proc_dim: int = 5,
             num_classes: int = 1, image_h: int = 128, image_w: int = 128,
             dims: Sequence[int] = (64, 128, 256), heads: Sequence[int] = (1, 2, 4)) -> CvtSpec:
    """The reference configuration, models/CvT(Par).py:66-72."""
    st = [
        StageSpec(dims[0], 7, 4, heads[0], 3, 1, projection_method, False),
        StageSpec(dims[1], 3, 2, heads[1], 3, 1, projection_method, False),
        StageSpec(dims[2], 3, 2, heads[2], 3, 1, projection_method, cls_token_switch),
    ]
    return CvtSpec(image_h, image_w, 1, st, proc_dim, num_classes)


def gradcam_spec(cls_token_switch: bool = False, projection_method: str = "dw_bn", **kw) -> CvtSpec:
    """The model tools/grad_cam_CvT.py:315-349 builds: cls token (when switched on) on EVERY stage and dropped after each block,
    proj_q/k/v never applied, attention-output dropout, `linear` = Dense(dim), GlobalAveragePooling2D head without LayerNorm."""
    spec = ref_spec(projection_method, cls_token_switch, **kw)
    for st in spec.stages:
        st.with_cls_token = cls_token_switch
        if projection_method == "linear":
            st.qkv_method = "linear_dense"
    spec.head, spec.apply_proj_qkv, spec.attn_drop_rate = "mean", False, 0.1
    return spec


def wide_spec(**kw) -> CvtSpec:
    """BASELINE.json configs[4]: 340x345 input, dims 64/192/384, heads 1/3/6."""
    return ref_spec(image_h=340, image_w=345, dims=(64, 192, 384), heads=(1, 3, 6), **kw)


def same_pad(n: int, k: int, s: int) -> Tuple[int, int, int]:
    """TF 'SAME': out = ceil(n/s); extra pad goes after (bottom/right)."""
    out = -(-n // s)
    total = max((out - 1) * s + k - n, 0)
    return out, total // 2, total - total // 2


def stage_geometry(spec: CvtSpec) -> List[dict]:
    h, w, c = spec.image_h, spec.image_w, spec.in_chans
    geo = []
    for st in spec.stages:
        ho, pt, pb = same_pad(h, st.patch_size, st.stride)
        wo, pl, pr = same_pad(w, st.patch_size, st.stride)
        t = ho * wo
        geo.append(dict(hin=h, win=w, cin=c, h=ho, w=wo, d=st.embed_dim, t=t,
                        tc=t + (1 if st.with_cls_token else 0), pad=(pt, pb, pl, pr),
                        heads=st.num_heads))
        h, w, c = ho, wo, st.embed_dim
    return geo


# --------------------------------------------------------------------------------------
# Parameters: canonical names, shapes, Keras initialisers
# --------------------------------------------------------------------------------------
def param_specs(spec: CvtSpec) -> List[Tuple[str, Tuple[int, ...], bool]]:
    """[(name, shape, trainable)] in creation order.  Names are structural (not Keras'
    auto-suffixed `dense_3` style); shapes are exactly Keras'."""
    out: List[Tuple[str, Tuple[int, ...], bool]] = []
    cin = spec.in_chans
    for i, st in enumerate(spec.stages, 1):
        d, n = st.embed_dim, st.num_heads
        hd = d // n
        p = f"stage{i}"
        out += [(f"{p}_ConvEmbed/kernel", (st.patch_size, st.patch_size, cin, d), True),
                (f"{p}_ConvEmbed/bias", (d,), True)]
        t = f"{p}_transformer"
        if st.with_cls_token:
            out.append((f"{t}/cls_token", (1, 1, 1, d), True))
        out += [(f"{t}/norm1/gamma", (d,), True), (f"{t}/norm1/beta", (d,), True)]
        if st.qkv_method == "dw_bn":
            for b in "qkv":
                out.append((f"{t}/attn/{b}_proj/depthwise_kernel", (3, 3, d, 1), True))
                out += [(f"{t}/attn/{b}_proj/bn/gamma", (d,), True),
                        (f"{t}/attn/{b}_proj/bn/beta", (d,), True),
                        (f"{t}/attn/{b}_proj/bn/moving_mean", (d,), False),
                        (f"{t}/attn/{b}_proj/bn/moving_variance", (d,), False)]
        if st.qkv_method == "linear_dense":
            for b in "qkv":
                out += [(f"{t}/attn/{b}_proj/dense/kernel", (d, d), True), (f"{t}/attn/{b}_proj/dense/bias", (d,), True)]
        if spec.apply_proj_qkv:
            for b in "qkv":
                out += [(f"{t}/attn/proj_{b}/kernel", (d, d), True), (f"{t}/attn/proj_{b}/bias", (d,), True)]
        for b in ("query", "key", "value"):
            out += [(f"{t}/attn/mha/{b}/kernel", (d, n, hd), True), (f"{t}/attn/mha/{b}/bias", (n, hd), True)]
        out += [(f"{t}/attn/mha/attention_output/kernel", (n, hd, d), True),
                (f"{t}/attn/mha/attention_output/bias", (d,), True),
                (f"{t}/attn/proj/kernel", (d, d), True), (f"{t}/attn/proj/bias", (d,), True),
                (f"{t}/mlp/fc1/kernel", (d, spec.mlp_ratio * d), True), (f"{t}/mlp/fc1/bias", (spec.mlp_ratio * d,), True),
                (f"{t}/mlp/fc2/kernel", (spec.mlp_ratio * d, d), True), (f"{t}/mlp/fc2/bias", (d,), True)]
        cin = d
    d = spec.stages[-1].embed_dim
    if spec.head_kind in ("cls_ln", "ln_mean"):
        out += [("head_norm/gamma", (d,), True), ("head_norm/beta", (d,), True)]
    feat = d
    if spec.proc_dim > 0:
        out += [("Proc_Dense_1/kernel", (spec.proc_dim, spec.proc_hidden), True), ("Proc_Dense_1/bias", (spec.proc_hidden,), True),
                ("Proc_Dense_2/kernel", (spec.proc_hidden, spec.proc_hidden), True), ("Proc_Dense_2/bias", (spec.proc_hidden,), True)]
        feat += spec.proc_hidden
    out += [("Final_Dense/kernel", (feat, spec.num_classes), True), ("Final_Dense/bias", (spec.num_classes,), True)]
    return out


def _glorot_fans(name: str, shape: Tuple[int, ...]) -> Tuple[int, int]:
    """Keras `_compute_fans`: receptive = prod(shape[:-2]), fan_in = shape[-2]*rec, fan_out = shape[-1]*rec."""
    if len(shape) == 2:
        return shape[0], shape[1]
    rec = int(np.prod(shape[:-2]))
    return shape[-2] * rec, shape[-1] * rec


def init_params(spec: CvtSpec, seed: int = 1234, mode: str = "parity") -> Dict[str, np.ndarray]:
    """mode='keras': the Keras initialisers (glorot_uniform kernels, zero biases/cls, LN/BN identity).
    mode='parity': same kernels, but every bias / gamma / beta / moving stat / cls token is made
    non-trivial so that no path is multiplied by 0 or 1 by accident."""
    rng = np.random.default_rng(seed)
    p: Dict[str, np.ndarray] = {}
    for name, shape, _ in param_specs(spec):
        leaf = name.rsplit("/", 1)[-1]
        if leaf in ("kernel", "depthwise_kernel"):
            fi, fo = _glorot_fans(name, shape)
            lim = math.sqrt(6.0 / (fi + fo))
            v = rng.uniform(-lim, lim, size=shape)
        elif leaf == "gamma":
            v = np.ones(shape) if mode == "keras" else rng.uniform(0.5, 1.5, size=shape)
        elif leaf == "moving_variance":
            v = np.ones(shape) if mode == "keras" else rng.uniform(0.5, 1.5, size=shape)
        elif leaf == "moving_mean":
            v = np.zeros(shape) if mode == "keras" else rng.normal(0, 0.1, size=shape)
        else:  # bias, beta, cls_token
            v = np.zeros(shape) if mode == "keras" else rng.normal(0, 0.02, size=shape)
        p[name] = v.astype(np.float64)
    return p


def synthetic_batch(spec: CvtSpec, batch: int, seed: int = 0):
    """SURVEY 8d synthetic inputs: images U[0,1), proc N(0,1), targets 1000 + 100 N(0,1)."""
    img = np.random.default_rng(seed).random((batch, spec.image_h, spec.image_w, spec.in_chans))
    proc = np.random.default_rng(seed + 1).standard_normal((batch, max(spec.proc_dim, 1)))[:, :spec.proc_dim]
    y = 1000.0 + 100.0 * np.random.default_rng(seed + 2).standard_normal((batch, spec.num_classes))
    return img.astype(np.float32), proc.astype(np.float32), y.astype(np.float32)


# --------------------------------------------------------------------------------------
# Dropout masks: Philox4x32-10 stream shared bit-for-bit with the CUDA kernels
# --------------------------------------------------------------------------------------
_PH_M0, _PH_M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_PH_W0, _PH_W1 = np.uint32(0x9E3779B9), np.uint32(0xBB67AE85)


def philox4x32(counter: np.ndarray, key: Tuple[int, int]) -> np.ndarray:
    """counter: uint32 [N,4]; returns uint32 [N,4] (10 rounds)."""
    c = counter.astype(np.uint32).copy()
    k0, k1 = np.uint32(key[0] & 0xFFFFFFFF), np.uint32(key[1] & 0xFFFFFFFF)
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = c[:, 0].astype(np.uint64) * _PH_M0
            p1 = c[:, 2].astype(np.uint64) * _PH_M1
            hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), p0.astype(np.uint32)
            hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), p1.astype(np.uint32)
            c = np.stack([hi1 ^ c[:, 1] ^ k0, lo1, hi0 ^ c[:, 3] ^ k1, lo0], axis=1)
            k0 = np.uint32(k0 + _PH_W0)
            k1 = np.uint32(k1 + _PH_W1)
    return c


def dropout_mask(n: int, rate: float, seed: int, step: int, site: int) -> np.ndarray:
    """keep-mask (uint8 0/1) for n elements; element i uses word i%4 of philox(counter=(i//4, site, step, 0),
    key=(seed_lo, seed_hi)); keep iff word >= rate * 2^32.  Same function as csrc/elementwise.cu:philox_mask."""
    nq = (n + 3) // 4
    ctr = np.zeros((nq, 4), dtype=np.uint32)
    ctr[:, 0] = np.arange(nq, dtype=np.uint64).astype(np.uint32)
    ctr[:, 1] = np.uint32(site)
    ctr[:, 2] = np.uint32(step & 0xFFFFFFFF)
    r = philox4x32(ctr, (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)).reshape(-1)[:n]
    thr = np.uint32(min(int(rate * 4294967296.0), 0xFFFFFFFF))
    return (r >= thr).astype(np.uint8)


def make_dropout_masks(spec: CvtSpec, batch: int, seed: int, step: int) -> Dict[str, np.ndarray]:
    """Masks for the three active dropout sites of every stage (models/CvT(Par).py:189,255,257).
    site id = stage_index*4 + {0: proj, 1: mlp hidden, 2: mlp out}."""
    masks = {}
    for i, g in enumerate(stage_geometry(spec)):
        d, tc = g["d"], g["tc"]
        for j, (nm, width) in enumerate((("proj", d), ("mlp1", spec.mlp_ratio * d), ("mlp2", d))):
            masks[f"s{i + 1}.{nm}"] = dropout_mask(batch * tc * width, spec.drop_rate, seed, step, i * 4 + j).reshape(batch, tc, width)
        if spec.attn_drop_rate > 0.0:                                # site 3: attention-output dropout of the grad-cam variant
            masks[f"s{i + 1}.attn"] = dropout_mask(batch * tc * d, spec.attn_drop_rate, seed, step, i * 4 + 3).reshape(batch, tc, d)
    return masks


# --------------------------------------------------------------------------------------
# Layers (Keras-2 semantics)
# --------------------------------------------------------------------------------------
def conv_same(x: torch.Tensor, kernel: torch.Tensor, bias: torch.Tensor, stride: int) -> torch.Tensor:
    """keras Conv2D(padding='same'), NHWC in/out, kernel HWIO.  models/CvT(Par).py:203-208,212."""
    kh, kw = kernel.shape[0], kernel.shape[1]
    _, pt, pb = same_pad(x.shape[1], kh, stride)
    _, pl, pr = same_pad(x.shape[2], kw, stride)
    xn = F.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb))
    y = F.conv2d(xn, kernel.permute(3, 2, 0, 1), bias, stride=stride)
    return y.permute(0, 2, 3, 1)


def layer_norm(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float) -> torch.Tensor:
    mu = x.mean(-1, keepdim=True)
    var = ((x - mu) ** 2).mean(-1, keepdim=True)
    return (x - mu) / torch.sqrt(var + eps) * gamma + beta


def depthwise3x3(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    """DepthwiseConv2D(3, 1, 'same', use_bias=False); x [B,H,W,C], w [3,3,C,1]. models/CvT(Par).py:93,105."""
    b, h, wd, c = x.shape
    xp = F.pad(x, (0, 0, 1, 1, 1, 1))
    y = torch.zeros_like(x)
    for di in range(3):
        for dj in range(3):
            y = y + xp[:, di:di + h, dj:dj + wd, :] * w[di, dj, :, 0]
    return y


def avgpool3x3_same(x: torch.Tensor) -> torch.Tensor:
    """AveragePooling2D(3, 1, 'same'): TF divides by the number of in-bounds taps. models/CvT(Par).py:96,108."""
    b, h, wd, c = x.shape
    xp = F.pad(x, (0, 0, 1, 1, 1, 1))
    ones = F.pad(torch.ones(1, h, wd, 1, dtype=x.dtype), (0, 0, 1, 1, 1, 1))
    s = torch.zeros_like(x)
    cnt = torch.zeros(1, h, wd, 1, dtype=x.dtype)
    for di in range(3):
        for dj in range(3):
            s = s + xp[:, di:di + h, dj:dj + wd, :]
            cnt = cnt + ones[:, di:di + h, dj:dj + wd, :]
    return s / cnt


def batch_norm(x: torch.Tensor, gamma, beta, mm, mv, training: bool, eps: float):
    """keras BatchNormalization(axis=-1); returns (y, batch_mean, batch_var_biased).  models/CvT(Par).py:94,106."""
    if training:
        mu = x.mean(dim=(0, 1, 2))
        var = ((x - mu) ** 2).mean(dim=(0, 1, 2))
    else:
        mu, var = mm, mv
    y = (x - mu) / torch.sqrt(var + eps) * gamma + beta
    return y, mu, var


def gelu_erf(x: torch.Tensor) -> torch.Tensor:
    return 0.5 * x * (1.0 + torch.erf(x / math.sqrt(2.0)))


def _apply_dropout(x, mask, rate, training):
    if not training or mask is None or rate <= 0.0:
        return x
    return x * mask.to(x.dtype) / (1.0 - rate)


# --------------------------------------------------------------------------------------
# Forward
# --------------------------------------------------------------------------------------
def _tp(params: Dict[str, np.ndarray], dtype, requires_grad: bool) -> Dict[str, torch.Tensor]:
    out = {}
    trainable = None
    for k, v in params.items():
        t = torch.tensor(np.asarray(v), dtype=dtype)
        leaf = k.rsplit("/", 1)[-1]
        if requires_grad and leaf not in ("moving_mean", "moving_variance"):
            t.requires_grad_(True)
        out[k] = t
    return out


def forward_torch(spec: CvtSpec, P: Dict[str, torch.Tensor], img: torch.Tensor, proc: Optional[torch.Tensor],
                  training: bool, masks: Optional[Dict[str, torch.Tensor]] = None, keep_taps: bool = True):
    """Returns (pred [B,nc], taps dict, bn_batch_stats dict).  img [B,H,W,C]."""
    taps: Dict[str, torch.Tensor] = {}
    bn_stats: Dict[str, Tuple[torch.Tensor, torch.Tensor, int]] = {}
    rate = spec.drop_rate
    x = img
    bsz = img.shape[0]
    cls_out = None
    for i, st in enumerate(spec.stages, 1):
        d, nh = st.embed_dim, st.num_heads
        hd = d // nh
        pfx, t = f"stage{i}", f"stage{i}_transformer"
        # ConvEmbed (LayerNorm branch is dead code: models/CvT(Par).py:209 vs :304)
        e = conv_same(x, P[f"{pfx}_ConvEmbed/kernel"], P[f"{pfx}_ConvEmbed/bias"], st.stride)
        _, h, w, _ = e.shape
        tok = e.reshape(bsz, h * w, d)
        if st.with_cls_token:                                        # :264-268
            cls = P[f"{t}/cls_token"].reshape(1, 1, d).expand(bsz, 1, d)
            x0 = torch.cat([cls, tok], dim=1)
        else:
            x0 = tok
        g1, b1 = P[f"{t}/norm1/gamma"], P[f"{t}/norm1/beta"]
        xn = layer_norm(x0, g1, b1, spec.ln_eps)                     # :272
        # ConvAttention.call :144-191
        if st.with_cls_token:
            cls_n, sp = xn[:, :1], xn[:, 1:]
        else:
            cls_n, sp = None, xn
        sp = sp.reshape(bsz, h, w, d)
        qkv0, convs = [], []
        for bi, b in enumerate("qkv"):
            method = st.qkv_method
            if method == "avg" and b == "q":
                method = "linear"                                    # :127
            if method == "dw_bn":
                c = depthwise3x3(sp, P[f"{t}/attn/{b}_proj/depthwise_kernel"])
                y, mu, var = batch_norm(c, P[f"{t}/attn/{b}_proj/bn/gamma"], P[f"{t}/attn/{b}_proj/bn/beta"],
                                        P[f"{t}/attn/{b}_proj/bn/moving_mean"], P[f"{t}/attn/{b}_proj/bn/moving_variance"],
                                        training, spec.bn_eps)
                bn_stats[f"{t}/attn/{b}_proj/bn"] = (mu, var, bsz * h * w)
                convs.append(c.reshape(bsz, h * w, d))
            elif method == "avg":
                y = avgpool3x3_same(sp)
            elif method == "linear_dense":                           # tools/grad_cam_CvT.py:103,117
                y = sp @ P[f"{t}/attn/{b}_proj/dense/kernel"] + P[f"{t}/attn/{b}_proj/dense/bias"]
            else:
                y = sp
            y = y.reshape(bsz, h * w, d)
            if cls_n is not None:
                y = torch.cat([cls_n, y], dim=1)                     # :168-170 (cls bypasses conv/BN)
            qkv0.append(y)
        if spec.apply_proj_qkv:
            qkv1 = [qkv0[j] @ P[f"{t}/attn/proj_{b}/kernel"] + P[f"{t}/attn/proj_{b}/bias"] for j, b in enumerate("qkv")]  # :180-182
        else:
            qkv1 = list(qkv0)                                        # tools/grad_cam_CvT.py:199 (q, k, v go straight into the attention)
        # keras MultiHeadAttention(num_heads, key_dim), called (query=q, value=v, key=k)  :185
        def heads(z, nm):
            return torch.einsum("btd,dnh->btnh", z, P[f"{t}/attn/mha/{nm}/kernel"]) + P[f"{t}/attn/mha/{nm}/bias"]
        Q = heads(qkv1[0], "query") * (1.0 / math.sqrt(hd))
        K = heads(qkv1[1], "key")
        V = heads(qkv1[2], "value")
        S = torch.einsum("btnh,bsnh->bnts", Q, K)
        Pm = torch.softmax(S, dim=-1)
        O = torch.einsum("bnts,bsnh->btnh", Pm, V)
        a1 = torch.einsum("btnh,nhd->btd", O, P[f"{t}/attn/mha/attention_output/kernel"]) + P[f"{t}/attn/mha/attention_output/bias"]
        if spec.attn_drop_rate > 0.0:                                                         # tools/grad_cam_CvT.py:200
            a1 = _apply_dropout(a1, masks.get(f"s{i}.attn") if masks else None, spec.attn_drop_rate, training)
        a2 = a1 @ P[f"{t}/attn/proj/kernel"] + P[f"{t}/attn/proj/bias"]                      # :188
        a2 = _apply_dropout(a2, masks.get(f"s{i}.proj") if masks else None, rate, training)   # :189
        x1 = a2 + x0                                                                          # :274
        y = layer_norm(x1, g1, b1, spec.ln_eps)                                               # :278 (same norm1)
        hpre = y @ P[f"{t}/mlp/fc1/kernel"] + P[f"{t}/mlp/fc1/bias"]
        hact = _apply_dropout(gelu_erf(hpre), masks.get(f"s{i}.mlp1") if masks else None, rate, training)
        m = hact @ P[f"{t}/mlp/fc2/kernel"] + P[f"{t}/mlp/fc2/bias"]
        m = _apply_dropout(m, masks.get(f"s{i}.mlp2") if masks else None, rate, training)
        x2 = x1 + m                                                                           # :281
        if st.with_cls_token:
            cls_out, sp_out = x2[:, :1], x2[:, 1:]
        else:
            cls_out, sp_out = None, x2
        x = sp_out.reshape(bsz, h, w, d)
        if keep_taps:
            s = f"s{i}."
            taps.update({s + "embed": e, s + "x0": x0, s + "xn": xn, s + "qkv0": torch.stack(qkv0),
                         s + "qkv1": torch.stack(qkv1),
                         s + "qkv2": torch.stack([Q.reshape(bsz, -1, d), K.reshape(bsz, -1, d), V.reshape(bsz, -1, d)]),
                         s + "attn_o": O.reshape(bsz, -1, d), s + "a1": a1, s + "x1": x1, s + "y": y,
                         s + "hpre": hpre, s + "hact": hact, s + "x2": x2, s + "out": x})
            if convs:
                taps[s + "conv_qkv"] = torch.stack(convs)
    # head  models/CvT(Par).py:325-350
    hk = spec.head_kind
    if hk in ("cls_ln", "cls") and cls_out is None:
        raise ValueError("head '%s' needs a cls token on the last stage" % hk)
    if hk == "cls_ln":
        feat = layer_norm(cls_out, P["head_norm/gamma"], P["head_norm/beta"], spec.ln_eps).squeeze(1)
    elif hk == "ln_mean":
        feat = layer_norm(x.reshape(bsz, -1, x.shape[-1]), P["head_norm/gamma"], P["head_norm/beta"], spec.ln_eps).mean(dim=1)
    elif hk == "cls":                                                # tools/model_plot.py:289-290
        feat = cls_out.squeeze(1)
    elif hk == "mean":                                               # tools/model_plot.py:291-292, tools/grad_cam_CvT.py:336 (GAP)
        feat = x.reshape(bsz, -1, x.shape[-1]).mean(dim=1)
    else:
        raise ValueError("unknown head kind %r" % hk)
    if spec.proc_dim > 0:
        p1 = torch.relu(proc @ P["Proc_Dense_1/kernel"] + P["Proc_Dense_1/bias"])
        p2 = torch.relu(p1 @ P["Proc_Dense_2/kernel"] + P["Proc_Dense_2/bias"])
        taps["head.proc1"], taps["head.proc2"] = p1, p2
        feat = torch.cat([feat, p2], dim=1)
    taps["head.feat"] = feat
    pred = feat @ P["Final_Dense/kernel"] + P["Final_Dense/bias"]
    taps["pred"] = pred
    return pred, taps, bn_stats


def _to_t(a, dtype):
    return None if a is None else torch.tensor(np.asarray(a), dtype=dtype)


def forward(spec: CvtSpec, params: Dict[str, np.ndarray], img, proc=None, training: bool = False,
            masks: Optional[Dict[str, np.ndarray]] = None, dtype=torch.float64):
    """numpy in / numpy out.  Returns (pred, taps)."""
    P = _tp(params, dtype, False)
    with torch.no_grad():
        pred, taps, _ = forward_torch(spec, P, _to_t(img, dtype), _to_t(proc, dtype) if spec.proc_dim > 0 else None, training,
                                      {k: _to_t(v, dtype) for k, v in masks.items()} if masks else None)
    return pred.numpy(), {k: v.numpy() for k, v in taps.items()}


def loss_and_grads(spec: CvtSpec, params: Dict[str, np.ndarray], img, proc, y, masks=None, dtype=torch.float64,
                   training: bool = True, tap_grads: Sequence[str] = ()):
    """One Keras train_step up to the gradients (models/CvT(Par).py:464-466,475):
    loss = mean((pred - y)^2), mae = mean(|pred - y|); grads of every trainable param via autograd
    (the reference gets them from TF autodiff).  Also returns the updated BN moving statistics."""
    P = _tp(params, dtype, True)
    pred, taps, bn_stats = forward_torch(spec, P, _to_t(img, dtype), _to_t(proc, dtype) if spec.proc_dim > 0 else None,
                                         training, {k: _to_t(v, dtype) for k, v in masks.items()} if masks else None)
    yt = _to_t(y, dtype).reshape(pred.shape)
    for nm in tap_grads:
        taps[nm].retain_grad()
    loss = ((pred - yt) ** 2).mean()
    mae = (pred - yt).abs().mean()
    loss.backward()
    grads = {k: (v.grad.numpy() if v.grad is not None else np.zeros(tuple(v.shape))) for k, v in P.items() if v.requires_grad}
    new_moving = {}
    if training:
        mom = spec.bn_momentum
        for name, (mu, var, n) in bn_stats.items():
            # TF fused batch-norm feeds the *unbiased* batch variance into the moving average
            unb = var.detach() * (n / max(n - 1, 1))
            new_moving[name + "/moving_mean"] = (mom * P[name + "/moving_mean"] + (1 - mom) * mu.detach()).numpy()
            new_moving[name + "/moving_variance"] = (mom * P[name + "/moving_variance"] + (1 - mom) * unb).numpy()
    out = dict(loss=float(loss.detach()), mae=float(mae.detach()), pred=pred.detach().numpy(), grads=grads, new_moving=new_moving,
               taps={k: v.detach().numpy() for k, v in taps.items()},
               tap_grads={nm: taps[nm].grad.numpy() for nm in tap_grads})
    return out


def adam_step(params, grads, m, v, t: int, lr: float, b1=0.9, b2=0.999, eps=1e-7):
    """Keras Adam (non-amsgrad): theta -= lr*sqrt(1-b2^t)/(1-b1^t) * m/(sqrt(v)+eps); t starts at 1."""
    alpha = lr * math.sqrt(1.0 - b2 ** t) / (1.0 - b1 ** t)
    for k, g in grads.items():
        m[k] = b1 * m[k] + (1 - b1) * g
        v[k] = b2 * v[k] + (1 - b2) * g * g
        params[k] = params[k] - alpha * m[k] / (np.sqrt(v[k]) + eps)
    return params, m, v


def lr_scheduler(epoch: int, lr: float) -> float:
    """models/CvT(Par).py:357-360."""
    if epoch > 0 and epoch % 50 == 0:
        return lr * 0.8
    return lr


def train_step(spec, params, opt_state, img, proc, y, lr, masks=None, dtype=torch.float64):
    """Full Keras train_step: grads -> Adam -> BN moving update.  opt_state = dict(m, v, t)."""
    r = loss_and_grads(spec, params, img, proc, y, masks, dtype)
    opt_state["t"] += 1
    tr = {k: params[k] for k in r["grads"]}
    tr, opt_state["m"], opt_state["v"] = adam_step(tr, r["grads"], opt_state["m"], opt_state["v"], opt_state["t"], lr)
    params.update(tr)
    params.update(r["new_moving"])
    return r


def new_opt_state(params):
    tr = [k for k in params if k.rsplit("/", 1)[-1] not in ("moving_mean", "moving_variance")]
    return dict(m={k: np.zeros_like(params[k]) for k in tr}, v={k: np.zeros_like(params[k]) for k in tr}, t=0)


def gradcam(spec, params, img, proc, layer: str = "s3.out", dtype=torch.float64):
    """tools/grad_cam_CvT.py:446-472: grads = d pred[:,0] / d feature map; weights = mean over (B,H,W);
    heatmap = relu(sum_c w_c * A) / max.  Returns (heatmap [B,h,w], feature map, grads)."""
    P = _tp(params, dtype, False)
    pred, taps, _ = _forward_with_tap_grad(spec, P, _to_t(img, dtype), _to_t(proc, dtype) if spec.proc_dim > 0 else None, layer)
    feat = taps[layer]
    pred[:, 0].sum().backward()
    g = feat.grad
    wts = g.mean(dim=(0, 1, 2))
    heat = torch.relu((feat.detach() * wts).sum(-1))
    heat = heat / heat.amax(dim=(1, 2), keepdim=True).clamp_min(1e-30)
    return heat.numpy(), feat.detach().numpy(), g.numpy()


def _forward_with_tap_grad(spec, P, img, proc, layer):
    img = img.clone().requires_grad_(True)     # makes every tap a non-leaf with grad_fn
    pred, taps, bn = forward_torch(spec, P, img, proc, False, None)
    taps[layer].retain_grad()
    return pred, taps, bn
