"""TEST INFRASTRUCTURE - never imported, linked or executed by the product (transformer-stm_b200/), bench.py's timed arms or smoke().

A define-by-run stand-in for the slice of TensorFlow / Keras 2 that the reference's model code touches (SURVEY 8b census:
tf.shape / reshape / split / concat / tile / squeeze / reduce_mean / cast / expand_dims / nn.gelu; layers.Layer / Dense / LayerNormalization /
Dropout / Conv2D / DepthwiseConv2D / BatchNormalization / AveragePooling2D / MultiHeadAttention / GlobalAveragePooling1D/2D / concatenate;
keras.Input / Model / Sequential), on float64 torch-CPU tensors.  Its purpose is to EXECUTE THE REFERENCE'S OWN MODEL SOURCE -
the classes Projection, ConvAttention, ConvEmbed, ConvTransformerBlock and create_cvt_model of models/CvT(Par).py:83-354 (and of CvT(Img).py,
tools/grad_cam_CvT.py, tools/model_plot.py), compiled from /root/reference as they lie there - so that the wiring of oracle/cvt_oracle.py
(which layer feeds which, the dead ConvEmbed LayerNorm, norm1 used twice, attention(q, v, k), the cls-token bypass, the heads) is
checked against what the reference source does when run, not against a reading of it (tests/test_oracle_vs_reference_source.py,
tests/golden/make_refsrc_golden.py).  TensorFlow itself is not available here: what stays restated is the arithmetic INSIDE each Keras
layer, written below from the Keras-2 layer definitions - independently of oracle/cvt_oracle.py, which restates the same definitions, is
checked layer by layer against torch.nn.functional in tests/test_oracle.py, and agrees with this file to 1e-10:
  Dense                 y = x W + b                                                  keras/layers/core/dense.py
  LayerNormalization    biased variance over the last axis, epsilon default 1e-3    keras/layers/normalization/layer_normalization.py
  BatchNormalization    epsilon 1e-3, momentum 0.99; batch statistics (biased variance) over N, H, W when training, moving ones otherwise
  Conv2D / DepthwiseConv2D   NHWC, HWIO / HWC1 kernels, TF 'same' padding (extra pixel at the bottom / right)
  AveragePooling2D      'same': the divisor counts valid pixels only
  MultiHeadAttention    the four einsums of keras/layers/attention/multi_head_attention.py, query scaled by 1/sqrt(key_dim) before QK^T
  Dropout               x * keep / (1 - rate) when training (keep masks are handed in: the reference's RNG stream cannot be reproduced)
Weights are not initialised here: every variable a layer creates - in the order Keras would create it, i.e. add_weight calls in __init__
and lazily-built layers at their first call - takes the next entry of the canonical parameter list handed to `session`, after a shape check."""
import ast
import math
import types
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

_S = None          # the running session


class session:
    """with session(weights, inputs, training, dropout_masks) as s: model = create_cvt_model(...)"""

    def __init__(self, weights: Sequence[Tuple[str, torch.Tensor]], inputs: Dict[str, torch.Tensor], training: bool = False,
                 dropout_masks: Optional[Sequence[torch.Tensor]] = None):
        self.weights, self.inputs, self.training = list(weights), inputs, training
        self.masks = list(dropout_masks) if dropout_masks is not None else None
        self.pos, self.mask_pos, self.taken, self.bn_batch_stats = 0, 0, [], []
        self.captured = {}                                   # outputs of named layers (what `model.get_layer(name).output` stands for)

    def __enter__(self):
        global _S
        assert _S is None, "nested sessions"
        _S = self
        return self

    def __exit__(self, *exc):
        global _S
        _S = None
        return False

    def next_weight(self, shape, who):
        if self.pos >= len(self.weights):
            raise AssertionError("%s asks for a variable of shape %s after the canonical list is exhausted" % (who, tuple(shape)))
        name, t = self.weights[self.pos]
        if tuple(t.shape) != tuple(int(s) for s in shape):
            raise AssertionError("%s creates a variable of shape %s where the canonical list has %s %s" % (who, tuple(shape), name, tuple(t.shape)))
        self.pos += 1
        self.taken.append(name)
        return t

    def assert_all_weights_used(self):
        assert self.pos == len(self.weights), "the reference model never created: %s" % [n for n, _ in self.weights[self.pos:]]
        if self.masks is not None:
            assert self.mask_pos == len(self.masks), "dropout sites called %d, masks given %d" % (self.mask_pos, len(self.masks))


# ---- keras.layers -------------------------------------------------------------------------------------------------------------
class Layer:
    def __init__(self, name=None, **kw):
        self.name = name

    def add_weight(self, shape=None, initializer=None, trainable=True, name=None, **kw):
        return _S.next_weight(shape, "%s.add_weight(%s)" % (type(self).__name__, name))

    def __call__(self, *args, **kw):
        out = self.call(*args, **kw)
        if self.name is not None and _S is not None:
            _S.captured[self.name] = out
        return out

    def get_config(self):
        return {}


def _activation(act):
    if act is None or act == "linear":
        return lambda x: x
    if act == "relu":
        return torch.relu
    if act == "gelu":
        return nn.gelu
    assert callable(act), act
    return act


class Dense(Layer):
    def __init__(self, units, activation=None, use_bias=True, name=None, **kw):
        super().__init__(name=name)
        self.units, self.act, self.use_bias, self.kernel = int(units), _activation(activation), use_bias, None

    def call(self, x):
        if self.kernel is None:
            self.kernel = _S.next_weight((x.shape[-1], self.units), "Dense(%s) kernel" % self.name)
            self.bias = _S.next_weight((self.units,), "Dense(%s) bias" % self.name) if self.use_bias else None
        y = x @ self.kernel
        return self.act(y + self.bias if self.bias is not None else y)


class LayerNormalization(Layer):
    def __init__(self, axis=-1, epsilon=1e-3, name=None, **kw):
        super().__init__(name=name)
        assert axis == -1
        self.eps, self.gamma = float(epsilon), None

    def call(self, x):
        if self.gamma is None:
            self.gamma = _S.next_weight((x.shape[-1],), "LayerNormalization gamma")
            self.beta = _S.next_weight((x.shape[-1],), "LayerNormalization beta")
        mu = x.mean(-1, keepdim=True)
        var = ((x - mu) ** 2).mean(-1, keepdim=True)
        return (x - mu) / torch.sqrt(var + self.eps) * self.gamma + self.beta


class Dropout(Layer):
    def __init__(self, rate, name=None, **kw):
        super().__init__(name=name)
        self.rate = float(rate)

    def call(self, x, training=None):
        if not _S.training or _S.masks is None or self.rate == 0.0:
            return x
        keep = _S.masks[_S.mask_pos]
        _S.mask_pos += 1
        return x * keep.reshape(x.shape) / (1.0 - self.rate)


def _same_pad(n, k, s):
    total = max((math.ceil(n / s) - 1) * s + k - n, 0)
    return total // 2, total - total // 2


def _pad_same_nchw(x, kh, kw, sh, sw):
    t, b = _same_pad(x.shape[2], kh, sh)
    l, r = _same_pad(x.shape[3], kw, sw)
    return F.pad(x, (l, r, t, b))


def _pair(v):
    return (int(v), int(v)) if not isinstance(v, (tuple, list)) else (int(v[0]), int(v[1]))


class Conv2D(Layer):
    def __init__(self, filters, kernel_size, strides=1, padding="valid", use_bias=True, name=None, **kw):
        super().__init__(name=name)
        self.filters, self.k, self.s, self.padding, self.use_bias, self.kernel = int(filters), _pair(kernel_size), _pair(strides), padding, use_bias, None

    def call(self, x):
        if self.kernel is None:
            self.kernel = _S.next_weight((self.k[0], self.k[1], x.shape[-1], self.filters), "Conv2D kernel")
            self.bias = _S.next_weight((self.filters,), "Conv2D bias") if self.use_bias else None
        h = x.permute(0, 3, 1, 2)
        if self.padding == "same":
            h = _pad_same_nchw(h, self.k[0], self.k[1], self.s[0], self.s[1])
        y = F.conv2d(h, self.kernel.permute(3, 2, 0, 1), self.bias, stride=self.s)
        return y.permute(0, 2, 3, 1)


class DepthwiseConv2D(Layer):
    def __init__(self, kernel_size, strides=1, padding="valid", use_bias=True, depth_multiplier=1, name=None, **kw):
        super().__init__(name=name)
        assert depth_multiplier == 1
        self.k, self.s, self.padding, self.use_bias, self.kernel = _pair(kernel_size), _pair(strides), padding, use_bias, None

    def call(self, x):
        c = x.shape[-1]
        if self.kernel is None:
            self.kernel = _S.next_weight((self.k[0], self.k[1], c, 1), "DepthwiseConv2D depthwise_kernel")
            self.bias = _S.next_weight((c,), "DepthwiseConv2D bias") if self.use_bias else None
        h = x.permute(0, 3, 1, 2)
        if self.padding == "same":
            h = _pad_same_nchw(h, self.k[0], self.k[1], self.s[0], self.s[1])
        y = F.conv2d(h, self.kernel.permute(2, 3, 0, 1), self.bias, stride=self.s, groups=c)
        return y.permute(0, 2, 3, 1)


class BatchNormalization(Layer):
    def __init__(self, axis=-1, momentum=0.99, epsilon=1e-3, name=None, **kw):
        super().__init__(name=name)
        assert axis == -1
        self.momentum, self.eps, self.gamma = float(momentum), float(epsilon), None

    def call(self, x, training=None):
        c = x.shape[-1]
        if self.gamma is None:
            self.gamma = _S.next_weight((c,), "BatchNormalization gamma")
            self.beta = _S.next_weight((c,), "BatchNormalization beta")
            self.moving_mean = _S.next_weight((c,), "BatchNormalization moving_mean")
            self.moving_variance = _S.next_weight((c,), "BatchNormalization moving_variance")
        if _S.training:
            dims = tuple(range(x.dim() - 1))
            mu = x.mean(dims)
            var = ((x - mu) ** 2).mean(dims)
            _S.bn_batch_stats.append((mu.detach(), var.detach(), x.numel() // c))
        else:
            mu, var = self.moving_mean, self.moving_variance
        return (x - mu) / torch.sqrt(var + self.eps) * self.gamma + self.beta


class AveragePooling2D(Layer):
    def __init__(self, pool_size=2, strides=None, padding="valid", name=None, **kw):
        super().__init__(name=name)
        self.k = _pair(pool_size)
        self.s = _pair(strides if strides is not None else pool_size)
        self.padding = padding

    def call(self, x):
        h = x.permute(0, 3, 1, 2)
        if self.padding == "same":
            ones = _pad_same_nchw(torch.ones_like(h[:1, :1]), self.k[0], self.k[1], self.s[0], self.s[1])
            h = _pad_same_nchw(h, self.k[0], self.k[1], self.s[0], self.s[1])
            y = F.avg_pool2d(h, self.k, self.s) / F.avg_pool2d(ones, self.k, self.s)      # divisor = number of valid pixels under the window
        else:
            y = F.avg_pool2d(h, self.k, self.s)
        return y.permute(0, 2, 3, 1)


class GlobalAveragePooling2D(Layer):
    def call(self, x):
        return x.mean((1, 2))


class GlobalAveragePooling1D(Layer):
    def call(self, x):
        return x.mean(1)


class MultiHeadAttention(Layer):
    def __init__(self, num_heads, key_dim, value_dim=None, dropout=0.0, use_bias=True, name=None, **kw):
        super().__init__(name=name)
        assert dropout == 0.0 and use_bias
        self.n, self.h, self.hv, self.wq = int(num_heads), int(key_dim), int(value_dim or key_dim), None

    def call(self, query, value, key=None, attention_mask=None, return_attention_scores=False, training=None):
        assert attention_mask is None and not return_attention_scores
        if key is None:
            key = value
        if self.wq is None:                                  # _build_from_signature: query, key, value, attention_output
            who = "MultiHeadAttention"
            self.wq = _S.next_weight((query.shape[-1], self.n, self.h), who + " query kernel"); self.bq = _S.next_weight((self.n, self.h), who + " query bias")
            self.wk = _S.next_weight((key.shape[-1], self.n, self.h), who + " key kernel"); self.bk = _S.next_weight((self.n, self.h), who + " key bias")
            self.wv = _S.next_weight((value.shape[-1], self.n, self.hv), who + " value kernel"); self.bv = _S.next_weight((self.n, self.hv), who + " value bias")
            self.wo = _S.next_weight((self.n, self.hv, query.shape[-1]), who + " output kernel"); self.bo = _S.next_weight((query.shape[-1],), who + " output bias")
        q = torch.einsum("abc,cde->abde", query, self.wq) + self.bq
        k = torch.einsum("abc,cde->abde", key, self.wk) + self.bk
        v = torch.einsum("abc,cde->abde", value, self.wv) + self.bv
        q = q * (1.0 / math.sqrt(float(self.h)))
        scores = torch.softmax(torch.einsum("aecd,abcd->acbe", k, q), dim=-1)
        o = torch.einsum("acbe,aecd->abcd", scores, v)
        return torch.einsum("abcd,cde->abe", o, self.wo) + self.bo


def concatenate(tensors, axis=-1):
    return torch.cat(list(tensors), dim=axis)


layers = types.SimpleNamespace(Layer=Layer, Dense=Dense, LayerNormalization=LayerNormalization, Dropout=Dropout, Conv2D=Conv2D,
                               DepthwiseConv2D=DepthwiseConv2D, BatchNormalization=BatchNormalization, AveragePooling2D=AveragePooling2D,
                               GlobalAveragePooling2D=GlobalAveragePooling2D, GlobalAveragePooling1D=GlobalAveragePooling1D,
                               MultiHeadAttention=MultiHeadAttention, concatenate=concatenate)


# ---- keras -------------------------------------------------------------------------------------------------------------------
def Input(shape=None, name=None, **kw):
    shape = tuple(int(s) for s in shape)
    if name not in _S.inputs:                                # tools/grad_cam_CvT.py:317-318 leaves its inputs unnamed: the one of this shape
        fits = [k for k, v in _S.inputs.items() if tuple(v.shape[1:]) == shape]
        assert len(fits) == 1, (name, shape, {k: tuple(v.shape) for k, v in _S.inputs.items()})
        name = fits[0]
    x = _S.inputs[name]
    assert tuple(x.shape[1:]) == shape, (name, tuple(x.shape), shape)
    return x


class Model:
    def __init__(self, inputs=None, outputs=None, **kw):
        self.inputs, self.outputs, self.output = inputs, outputs, outputs


class Sequential(Layer):
    def __init__(self, layers_=None, name=None):
        super().__init__(name=name)
        self.layers = list(layers_ or [])

    def call(self, x):
        for l in self.layers:
            x = l(x)
        return x


keras = types.SimpleNamespace(Input=Input, Model=Model, Sequential=Sequential, layers=layers)


# ---- tf -----------------------------------------------------------------------------------------------------------------------
def _gelu(x, approximate=False):
    assert not approximate
    return 0.5 * x * (1.0 + torch.erf(x / math.sqrt(2.0)))


nn = types.SimpleNamespace(gelu=_gelu, relu=torch.relu)
tf = types.SimpleNamespace(
    shape=lambda x: tuple(int(s) for s in x.shape),
    reshape=lambda x, shape: x.reshape(tuple(int(s) for s in shape)),
    split=lambda x, sizes, axis=0: list(torch.split(x, [int(s) for s in sizes], dim=axis)),
    concat=lambda xs, axis: torch.cat(list(xs), dim=axis),
    tile=lambda x, multiples: x.repeat(*[int(m) for m in multiples]),
    squeeze=lambda x, axis=None: x.squeeze(axis) if axis is not None else x.squeeze(),
    reduce_mean=lambda x, axis=None: x.mean(axis) if axis is not None else x.mean(),
    expand_dims=lambda x, axis: x.unsqueeze(axis),
    cast=lambda x, dtype: x,
    float32="float32", nn=nn, keras=keras)


# ---- the slice of tf that tools/grad_cam_CvT.py:422-481 (make_gradcam_heatmap) uses on top of the above ---------------------------------
class _Ref:
    """A symbolic tensor of a ReferenceModel: its final output, or the output of one named layer."""

    def __init__(self, model, layer):
        self.model, self.layer = model, layer


class ReferenceModel:
    """What `create_cvt_model(...)` stands for when the model has to be called more than once (Grad-CAM): every call re-executes the
    reference's create_cvt_model on the given inputs with the same weights."""

    def __init__(self, path, script_globals, args, weights, input_names=("Image_inputs", "Proc_inputs")):
        self.path, self.script_globals, self.args, self.weights = path, dict(script_globals), list(args), list(weights)
        self.inputs = list(input_names)
        self.output = _Ref(self, None)

    def get_layer(self, name):
        return types.SimpleNamespace(output=_Ref(self, name))

    def run(self, tensors, layers_wanted):
        inputs = dict(zip(self.inputs, tensors))
        out, s = run_reference_model(self.path, self.script_globals, self.args, self.weights, inputs)
        return [out if l is None else s.captured[l] for l in layers_wanted]


class _SubModel:
    def __init__(self, inputs=None, outputs=None):
        assert all(isinstance(o, _Ref) for o in outputs) and len({id(o.model) for o in outputs}) == 1
        self.model, self.wanted = outputs[0].model, [o.layer for o in outputs]

    def __call__(self, tensors):
        tensors = [t.detach().clone().requires_grad_(True) for t in tensors]      # so that an intermediate output is part of a differentiable graph
        return self.model.run(tensors, self.wanted)


class GradientTape:
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False

    def gradient(self, target, source):                      # tf: the gradient of sum(target)
        return torch.autograd.grad(target.sum(), source, retain_graph=True, allow_unused=True)[0]


class _Dataset:
    def __init__(self, tensors):
        self.tensors, self.bs = tensors, 1

    def batch(self, n):
        self.bs = int(n)
        return self

    def __iter__(self):
        n = self.tensors[0].shape[0]
        for i in range(0, n, self.bs):
            yield tuple(t[i:i + self.bs] for t in self.tensors)

    def __len__(self):
        return -(-self.tensors[0].shape[0] // self.bs)


tf.GradientTape = GradientTape
tf.data = types.SimpleNamespace(Dataset=types.SimpleNamespace(from_tensor_slices=lambda ts: _Dataset([torch.as_tensor(t, dtype=torch.float64) for t in ts])))
tf.config = types.SimpleNamespace(experimental=types.SimpleNamespace(get_memory_usage=lambda dev: 0))
tf.reduce_sum = lambda x, axis=None: x.sum(axis) if axis is not None else x.sum()
tf.maximum = lambda x, y: torch.clamp_min(x.detach(), float(y))      # (only used for the heat-map post-processing: no gradient wanted)
tf.math = types.SimpleNamespace(reduce_max=lambda x: x.detach().max())
keras.models = types.SimpleNamespace(Model=_SubModel)


# ---- running a reference script's model section ------------------------------------------------------------------------------------
MODEL_NAMES = ("extract_dimensions", "Projection", "ConvAttention", "ConvEmbed", "ConvTransformerBlock", "create_cvt_model")


def compile_model_section(path: str, names: Iterable[str] = MODEL_NAMES):
    """The model classes, create_cvt_model and the `spec = {...}` assignment of one reference script, unmodified, as a code object."""
    tree = ast.parse(open(path, encoding="utf-8").read(), path)
    names = set(names)
    nodes = [n for n in tree.body
             if (isinstance(n, (ast.FunctionDef, ast.ClassDef)) and n.name in names)
             or (isinstance(n, ast.Assign) and any(isinstance(t, ast.Name) and t.id == "spec" for t in n.targets))]
    found = {n.name for n in nodes if not isinstance(n, ast.Assign)}
    assert {"create_cvt_model", "ConvTransformerBlock", "ConvAttention"} <= found, found
    return compile(ast.Module(body=nodes, type_ignores=[]), path, "exec")


def run_reference_model(path: str, script_globals: dict, args: Sequence, weights: List[Tuple[str, torch.Tensor]], inputs: Dict[str, torch.Tensor],
                        training: bool = False, dropout_masks: Optional[Sequence[torch.Tensor]] = None):
    """Executes the script's model section with tf / keras / layers bound to this module and calls its create_cvt_model(*args) on `inputs`
    ({'Image_inputs': [B,H,W,C], 'Proc_inputs': [B,P]}).  Returns (outputs, session)."""
    ns = dict(tf=tf, keras=keras, layers=layers, print=lambda *a, **k: None)
    ns.update(script_globals)
    with session(weights, inputs, training, dropout_masks) as s:
        exec(compile_model_section(path), ns)
        model = ns["create_cvt_model"](*args)
        s.assert_all_weights_used()
    return model.outputs, s
