"""Self-checks of the arithmetic inside each layer of the oracle (the level that stays unpinned without TensorFlow; the wiring is pinned
to the reference's own model source by tests/test_oracle_vs_reference_source.py - see oracle/cvt_oracle.py header)."""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import cvt_oracle as O


def small_spec(method="dw_bn", cls=True, proc_dim=5, nc=1, hw=(32, 32)):
    return O.ref_spec(method, cls, proc_dim, nc, image_h=hw[0], image_w=hw[1])


def test_shapes_match_model_plot_png():
    # model_plot_Cls.png: 128x128x1 -> 32x32x64 -> 16x16x128 -> 8x8x256, cls (None,1,256), concat 512 -> 1
    spec = O.ref_spec()
    geo = O.stage_geometry(spec)
    assert [(g["h"], g["w"], g["d"]) for g in geo] == [(32, 32, 64), (16, 16, 128), (8, 8, 256)]
    assert [g["tc"] for g in geo] == [1024, 256, 65]
    assert [g["pad"] for g in geo] == [(1, 2, 1, 2), (0, 1, 0, 1), (0, 1, 0, 1)]
    p = O.init_params(spec)
    img, proc, _ = O.synthetic_batch(spec, 2)
    pred, taps = O.forward(spec, p, img, proc)
    assert pred.shape == (2, 1)
    assert taps["s3.out"].shape == (2, 8, 8, 256) and taps["s3.x2"].shape == (2, 65, 256)
    assert p["Final_Dense/kernel"].shape == (512, 1)


def test_param_counts():
    spec = O.ref_spec()
    tot = sum(int(np.prod(s)) for _, s, _ in O.param_specs(spec))
    tr = sum(int(np.prod(s)) for _, s, t in O.param_specs(spec) if t)
    assert (tot, tr) == (1841281, 1838593)
    img_spec = O.ref_spec(proc_dim=0)
    assert sum(int(np.prod(s)) for _, s, _ in O.param_specs(img_spec)) == 1773697
    wide = O.wide_spec()
    g = O.stage_geometry(wide)
    assert [(x["h"], x["w"]) for x in g] == [(85, 87), (43, 44), (22, 22)]
    assert [x["pad"] for x in g] == [(1, 2, 3, 3), (1, 1, 1, 1), (1, 1, 0, 1)]


def test_layers_against_torch_library_ops():
    rng = np.random.default_rng(0)
    x = torch.tensor(rng.standard_normal((2, 9, 11, 8)))
    # LayerNorm
    g, b = torch.tensor(rng.standard_normal(8)), torch.tensor(rng.standard_normal(8))
    assert torch.allclose(O.layer_norm(x, g, b, 1e-6), F.layer_norm(x, (8,), g, b, 1e-6), atol=1e-12)
    # depthwise conv vs grouped conv2d
    w = torch.tensor(rng.standard_normal((3, 3, 8, 1)))
    ref = F.conv2d(x.permute(0, 3, 1, 2), w.permute(2, 3, 0, 1), padding=1, groups=8).permute(0, 2, 3, 1)
    assert torch.allclose(O.depthwise3x3(x, w), ref, atol=1e-12)
    # avg pool, count_include_pad=False == TF SAME average
    ref = F.avg_pool2d(x.permute(0, 3, 1, 2), 3, 1, 1, count_include_pad=False).permute(0, 2, 3, 1)
    assert torch.allclose(O.avgpool3x3_same(x), ref, atol=1e-12)
    # batch norm (training): library functional with biased batch variance
    y, mu, var = O.batch_norm(x, g, b, None, None, True, 1e-3)
    ref = F.batch_norm(x.permute(0, 3, 1, 2), None, None, g, b, True, 0.0, 1e-3).permute(0, 2, 3, 1)
    assert torch.allclose(y, ref, atol=1e-12)
    # gelu
    assert torch.allclose(O.gelu_erf(x), F.gelu(x), atol=1e-12)


def test_conv_same_explicit_loops():
    rng = np.random.default_rng(1)
    for (h, w, k, s) in [(12, 12, 7, 4), (9, 10, 3, 2), (8, 8, 3, 2)]:
        x = rng.standard_normal((1, h, w, 2)); ker = rng.standard_normal((k, k, 2, 3)); bias = rng.standard_normal(3)
        y = O.conv_same(torch.tensor(x), torch.tensor(ker), torch.tensor(bias), s).numpy()
        ho, pt, _ = O.same_pad(h, k, s); wo, pl, _ = O.same_pad(w, k, s)
        assert y.shape == (1, ho, wo, 3)
        ref = np.zeros_like(y)
        for i in range(ho):
            for j in range(wo):
                acc = bias.copy()
                for a in range(k):
                    for b in range(k):
                        yy, xx = s * i + a - pt, s * j + b - pl
                        if 0 <= yy < h and 0 <= xx < w:
                            acc += x[0, yy, xx] @ ker[a, b]
                ref[0, i, j] = acc
        assert np.allclose(y, ref, atol=1e-12)


def test_mha_against_torch_multi_head_attention():
    """The keras MHA restatement (einsum kernels [D,N,H]) equals torch's fused MHA with the same weights."""
    spec = small_spec(hw=(16, 16))
    p = O.init_params(spec, 7)
    img, proc, _ = O.synthetic_batch(spec, 2, 3)
    _, taps = O.forward(spec, p, img, proc)
    t = "stage2_transformer"
    d, n = 128, 2
    q1, k1, v1 = [torch.tensor(a) for a in taps["s2.qkv1"]]
    wq = p[f"{t}/attn/mha/query/kernel"].reshape(d, d); wk = p[f"{t}/attn/mha/key/kernel"].reshape(d, d)
    wv = p[f"{t}/attn/mha/value/kernel"].reshape(d, d); wo = p[f"{t}/attn/mha/attention_output/kernel"].reshape(d, d)
    in_w = torch.tensor(np.concatenate([wq.T, wk.T, wv.T], 0))
    in_b = torch.tensor(np.concatenate([p[f"{t}/attn/mha/{x}/bias"].reshape(-1) for x in ("query", "key", "value")]))
    out, _ = F.multi_head_attention_forward(
        q1.transpose(0, 1), k1.transpose(0, 1), v1.transpose(0, 1), d, n, in_w, in_b, None, None, False, 0.0,
        torch.tensor(wo.T), torch.tensor(p[f"{t}/attn/mha/attention_output/bias"]), need_weights=False)
    assert np.allclose(out.transpose(0, 1).numpy(), taps["s2.a1"], atol=1e-10)


@pytest.mark.parametrize("method,cls,proc_dim", [("dw_bn", True, 5), ("avg", False, 0), ("linear", True, 5)])
def test_finite_difference_gradients(method, cls, proc_dim):
    spec = small_spec(method, cls, proc_dim, hw=(16, 16))
    spec.drop_rate = 0.1
    p = O.init_params(spec, 11)
    img, proc, y = O.synthetic_batch(spec, 3, 5)
    y = y / 1000.0
    masks = O.make_dropout_masks(spec, 3, seed=99, step=1)
    r = O.loss_and_grads(spec, p, img, proc, y, masks)
    rng = np.random.default_rng(0)
    names = [k for k in r["grads"]]
    for name in names:
        g = r["grads"][name]
        idx = tuple(rng.integers(0, s) for s in g.shape)
        eps = 1e-6
        pp = {k: v.copy() for k, v in p.items()}; pp[name][idx] += eps
        pm = {k: v.copy() for k, v in p.items()}; pm[name][idx] -= eps
        lp = O.loss_and_grads(spec, pp, img, proc, y, masks)["loss"]
        lm = O.loss_and_grads(spec, pm, img, proc, y, masks)["loss"]
        fd = (lp - lm) / (2 * eps)
        assert abs(fd - g[idx]) <= 1e-6 + 1e-4 * abs(fd), (name, idx, fd, g[idx])


def test_inference_bn_fold_and_linear_identity():
    spec = small_spec("dw_bn", True, hw=(16, 16))
    p = O.init_params(spec, 3)
    img, proc, _ = O.synthetic_batch(spec, 2, 1)
    _, taps = O.forward(spec, p, img, proc, training=False)
    t = "stage1_transformer"
    # BN folded into the depthwise weights
    xn = torch.tensor(taps["s1.xn"]).reshape(2, 4, 4, 64)
    sc = p[f"{t}/attn/k_proj/bn/gamma"] / np.sqrt(p[f"{t}/attn/k_proj/bn/moving_variance"] + spec.bn_eps)
    wf = torch.tensor(p[f"{t}/attn/k_proj/depthwise_kernel"] * sc[None, None, :, None])
    bf = p[f"{t}/attn/k_proj/bn/beta"] - p[f"{t}/attn/k_proj/bn/moving_mean"] * sc
    folded = O.depthwise3x3(xn, wf).numpy() + bf
    assert np.allclose(folded.reshape(2, 16, 64), taps["s1.qkv0"][1], atol=1e-10)
    spec_l = small_spec("linear", True, hw=(16, 16))
    pl = O.init_params(spec_l, 3)
    _, tl = O.forward(spec_l, pl, img, proc)
    assert np.array_equal(tl["s3.qkv0"][0], tl["s3.xn"]) and np.array_equal(tl["s3.qkv0"][2], tl["s3.xn"])


def test_philox_known_answer():
    # Random123 KAT for philox4x32-10: ctr=0,key=0 and ctr=ff..,key=ff..
    z = O.philox4x32(np.zeros((1, 4), np.uint32), (0, 0))[0]
    assert [hex(int(v)) for v in z] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    f = O.philox4x32(np.full((1, 4), 0xFFFFFFFF, np.uint32), (0xFFFFFFFF, 0xFFFFFFFF))[0]
    assert [hex(int(v)) for v in f] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    m = O.dropout_mask(100000, 0.1, 5, 1, 2)
    assert abs(m.mean() - 0.9) < 0.005


def test_adam_and_lr_schedule():
    p = {"a": np.array([1.0, -2.0])}; g = {"a": np.array([0.5, -0.25])}
    m = {"a": np.zeros(2)}; v = {"a": np.zeros(2)}
    p, m, v = O.adam_step(p, g, m, v, 1, 1e-3)
    # first Adam step moves every weight by ~lr against the gradient sign
    assert np.allclose(p["a"], [1.0 - 1e-3, -2.0 + 1e-3], atol=1e-8)
    lr = 1e-3
    for e in range(101):
        lr = O.lr_scheduler(e, lr)
    assert math.isclose(lr, 1e-3 * 0.64)


def test_bn_moving_update_uses_unbiased_variance():
    spec = small_spec(hw=(16, 16))
    p = O.init_params(spec, 2)
    img, proc, y = O.synthetic_batch(spec, 2, 0)
    r = O.loss_and_grads(spec, p, img, proc, y, None)
    name = "stage3_transformer/attn/v_proj/bn"
    c = r["taps"]["s3.conv_qkv"][2].reshape(-1, 256)
    n = c.shape[0]
    exp = 0.99 * p[name + "/moving_variance"] + 0.01 * c.var(0) * n / (n - 1)
    assert np.allclose(r["new_moving"][name + "/moving_variance"], exp, atol=1e-12)
