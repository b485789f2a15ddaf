"""Pins the WIRING of oracle/cvt_oracle.py to the reference's own model source: the classes Projection / ConvAttention / ConvEmbed /
ConvTransformerBlock and create_cvt_model are compiled from the files under /root/reference as they lie there and executed over
oracle/mini_keras.py (a define-by-run stand-in for the Keras layers they use, float64); predictions, and the gradients of the MSE loss with
respect to every weight (torch autograd through the reference's graph code), have to agree with the oracle's to rounding.  What this cannot
pin is the arithmetic inside TensorFlow's kernels - TensorFlow is not available here - see the header of oracle/mini_keras.py.
Runs only where the reference tree exists (this container); tests/golden/refsrc_*.npz carries its outputs to the GPU box."""
import os

import numpy as np
import pytest
import torch

from oracle import cvt_oracle as O
from oracle import mini_keras as MK

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not present (it does not travel to the GPU box)")


def weights_of(spec, params, requires_grad=False):
    out = []
    for name, shape, trainable in O.param_specs(spec):
        t = torch.tensor(np.asarray(params[name]), dtype=torch.float64)
        if requires_grad and trainable:
            t.requires_grad_(True)
        out.append((name, t))
    return out


def mask_queue(spec, masks):
    """The keep masks in the order the reference calls its Dropout layers: per stage [attention-output (grad-cam variant)], proj, Mlp 1, Mlp 2."""
    q = []
    for i in range(1, len(spec.stages) + 1):
        for nm in ("attn", "proj", "mlp1", "mlp2"):
            if "s%d.%s" % (i, nm) in masks:
                q.append(torch.tensor(masks["s%d.%s" % (i, nm)], dtype=torch.float64))
    return q


def run(script, spec, params, img, proc, *, method, cls, training=False, masks=None, requires_grad=False):
    inputs = {"Image_inputs": torch.tensor(img, dtype=torch.float64)}
    args = [spec.image_h, spec.image_w, spec.in_chans]
    if spec.proc_dim > 0:
        inputs["Proc_inputs"] = torch.tensor(proc, dtype=torch.float64)
        args.append(spec.proc_dim)
    args.append(spec.num_classes)
    w = weights_of(spec, params, requires_grad)
    out, s = MK.run_reference_model(os.path.join(REF, script), dict(projection_method=method, cls_token_switch=cls), args, w, inputs,
                                    training=training, dropout_masks=mask_queue(spec, masks) if masks else None)
    assert s.taken == [n for n, _, _ in O.param_specs(spec)]      # the reference creates its variables in the canonical order
    return out, dict(w)


CASES = [("dw_bn", True, (32, 32)), ("dw_bn", False, (32, 32)), ("avg", True, (32, 32)), ("avg", False, (36, 44)), ("linear", True, (36, 44)),
         ("linear", False, (32, 32)), ("dw_bn", True, (36, 44))]


@pytest.mark.parametrize("method,cls,hw", CASES)
def test_cvt_par_inference_matches_the_reference_source(method, cls, hw):
    spec = O.ref_spec(method, cls, proc_dim=5, num_classes=2, image_h=hw[0], image_w=hw[1])
    params = O.init_params(spec, 7, "parity")
    img, proc, _ = O.synthetic_batch(spec, 3, 1)
    want, _ = O.forward(spec, params, img, proc, training=False)
    got, _ = run("models/CvT(Par).py", spec, params, img, proc, method=method, cls=cls)
    np.testing.assert_allclose(got.numpy(), want, rtol=1e-10, atol=1e-10)


@pytest.mark.parametrize("method,cls,hw", CASES[:5])
def test_cvt_par_train_step_gradients_match_the_reference_source(method, cls, hw):
    spec = O.ref_spec(method, cls, proc_dim=5, num_classes=1, image_h=hw[0], image_w=hw[1])
    params = O.init_params(spec, 11, "parity")
    img, proc, y = O.synthetic_batch(spec, 3, 2)
    masks = O.make_dropout_masks(spec, 3, seed=5, step=0)
    want = O.loss_and_grads(spec, params, img, proc, y, masks)                    # Keras train_step: batch-statistics BN, dropout, MSE, autodiff
    out, w = run("models/CvT(Par).py", spec, params, img, proc, method=method, cls=cls, training=True, masks=masks, requires_grad=True)
    loss = ((out - torch.tensor(y, dtype=torch.float64).reshape(out.shape)) ** 2).mean()
    loss.backward()
    assert float(loss.detach()) == pytest.approx(want["loss"], rel=1e-12)
    scale = max(float(np.abs(g).max()) for g in want["grads"].values())
    for name, g in want["grads"].items():
        got = w[name].grad
        got = np.zeros_like(g) if got is None else got.numpy()
        assert np.abs(got - g).max() <= 1e-10 * scale + 1e-9 * np.abs(g).max(), name


@pytest.mark.parametrize("method,cls", [("dw_bn", True), ("avg", False), ("linear", True)])
def test_cvt_img_matches_the_reference_source(method, cls):
    spec = O.ref_spec(method, cls, proc_dim=0, num_classes=1, image_h=32, image_w=32)
    params = O.init_params(spec, 3, "parity")
    img, _, _ = O.synthetic_batch(spec, 2, 3)
    want, _ = O.forward(spec, params, img, None, training=False)
    got, _ = run("models/CvT(Img).py", spec, params, img, None, method=method, cls=cls)
    np.testing.assert_allclose(got.numpy(), want, rtol=1e-10, atol=1e-10)


@pytest.mark.parametrize("script", ["models/CvT(Par)(many).py", "models/CvT_test(Par).py", "models/CvT_test(Par)(many).py"])
def test_the_other_scripts_define_the_same_network(script):
    spec = O.ref_spec("dw_bn", True, proc_dim=5, num_classes=1, image_h=32, image_w=32)
    params = O.init_params(spec, 5, "parity")
    img, proc, _ = O.synthetic_batch(spec, 2, 4)
    want, _ = O.forward(spec, params, img, proc, training=False)
    got, _ = run(script, spec, params, img, proc, method="dw_bn", cls=True)
    np.testing.assert_allclose(got.numpy(), want, rtol=1e-10, atol=1e-10)


def test_gradcam_tool_with_avg_is_a_mixed_projection_and_is_refused():
    """Executing the tool's source shows what a reading missed: with projection_method = 'avg' its q-projection is the 'linear' Projection (:141),
    which in THIS script is a Dense(dim) (:103).  The engine has no such mixed mode; keras_api refuses to build that model instead of building another."""
    from transformer_stm_b200 import keras_api as K
    from tests.fake_engine import OracleEngine
    spec = O.gradcam_spec(False, "avg", proc_dim=5, num_classes=1, image_h=32, image_w=32)      # q identity: NOT the tool's 'avg' model
    params = O.init_params(spec, 9, "parity")
    img, proc, _ = O.synthetic_batch(spec, 2, 5)
    with pytest.raises(AssertionError, match=r"Dense\(None\) kernel creates a variable of shape \(64, 64\)"):
        run("tools/grad_cam_CvT.py", spec, params, img, proc, method="avg", cls=False)
    K.configure(projection="avg", cls_token=False, variant="gradcam")
    try:
        with pytest.raises(NotImplementedError, match="mixes a Dense q-projection"):
            K.create_cvt_model(32, 32, 1, 5, 1, engine_factory=OracleEngine)
    finally:
        K.configure(projection="dw_bn", cls_token=True, variant="reference")


@pytest.mark.parametrize("method,cls", [("dw_bn", False), ("dw_bn", True), ("linear", True), ("linear", False)])
def test_gradcam_tool_variant_matches_its_source(method, cls):
    """tools/grad_cam_CvT.py:89-349: cls token on every stage and dropped after each block, proj_q/k/v built but never applied, attention-output
    dropout, a Dense inside the 'linear' Projection, GlobalAveragePooling2D head."""
    spec = O.gradcam_spec(cls, method, proc_dim=5, num_classes=1, image_h=32, image_w=32)
    params = O.init_params(spec, 9, "parity")
    img, proc, y = O.synthetic_batch(spec, 2, 5)
    want, _ = O.forward(spec, params, img, proc, training=False)
    got, _ = run("tools/grad_cam_CvT.py", spec, params, img, proc, method=method, cls=cls)
    np.testing.assert_allclose(got.numpy(), want, rtol=1e-10, atol=1e-10)
    masks = O.make_dropout_masks(spec, 2, seed=6, step=0)
    want = O.loss_and_grads(spec, params, img, proc, y, masks)
    out, w = run("tools/grad_cam_CvT.py", spec, params, img, proc, method=method, cls=cls, training=True, masks=masks, requires_grad=True)
    ((out - torch.tensor(y, dtype=torch.float64).reshape(out.shape)) ** 2).mean().backward()
    scale = max(float(np.abs(g).max()) for g in want["grads"].values())
    for name, g in want["grads"].items():
        got = w[name].grad
        got = np.zeros_like(g) if got is None else got.numpy()
        assert np.abs(got - g).max() <= 1e-10 * scale + 1e-9 * np.abs(g).max(), name


@pytest.mark.parametrize("cls", [False, True])
def test_model_plot_tool_variant_matches_its_source(cls):
    """tools/model_plot.py:243-300: the reference network with un-normalised heads (the bare cls token, or the mean over the last feature map).
    Its spec (:17-23) hard-codes qkv_method 'linear' whatever its projection_method constant says - executing the source shows it (a dw_bn
    parameter list is refused at the first variable); keras_api mirrors it with configure(projection='linear', variant='model_plot')."""
    method = "linear"
    wrong = O.ref_spec("dw_bn", cls, proc_dim=5, num_classes=1, image_h=32, image_w=32)
    wrong.head = "cls" if cls else "mean"
    img, proc, y = O.synthetic_batch(wrong, 2, 6)
    with pytest.raises(AssertionError, match="q_proj/depthwise_kernel"):
        run("tools/model_plot.py", wrong, O.init_params(wrong, 13, "parity"), img, proc, method="dw_bn", cls=cls)
    spec = O.ref_spec(method, cls, proc_dim=5, num_classes=1, image_h=32, image_w=32)
    spec.head = "cls" if cls else "mean"
    params = O.init_params(spec, 13, "parity")
    img, proc, y = O.synthetic_batch(spec, 2, 6)
    want, _ = O.forward(spec, params, img, proc, training=False)
    got, _ = run("tools/model_plot.py", spec, params, img, proc, method=method, cls=cls)
    np.testing.assert_allclose(got.numpy(), want, rtol=1e-10, atol=1e-10)
    assert not any(n.startswith("head_norm") for n, _, _ in O.param_specs(spec))


@pytest.mark.parametrize("method", ["dw_bn", "linear"])
def test_gradcam_heatmaps_match_the_tools_own_function(method):
    """tools/grad_cam_CvT.py:422-481 make_gradcam_heatmap as written (tf.data batches, a sub-model to the named block's output, GradientTape,
    pooled gradients, relu / max normalisation) over the tool's own model source - against oracle.gradcam and keras_api.make_gradcam_heatmap."""
    import ast
    from tqdm import tqdm
    from transformer_stm_b200 import keras_api as K
    from tests.fake_engine import OracleEngine
    path = os.path.join(REF, "tools/grad_cam_CvT.py")
    spec = O.gradcam_spec(False, method, proc_dim=5, num_classes=1, image_h=32, image_w=32)
    params = O.init_params(spec, 21, "parity")
    img, proc, _ = O.synthetic_batch(spec, 3, 8)
    tree = ast.parse(open(path, encoding="utf-8").read(), path)
    fn = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "make_gradcam_heatmap"]
    ns = dict(tf=MK.tf, np=np, tqdm=tqdm, print=lambda *a, **k: None)
    exec(compile(ast.Module(body=fn, type_ignores=[]), path, "exec"), ns)
    model = MK.ReferenceModel(path, dict(projection_method=method, cls_token_switch=False), [32, 32, 1, 5, 1], weights_of(spec, params))
    heat = ns["make_gradcam_heatmap"](img[..., 0].astype(np.float64), proc.astype(np.float64), model, "stage3_transformer")      # [n, h3, w3]
    assert heat.shape == (3, 2, 2)
    for i in range(3):
        ref, _, _ = O.gradcam(spec, params, img[i:i + 1], proc[i:i + 1], "s3.out")
        np.testing.assert_allclose(heat[i], ref[0], rtol=1e-8, atol=1e-10)
    K.configure(projection=method, cls_token=False, variant="gradcam")
    try:
        m = K.create_cvt_model(32, 32, 1, 5, 1, engine_factory=OracleEngine, max_batch=2)
        m.set_weights_dict(params)
        mine = K.make_gradcam_heatmap(img[..., 0], proc, m, "stage3_transformer")
        np.testing.assert_allclose(mine, heat, rtol=1e-4, atol=1e-6)
    finally:
        K.configure(projection="dw_bn", cls_token=True, variant="reference")
