"""Model variants of the tools/ scripts against the oracle, through the C-ABI:
  tools/grad_cam_CvT.py:89-349  cls token on EVERY stage (dropped after each block), proj_q/k/v built but never applied, Dropout on the
                                attention output, Projection 'linear' = Dense(dim), GlobalAveragePooling2D head without LayerNorm;
  tools/model_plot.py:289-292   un-normalised heads: the bare cls token / the bare token mean.
Tolerances as in test_gpu_parity.py: fp32 mode 2e-4 activations / 1e-3 gradients, TF32 mode 5e-3 / 2e-2 (max|a-b| / max|b| per tensor)."""
import numpy as np
import pytest
import torch

import transformer_stm_b200 as T
from oracle import cvt_oracle as O
from tests.parity_utils import compare_grads, compare_taps, fmt, rel_err

pytestmark = pytest.mark.gpu

TAPS = ("embed", "x0", "xn", "qkv2", "attn_o", "x1", "y", "hact", "x2", "out")


def _dev(a):
    return torch.tensor(np.ascontiguousarray(a, dtype=np.float32), device="cuda")


def _pair(spec, cfg, batch, seed=1234):
    eng = T.CvtEngine(cfg)
    params = O.init_params(spec, seed, "parity")
    eng.set_params(params)
    img, proc, y = O.synthetic_batch(spec, batch, seed=7)
    return eng, params, img, proc, y


def _check(spec, eng, params, img, proc, y, tf32, masks_seed=5):
    B = img.shape[0]
    at, gt = (5e-3, 2e-2) if tf32 else (2e-4, 1e-3)
    # inference
    pred = eng.predict(img, proc if spec.proc_dim > 0 else None)
    ref, taps = O.forward(spec, params, img, proc, training=False)
    # TF32 inference: stages with D = 64 / 128 run the fused Mlp kernel, which keeps y / hact on the SM
    bad, report = compare_taps(eng, taps, tol=at, names=[t for t in TAPS if not (tf32 and t in ("y", "hact"))])
    assert not bad, "inference mismatches:\n" + fmt(report)
    assert rel_err(pred, ref) <= at
    # training forward + backward with the oracle's dropout masks
    masks = O.make_dropout_masks(spec, B, seed=masks_seed, step=0)
    eng.set_dropout("masks")
    eng.set_dropout_masks(masks)
    d = (_dev(img), _dev(proc) if spec.proc_dim > 0 else None, _dev(y))
    torch.cuda.synchronize()
    eng.forward_dev(d[0], d[1], B, True)
    loss, mae = eng.loss_backward_dev(d[2])
    r = O.loss_and_grads(spec, params, img, proc, y, masks)
    bad, report = compare_taps(eng, r["taps"], tol=at, names=TAPS)
    assert not bad, "training forward mismatches:\n" + fmt(report)
    assert abs(loss - r["loss"]) <= (2e-3 if tf32 else 1e-5) * abs(r["loss"])
    bad, report = compare_grads(eng, r["grads"], tol=gt, floor_frac=1e-3 if tf32 else 1e-4)
    assert not bad, "gradient mismatches:\n" + fmt(report)


@pytest.mark.parametrize("precision", ["fp32", "tf32"])
@pytest.mark.parametrize("cls,method", [(False, "dw_bn"), (True, "dw_bn"), (True, "linear"), (False, "avg")])
def test_gradcam_script_model(cls, method, precision):
    hw = (32, 32)
    spec = O.gradcam_spec(cls, method, image_h=hw[0], image_w=hw[1])
    cfg = T.CvtConfig.gradcam(cls, method, image_h=hw[0], image_w=hw[1], max_batch=3, precision=precision)
    eng, params, img, proc, y = _pair(spec, cfg, 3)
    infos = {p["name"]: tuple(p["shape"]) for p in eng.param_infos()}
    assert infos == {n: tuple(s) for n, s, _ in O.param_specs(spec)}
    assert not any("proj_q" in n or "head_norm" in n for n in infos)
    _check(spec, eng, params, img, proc, y, precision == "tf32")
    eng.close()


@pytest.mark.parametrize("precision", ["fp32", "tf32"])
@pytest.mark.parametrize("head,cls", [("cls", True), ("mean", True), ("mean", False), ("ln_mean", True)])
def test_head_variants(head, cls, precision):
    """tools/model_plot.py:289-292 (bare cls token / bare mean) on the otherwise unchanged models/CvT(Par).py network (folded TF32 path)."""
    hw = (36, 44)
    spec = O.ref_spec("dw_bn", cls, image_h=hw[0], image_w=hw[1]); spec.head = head
    cfg = T.CvtConfig.reference("dw_bn", cls, image_h=hw[0], image_w=hw[1], max_batch=3, precision=precision, head=head)
    eng, params, img, proc, y = _pair(spec, cfg, 3)
    _check(spec, eng, params, img, proc, y, precision == "tf32")
    eng.close()


@pytest.mark.parametrize("precision", ["fp32", "tf32"])
def test_cls_token_on_every_stage_reference_network(precision):
    """with_cls_token honoured per stage (tools/grad_cam_CvT.py:324-332) with everything else as models/CvT(Par).py: in TF32 mode this is
    the folded path, whose ConvEmbed reads the previous block's output past its cls row."""
    hw = (32, 32)
    spec = O.ref_spec("dw_bn", True, image_h=hw[0], image_w=hw[1])
    cfg = T.CvtConfig.reference("dw_bn", True, image_h=hw[0], image_w=hw[1], max_batch=4, precision=precision)
    for st in spec.stages: st.with_cls_token = True
    for st in cfg.stages: st.with_cls_token = True
    eng, params, img, proc, y = _pair(spec, cfg, 4)
    _check(spec, eng, params, img, proc, y, precision == "tf32")
    eng.close()


def test_variant_errors_are_loud():
    with pytest.raises(T.CvtError):
        T.CvtEngine(T.CvtConfig.reference("dw_bn", False, head="cls"))            # cls head without a cls token
    with pytest.raises(ValueError):
        T.CvtConfig.reference(head="nope").to_c()


def test_which_contraction_kernels_the_layer_by_layer_tf32_path_uses():
    """The Grad-CAM script's model in TF32 mode runs the unfolded (layer-by-layer) path.  Pins where its dense products go: everything the
    TMA-fed kernel takes goes there; the ConvEmbeds (implicit GEMM with a gathered A operand: stage 1 has Cin = 1, the unfolded path keeps
    no patch matrix) and the split-K weight gradients go to the older tcgen05 kernel (gemm_tc.cu); nothing lands on the FFMA kernel."""
    hw = (64, 64)
    spec = O.gradcam_spec(False, "dw_bn", image_h=hw[0], image_w=hw[1])
    cfg = T.CvtConfig.gradcam(False, "dw_bn", image_h=hw[0], image_w=hw[1], max_batch=4, precision="tf32")
    eng, params, img, proc, y = _pair(spec, cfg, 4)
    eng.set_dropout("off")
    eng.set_graphs(False)
    eng.gemm_dispatch_counts(reset=True)
    eng.train_step(img, proc, y, 1e-3)
    tma, tc, simt = eng.gemm_dispatch_counts(reset=True)
    print("\n[dispatch] gradcam variant, TF32 unfolded: gemm_tma %d, gemm_tc %d, gemm_simt %d" % (tma, tc, simt))
    assert tma > 0 and tc > 0, (tma, tc, simt)
    assert simt == 0, (tma, tc, simt)
    eng.close()
