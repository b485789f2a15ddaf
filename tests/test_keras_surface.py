"""The Keras-2 call surface the reference scripts use, exercised in the scripts' own order:

  models/CvT(Par).py:463-494      create_cvt_model -> compile -> fit(validation_data, callbacks=[TensorBoard, LearningRateScheduler])
                                  -> save_weights(.h5) -> history -> xlsx record
  models/CvT_test(Par).py:510-530 create_cvt_model -> load_weights -> predict -> r2 / mse / mae -> xlsx
  tools/grad_cam_CvT.py:422-481   make_gradcam_heatmap

CPU tests drive keras_api.Model over the oracle-backed engine of tests/fake_engine.py (host logic only); the GPU tests run the same
sequence on libcvt_b200 and compare the trajectory with that engine step by step (same permutations, same Philox dropout stream)."""
import os

import numpy as np
import pytest

import transformer_stm_b200 as T
from oracle import cvt_oracle as O
from tests.fake_engine import OracleEngine
from transformer_stm_b200 import keras_api as K
from transformer_stm_b200 import xlsxlite

H = W = 32


def _data(n, seed, f64=True):
    rng = np.random.default_rng(seed)
    x = rng.random((n, H, W))                              # float64 [N,H,W]: what preprocess_data returns (models/CvT(Par).py:425)
    p = rng.standard_normal((n, 5))
    y = 1000.0 + 100.0 * rng.standard_normal(n)
    if not f64:
        x, p, y = x.astype(np.float32), p.astype(np.float32), y.astype(np.float32)
    return x, p, y


def _script_sequence(factory, tmp_path, epochs, precision="fp32", seed=3, n=10, nv=4, bs=4, f64=True):
    K.configure(projection="dw_bn", cls_token=True)
    x, p, y = _data(n, 0, f64)
    xv, pv, yv = _data(nv, 1, f64)
    model = K.create_cvt_model(H, W, 1, p.shape[1], 1, precision=precision, max_batch=bs, seed=seed, engine_factory=factory)
    model.compile(optimizer=K.Adam(learning_rate=1e-3), loss="mean_squared_error", metrics=["mae"])
    cbs = [K.TensorBoard(log_dir=str(tmp_path / "logs"), histogram_freq=1), K.LearningRateScheduler(K.lr_scheduler)]
    model.fit([x, p], y, epochs=epochs, batch_size=bs, validation_data=([xv, pv], yv), callbacks=cbs, verbose=0)
    return model, (x, p, y), (xv, pv, yv)


def _check_history(model, epochs):
    h = model.history.history
    assert set(h) == {"loss", "mae", "val_loss", "val_mae", "lr"}
    assert all(len(v) == epochs for v in h.values())
    want_lr = [1e-3 * 0.8 ** (e // 50) for e in range(epochs)]                 # models/CvT(Par).py:357-360
    np.testing.assert_allclose(h["lr"], want_lr, rtol=1e-6)
    assert np.isfinite(h["loss"]).all() and np.isfinite(h["val_loss"]).all()


def test_fit_history_lr_schedule_weights_and_sheets_cpu(tmp_path):
    epochs = 52
    model, (x, p, y), (xv, pv, yv) = _script_sequence(OracleEngine, tmp_path, epochs)
    _check_history(model, epochs)
    eng = model.engine
    # first epoch streams from the host arrays and fills the resident set, every later one runs from it
    assert eng.calls[0] == ("stream", 0) and all(c == ("resident", 0) for c in eng.calls[1:]) and len(eng.calls) == epochs
    assert eng.optimizer_steps() == epochs * 3                                     # 10 samples / batch 4: 4 + 4 + 2 (partial batch kept)
    assert model.history.history["loss"][-1] < model.history.history["loss"][0]
    # ---- save_weights(.h5) -> new model -> load_weights -> predict  (models/CvT(Par).py:489, CvT_test(Par).py:510-522) ----
    wpath = str(tmp_path / "Weight" / "cvt_model_weights_50HZ_Bm_dw_bn_clsTrue.h5")
    model.save_weights(wpath)
    assert open(wpath, "rb").read(8) == b"\x89HDF\r\n\x1a\n"
    m2 = K.create_cvt_model(H, W, 1, 5, 1, max_batch=4, engine_factory=OracleEngine)
    m2.load_weights(wpath)
    pred, pred2 = model.predict([xv, pv]), m2.predict([xv, pv])
    assert pred.shape == (4, 1) and pred.dtype == np.float32
    np.testing.assert_array_equal(pred, pred2)
    w1, w2 = model.get_weights_dict(), m2.get_weights_dict()
    assert all(np.array_equal(w1[k], w2[k]) for k in w1)
    # ---- records + prediction sheets (models/CvT(Par).py:492-494, CvT_test(Par).py:525-530) ----
    rpath = K.save_history_to_excel(model.history, str(tmp_path / "Records" / "cvt_records.xlsx"))
    rec = xlsxlite.read_xlsx(rpath)
    assert list(rec) == ["epoch", "loss", "mae", "val_loss", "val_mae", "lr"] and rec["epoch"] == list(range(1, epochs + 1))
    np.testing.assert_allclose(rec["loss"], model.history.history["loss"], rtol=1e-12)
    r2, mse, mae = K.regression_metrics(yv, pred.flatten())
    spath = K.save_predictions_to_excel(pred, yv, 10, 4, r2, mse, mae, str(tmp_path / "Excel" / "Predictions_Metrics_50HZ_Bm.xlsx"))
    sh = xlsxlite.read_xlsx(spath)
    assert list(sh) == ["Predictions", "Actual", "Errors(%)", "Train mounts", "Test mounts", "R2 Score", "MSE", "MAE"]   # tools/Compare_r.py:21-24
    assert sh["R2 Score"][0] == pytest.approx(r2) and sh["R2 Score"][1:] == [None] * 3 and sh["Train mounts"][0] == 10
    np.testing.assert_allclose(sh["Predictions"], pred.flatten().astype(np.float64), rtol=1e-7)
    import sklearn.metrics as SM
    assert r2 == pytest.approx(SM.r2_score(yv, pred.flatten())) and mse == pytest.approx(SM.mean_squared_error(yv, pred.flatten()))


def test_img_variant_and_argument_forms_cpu(tmp_path):
    """models/CvT(Img).py:291: create_cvt_model(image_height, image_width, num_channels, num_classes); bare-array inputs."""
    K.configure(projection="avg", cls_token=False)
    try:
        m = K.create_cvt_model(H, W, 1, 1, engine_factory=OracleEngine, max_batch=4, seed=0)
        assert m.cfg.proc_dim == 0 and m.cfg.stages[0].qkv_method == "avg" and not m.cfg.stages[2].with_cls_token
        m.compile(optimizer=K.Adam(learning_rate=1e-3), loss="mean_squared_error", metrics=["mae"])
        x, p, y = _data(6, 5)
        m.fit(x, y, epochs=2, batch_size=4, verbose=0)
        assert set(m.history.history) == {"loss", "mae"}
        assert m.predict(x).shape == (6, 1)
        assert m.engine.calls == [("stream", 0), ("resident", 0)]
        m.fit(x, y, epochs=1, batch_size=4, verbose=0)                             # a single epoch is streamed, nothing is kept resident
        assert m.engine.calls[-1] == ("stream", -1)
    finally:
        K.configure(projection="dw_bn", cls_token=True)
    x, p, y = _data(2, 0)
    with pytest.raises(RuntimeError):
        K.create_cvt_model(H, W, 1, 5, 1, engine_factory=OracleEngine).fit([x, p], y)                 # not compiled
    with pytest.raises(ValueError):
        K.create_cvt_model(H, W, 1, 5, 1, engine_factory=OracleEngine).compile(optimizer="sgd", loss="mean_squared_error")
    par = K.create_cvt_model(H, W, 1, 5, 1, engine_factory=OracleEngine)
    with pytest.raises(ValueError):
        par.predict(x)                                                             # the Par model takes [images, process parameters]
    with pytest.raises(TypeError):
        K.create_cvt_model(H, W, 1)


def test_batch_larger_than_workspace_and_repeat_models_cpu():
    """_ensure_batch used an engine method that did not exist; seed=None must give every model its own streams."""
    K.configure(projection="dw_bn", cls_token=True)
    x, p, y = _data(8, 2)
    m = K.create_cvt_model(H, W, 1, 5, 1, max_batch=2, engine_factory=OracleEngine, seed=1)
    m.compile(optimizer=K.Adam(1e-3), loss="mean_squared_error", metrics=["mae"])
    m.count_params()                                                               # creates the engine at max_batch 2
    w0 = m.get_weights_dict()
    e0 = m.engine
    m._ensure_batch(8)                                                             # grows the workspace, weights carried over
    assert m.cfg.max_batch == 8 and m.engine is not e0
    assert all(np.array_equal(w0[k], v) for k, v in m.get_weights_dict().items())
    m.fit([x, p], y, epochs=1, batch_size=8, verbose=0)
    assert m.engine.optimizer_steps() == 1
    m2 = K.create_cvt_model(H, W, 1, 5, 1, max_batch=2, engine_factory=OracleEngine, seed=1)
    m2.compile(optimizer=K.Adam(1e-3), loss="mean_squared_error")
    m2.train_on_batch([x[:2], p[:2]], y[:2])
    with pytest.raises(ValueError):
        m2.fit([x, p], y, epochs=1, batch_size=8, verbose=0)                       # after training started the workspace cannot grow
    a, b = K.create_cvt_model(H, W, 1, 5, 1, engine_factory=OracleEngine), K.create_cvt_model(H, W, 1, 5, 1, engine_factory=OracleEngine)
    assert a.dropout_seed != b.dropout_seed
    assert not np.array_equal(a._rng.permutation(50), b._rng.permutation(50))


def test_gradcam_heatmap_cpu():
    K.configure(projection="dw_bn", cls_token=False)
    try:
        m = K.create_cvt_model(H, W, 1, 5, 1, engine_factory=OracleEngine, max_batch=2, seed=4)
        spec = m.engine.spec
        m.set_weights_dict(O.init_params(spec, 9, "parity"))
        x, p, _ = _data(3, 8)
        heat = K.make_gradcam_heatmap(x, p, m, "stage3_transformer")
        assert heat.shape == (3, 1, 1) or heat.shape[0] == 3
        for i in range(3):
            ref, _, _ = O.gradcam(spec, m.engine.params, x[i:i + 1, ..., None].astype(np.float32), p[i:i + 1].astype(np.float32), "s3.out")
            np.testing.assert_allclose(heat[i], ref[0], rtol=1e-5, atol=1e-6)
    finally:
        K.configure(projection="dw_bn", cls_token=True)


# =====================================================================================================================
# the same on the GPU
# =====================================================================================================================
@pytest.mark.gpu
@pytest.mark.parametrize("f64", [True, False])
def test_fit_trajectory_matches_oracle_gpu(tmp_path, f64):
    """fp32 mode, 4 epochs with shuffling, Philox dropout, a partial last batch, float64 (as the scripts pass) and float32 inputs: the
    per-epoch loss / mae / val_loss / val_mae equal the oracle-engine run of the same call sequence (same permutations, same masks)."""
    epochs = 4
    ref, _, _ = _script_sequence(OracleEngine, tmp_path, epochs, f64=f64)
    got, (x, p, y), (xv, pv, yv) = _script_sequence(T.CvtEngine, tmp_path, epochs, precision="fp32", f64=f64)
    try:
        _check_history(got, epochs)
        for k in ("loss", "mae", "val_loss", "val_mae"):
            np.testing.assert_allclose(got.history.history[k], ref.history.history[k], rtol=2e-3, err_msg=k)
        np.testing.assert_allclose(got.predict([xv, pv]), ref.predict([xv, pv]), rtol=2e-3)
    finally:
        got.close()


@pytest.mark.gpu
def test_script_sequence_tf32_52_epochs_gpu(tmp_path):
    """The full sequence on the tensor-core path, long enough for the x0.8 learning-rate step at epoch 50."""
    epochs = 52
    model, (x, p, y), (xv, pv, yv) = _script_sequence(T.CvtEngine, tmp_path, epochs, precision="tf32")
    try:
        _check_history(model, epochs)
        assert model.history.history["loss"][-1] < model.history.history["loss"][0]
        assert model.engine.optimizer_steps() == epochs * 3
        wpath = str(tmp_path / "w.h5")
        model.save_weights(wpath)
        m2 = K.create_cvt_model(H, W, 1, 5, 1, precision="tf32", max_batch=4)
        try:
            m2.load_weights(wpath)
            np.testing.assert_array_equal(m2.predict([xv, pv]), model.predict([xv, pv]))
            l1, a1 = model.evaluate([xv, pv], yv)
            assert l1 == pytest.approx(model.history.history["val_loss"][-1], rel=1e-5)
        finally:
            m2.close()
    finally:
        model.close()


@pytest.mark.gpu
@pytest.mark.parametrize("variant,cls", [("reference", False), ("gradcam", False), ("gradcam", True), ("model_plot", False)])
def test_gradcam_heatmap_gpu(variant, cls):
    """make_gradcam_heatmap (tools/grad_cam_CvT.py:422-481) on the models the tools build (configure(variant=...)) against oracle.gradcam."""
    K.configure(projection="dw_bn", cls_token=cls, variant=variant)
    try:
        m = K.create_cvt_model(H, W, 1, 5, 1, precision="fp32", max_batch=2, seed=4)
        spec = O.gradcam_spec(cls, "dw_bn", image_h=H, image_w=W) if variant == "gradcam" else O.ref_spec("dw_bn", cls, 5, 1, H, W)
        if variant == "model_plot":
            spec.head = "mean"
        params = O.init_params(spec, 9, "parity")
        m.set_weights_dict(params)
        x, p, _ = _data(3, 8)
        heat = K.make_gradcam_heatmap(x, p, m, "stage3_transformer")
        for i in range(3):
            ref, feat, g = O.gradcam(spec, params, x[i:i + 1, ..., None].astype(np.float32), p[i:i + 1].astype(np.float32), "s3.out")
            np.testing.assert_allclose(heat[i], ref[0], rtol=2e-3, atol=2e-4)
        m.close()
    finally:
        K.configure(projection="dw_bn", cls_token=True, variant="reference")


@pytest.mark.gpu
def test_gradcam_on_the_cls_head_fails_like_the_reference():
    """With the cls-token head of models/CvT(Par).py:326-329 the prediction never reads the spatial map: tape.gradient returns None in the
    reference (tools/grad_cam_CvT.py:454 then raises); here the zero gradient is reported instead of a NaN heat-map."""
    K.configure(projection="dw_bn", cls_token=True, variant="reference")
    m = K.create_cvt_model(H, W, 1, 5, 1, precision="fp32", max_batch=2, seed=4)
    x, p, _ = _data(2, 8)
    with pytest.raises(ValueError, match="No gradients"):
        K.make_gradcam_heatmap(x, p, m, "stage3_transformer")
    m.close()


@pytest.mark.parametrize("variant,cls,proj", [("gradcam", False, "dw_bn"), ("gradcam", True, "linear"), ("model_plot", True, "dw_bn"),
                                               ("model_plot", False, "avg")])
def test_configure_variant_builds_the_tools_models(variant, cls, proj):
    """keras_api.configure(variant=...) (CPU, oracle-backed engine): create_cvt_model then builds the model of tools/grad_cam_CvT.py:315-349
    / tools/model_plot.py:254-300 - parameter table, one fit epoch and predict go through."""
    K.configure(projection=proj, cls_token=cls, variant=variant)
    try:
        m = K.create_cvt_model(H, W, 1, 5, 1, precision="fp32", max_batch=4, seed=1, engine_factory=OracleEngine)
        names = [p["name"] for p in m.engine.param_infos()]
        if variant == "gradcam":
            assert m.cfg.head == "mean" and not m.cfg.apply_proj_qkv and m.cfg.attn_drop_rate == pytest.approx(0.1)
            assert all(st.with_cls_token == cls for st in m.cfg.stages)
            assert not any("/proj_q/" in n or n.startswith("head_norm") for n in names)
            assert any("q_proj/dense/kernel" in n for n in names) == (proj == "linear")
        else:
            assert m.cfg.head == ("cls" if cls else "mean") and m.cfg.apply_proj_qkv
            assert not any(n.startswith("head_norm") for n in names) and any("/proj_q/" in n for n in names)
        x, p, y = _data(6, 3)
        m.compile(optimizer=K.Adam(learning_rate=1e-3), loss="mean_squared_error", metrics=["mae"])
        h = m.fit([x, p], y, epochs=1, batch_size=4, verbose=0)
        assert np.isfinite(h.history["loss"][0])
        assert m.predict([x, p]).shape == (6, 1)
        m.close()
    finally:
        K.configure(projection="dw_bn", cls_token=True, variant="reference")
