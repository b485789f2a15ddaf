"""The drop-in claim of INTEGRATION.md, executed: the reference's OWN functions - `preprocess_data`, `lr_scheduler`, `train_and_save_model`
of models/CvT(Par).py, CvT(Par)(many).py and CvT(Img).py, `test_and_save_results` and `save_predictions_to_excel` of the three CvT_test scripts and of tools/grad_cam_CvT.py -
are taken from /root/reference as they lie there (parsed, compiled and run from those files: nothing of them is copied into this repository)
and run against transformer_stm_b200.keras_api in place of TensorFlow / Keras: real cv2 on JPEG files written to a temporary tree, real pandas
frames for the two Excel sheets, real sklearn scalers and metrics; `create_cvt_model`, `keras.optimizers.Adam`,
`keras.callbacks.LearningRateScheduler` and `TensorBoard` come from keras_api.  Two stand-ins for packages this image lacks: pandas' Excel
writer (openpyxl) goes through xlsxlite - the writer keras_api itself uses - and matplotlib's `plt` is a no-op (plots are out of scope).
The engine underneath is the oracle-backed CPU stand-in of tests/fake_engine.py (host logic only; the GPU twin of these call sequences is
tests/test_keras_surface.py).  Runs only where the reference tree exists (this container): the GPU box does not have it."""
import ast
import os
import types

import numpy as np
import pytest

REF = "/root/reference/models"
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not present (it does not travel to the GPU box)")

H = W = 32
LAYERS, GROUPS, PIECES = 2, 2, 5
FREQ = "50HZ_Bm"


def reference_functions(script, *names):
    """Code object holding the named top-level functions of one reference script, unmodified."""
    path = os.path.join(REF, script)
    tree = ast.parse(open(path, encoding="utf-8").read(), path)
    nodes = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in names]
    assert sorted(n.name for n in nodes) == sorted(names), [n.name for n in nodes]
    return compile(ast.Module(body=nodes, type_ignores=[]), path, "exec")


class World:
    """The file tree, the two Excel sheets (as pandas frames) and the name space a reference script sees after its import block."""

    def __init__(self, tmp_path, proc_dim=5):
        import cv2
        import pandas as pd
        from sklearn.metrics import mean_absolute_error, mean_squared_error, r2_score
        from sklearn.preprocessing import StandardScaler
        from transformer_stm_b200 import keras_api as K, xlsxlite
        from tests.fake_engine import OracleEngine

        class Frame(pd.DataFrame):
            @property
            def _constructor(self):
                return Frame

            def to_excel(self, path, index=True, **kw):
                try:
                    return super().to_excel(path, index=index, **kw)
                except (ImportError, ModuleNotFoundError):                        # openpyxl is not installed here
                    assert index is False
                    xlsxlite.write_xlsx(str(path), {str(c): [x.item() if hasattr(x, "item") else x for x in self[c]] for c in self.columns})

        self.K, self.xlsxlite, self.tmp, self.proc_dim = K, xlsxlite, tmp_path, proc_dim
        self.script_dir = tmp_path / "models"
        self.script_dir.mkdir()
        rng = np.random.default_rng(0)
        for g in range(1, GROUPS + 1):                                            # ../data/circle(340x345)/trail{group}_{piece:02d}/layer_{i:02d}.jpg
            for p in range(1, PIECES + 1):
                d = tmp_path / "data" / "circle(340x345)" / ("trail%d_%02d" % (g, p))
                d.mkdir(parents=True)
                for i in range(1, LAYERS + 1):
                    assert cv2.imwrite(str(d / ("layer_%02d.jpg" % i)), rng.integers(0, 256, (44, 40, 3), dtype=np.uint8))
        labels = 1000.0 + 100.0 * rng.standard_normal(GROUPS * PIECES)
        labels[3] = np.nan                                                        # a piece without a measurement: the scripts drop it
        self.excel_data = pd.DataFrame({FREQ: labels})
        self.excel_process = pd.DataFrame({k: rng.random(GROUPS) * s for k, s in
                                           (("氧濃度", 1.0), ("雷射掃描速度", 900.0), ("雷射功率", 200.0), ("線間距", 0.1), ("能量密度", 60.0))})
        self.made = []

        def create(*a, **kw):
            m = K.create_cvt_model(*a, engine_factory=OracleEngine, max_batch=4, seed=3 + len(self.made), **kw)
            self.made.append(m)
            return m

        noop = lambda *a, **kw: None
        K.configure(projection="dw_bn", cls_token=True)
        self.ns = dict(
            np=np, os=os, cv2=cv2, StandardScaler=StandardScaler, r2_score=r2_score, mean_squared_error=mean_squared_error,
            mean_absolute_error=mean_absolute_error,
            pd=types.SimpleNamespace(isnull=pd.isnull, DataFrame=Frame, concat=pd.concat),
            plt=types.SimpleNamespace(scatter=noop, plot=noop, title=noop, xlabel=noop, ylabel=noop, legend=noop, savefig=noop, clf=noop),
            keras=types.SimpleNamespace(optimizers=types.SimpleNamespace(Adam=K.Adam), callbacks=types.SimpleNamespace(LearningRateScheduler=K.LearningRateScheduler)),
            TensorBoard=K.TensorBoard, create_cvt_model=create,
            # the module-level constants of the scripts (models/CvT(Par).py:25-53), scaled down
            freq=FREQ, projection_method="dw_bn", cls_token_switch=True, image_layers=LAYERS, image_height=H, image_width=W,
            num_channels=1, num_classes=1, batch_size=4, train_epochs=2, script_dir=str(self.script_dir))

    def run(self, script, *names):
        exec(reference_functions(script, *names), self.ns)
        return [self.ns[n] for n in names]

    def result(self, *parts):
        return self.tmp.joinpath("Result", *parts)


def check_history_sheet(w, model, path):
    hist = model.history.history
    assert all(len(hist[k]) == 2 for k in ("loss", "mae", "val_loss", "val_mae")) and np.isfinite(hist["loss"]).all()
    sheet = w.xlsxlite.read_xlsx(str(path))
    assert list(sheet)[0] == "epoch" and [int(e) for e in sheet["epoch"]] == [1, 2]
    assert {"loss", "mae", "val_loss", "val_mae"} <= set(sheet)
    np.testing.assert_allclose(np.asarray(sheet["loss"], dtype=np.float64), np.asarray(hist["loss"], dtype=np.float64), rtol=1e-6)


def test_cvt_par_training_script_then_the_test_script(tmp_path):
    w = World(tmp_path)
    preprocess, train = w.run("CvT(Par).py", "preprocess_data", "train_and_save_model", "lr_scheduler")[:2]
    labels_dict, proc_scaled, images, valid_dict, count = preprocess(w.excel_data, w.excel_process, 1, GROUPS, 1, PIECES, LAYERS, H, W)
    assert count == GROUPS * PIECES and len(valid_dict) == GROUPS * PIECES - 1
    assert images.shape == (len(valid_dict) * LAYERS, H, W) and images.dtype == np.float64 and 0.0 <= images.min() and images.max() <= 1.0
    assert proc_scaled.shape == (len(images), 5) and labels_dict.shape == (len(images),)

    train(FREQ, labels_dict, proc_scaled, images, valid_dict, count)              # models/CvT(Par).py:431-494, as written

    assert len(w.made) == 1
    model = w.made[0]
    wpath = w.result("Weight", "Images & Parameters", "cvt_model_weights_%s_dw_bn_clsTrue.h5" % FREQ)
    assert wpath.exists()
    check_history_sheet(w, model, w.result("Records", "Images & Parameters", "cvt_records_%s_dw_bn_clsTrue.xlsx" % FREQ))

    # the companion script: models/CvT_test(Par).py:480-558 (new model, load_weights, predict, sklearn metrics, the sheet tools/Compare_r.py reads)
    test_results, _ = w.run("CvT_test(Par).py", "test_and_save_results", "save_predictions_to_excel")
    test_results(FREQ, labels_dict, proc_scaled, images, valid_dict, count)
    assert len(w.made) == 2
    sheet = w.xlsxlite.read_xlsx(str(w.result("Excel", "Images & Parameters", "Predictions_Metrics_%s.xlsx" % FREQ)))
    assert list(sheet) == ["Predictions", "Actual", "Errors(%)", "Train mounts", "Test mounts", "R2 Score", "MSE", "MAE"]      # tools/Compare_r.py:21-24
    n_val = 2 * LAYERS                                                            # the first valid piece of each group of five
    assert len(sheet["Predictions"]) == n_val and int(sheet["Test mounts"][0]) == n_val and int(sheet["Train mounts"][0]) == len(images) - n_val
    assert all(v is None for v in sheet["MSE"][1:])                               # metrics in the first row only
    # ... and the predictions in it are those of the trained model (different initial weights, same file)
    val_idx = [i * LAYERS + k for i, v in enumerate(valid_dict) if v in (0, 5) for k in range(LAYERS)]
    want = model.predict([images[val_idx], proc_scaled[val_idx]]).ravel()
    np.testing.assert_allclose(np.asarray(sheet["Predictions"], dtype=np.float64), want, rtol=1e-6)
    np.testing.assert_allclose(np.asarray(sheet["Actual"], dtype=np.float64), labels_dict[val_idx], rtol=1e-12)

    # the consumer of that sheet, tools/Compare_r.py:12-62 as written: reads 'R2 Score' of Sheet1 and plots it (pandas' reader needs openpyxl too:
    # read_excel goes through xlsxlite; the plot calls are recorded instead of drawn)
    import pandas
    from sklearn.metrics import r2_score
    plotted = []
    w.ns["pd"].read_excel = lambda path, sheet_name=None: pandas.DataFrame({k: pandas.Series(v, dtype="float64") for k, v in w.xlsxlite.read_xlsx(str(path)).items()})
    noop = lambda *a, **kw: None
    w.ns["plt"] = types.SimpleNamespace(figure=noop, plot=lambda x, yv, **kw: plotted.append((list(x), list(yv), kw.get("label"))), title=noop, xlabel=noop,
                                        ylabel=noop, legend=noop, grid=noop, savefig=noop, close=noop)
    from collections import OrderedDict
    w.ns["OrderedDict"] = OrderedDict
    w.ns["script_dir"] = str(tmp_path / "tools")
    (tmp_path / "tools").mkdir()
    collect, = w.run(os.path.join("..", "tools", "Compare_r.py"), "collect_data_and_plot")
    collect(["CvT"], [FREQ.split("_")[1]], [FREQ])
    assert len(plotted) == 1 and plotted[0][0] == [FREQ.split("_")[0]] and plotted[0][2] == "CvT"
    assert plotted[0][1][0] == pytest.approx(r2_score(labels_dict[val_idx], want), rel=1e-6)


def test_cvt_par_many_script_repeats(tmp_path):
    w = World(tmp_path)
    preprocess, train = w.run("CvT(Par)(many).py", "preprocess_data", "train_and_save_model", "lr_scheduler")[:2]
    for t in (1, 2):                                                              # models/CvT(Par)(many).py:501-505: eight independent repeats
        w.ns["time"] = t
        data = preprocess(w.excel_data, w.excel_process, 1, GROUPS, 1, PIECES, LAYERS, H, W)
        train(FREQ, *data)
        assert w.result("Weight", "Images & Parameters", "cvt_model_weights_%s_%d_dw_bn_clsTrue.h5" % (FREQ, t)).exists()
        check_history_sheet(w, w.made[-1], w.result("Records", "Images & Parameters", "cvt_records_%s_%d_dw_bn_clsTrue.xlsx" % (FREQ, t)))
    assert len(w.made) == 2
    # models/CvT_test(Par)(many).py:450-530 for the second repeat
    test_results, _ = w.run("CvT_test(Par)(many).py", "test_and_save_results", "save_predictions_to_excel")
    test_results(FREQ, *data)
    sheet = w.xlsxlite.read_xlsx(str(w.result("Excel", "Images & Parameters", "Predictions_Metrics_%s_2.xlsx" % FREQ)))
    labels_dict, proc_scaled, images, valid_dict, _ = data
    val_idx = [i * LAYERS + k for i, v in enumerate(valid_dict) if v in (0, 5) for k in range(LAYERS)]
    np.testing.assert_allclose(np.asarray(sheet["Predictions"], dtype=np.float64), w.made[1].predict([images[val_idx], proc_scaled[val_idx]]).ravel(), rtol=1e-6)


def test_cvt_img_training_script(tmp_path):
    w = World(tmp_path, proc_dim=0)
    preprocess, train = w.run("CvT(Img).py", "preprocess_data", "train_and_save_model", "lr_scheduler")[:2]
    labels_dict, images, valid_dict, count = preprocess(w.excel_data, 1, GROUPS, 1, PIECES, LAYERS, H, W)
    train(FREQ, labels_dict, images, valid_dict, count)                           # the four-argument create_cvt_model, fit([x_train], ...)
    model = w.made[0]
    assert model.cfg.proc_dim == 0
    assert w.result("Weight", "Images", "cvt_model_weights_%s_dw_bn_clsTrue.h5" % FREQ).exists()
    check_history_sheet(w, model, w.result("Records", "Images", "cvt_records_%s_dw_bn_clsTrue.xlsx" % FREQ))
    assert model.predict([images[:3]]).shape == (3, 1)
    # models/CvT_test(Img).py:417-480
    test_results, _ = w.run("CvT_test(Img).py", "test_and_save_results", "save_predictions_to_excel")
    test_results(FREQ, labels_dict, images, valid_dict, count)
    sheet = w.xlsxlite.read_xlsx(str(w.result("Excel", "Images", "Predictions_Metrics_%s.xlsx" % FREQ)))
    assert list(sheet) == ["Predictions", "Actual", "Errors(%)", "Train mounts", "Test mounts", "R2 Score", "MSE", "MAE"]
    val_idx = [i * LAYERS + k for i, v in enumerate(valid_dict) if v in (0, 5) for k in range(LAYERS)]
    np.testing.assert_allclose(np.asarray(sheet["Predictions"], dtype=np.float64), model.predict(images[val_idx]).ravel(), rtol=1e-6)


def test_gradcam_tool_script(tmp_path):
    """tools/grad_cam_CvT.py:484-610 as written (validation split, load_weights, cv2 colour maps and overlays, predict, sklearn metrics), with the
    one function that is TensorFlow through and through - make_gradcam_heatmap (:422-481: tf.data, tf.GradientTape) - supplied by keras_api."""
    w = World(tmp_path)
    K = w.K
    from tqdm import tqdm
    K.configure(projection="dw_bn", cls_token=False, variant="gradcam")            # the constants of tools/grad_cam_CvT.py:25-31 and its model variant
    try:
        calls = []

        def heatmaps(img_array, proc_array, model, last_conv_layer_name, **kw):
            h = K.make_gradcam_heatmap(img_array, proc_array, model, last_conv_layer_name, **kw)
            calls.append((last_conv_layer_name, h))
            return h

        noop = lambda *a, **kw: None
        w.ns.update(cls_token_switch=False, last_conv_layer_name="stage3_transformer", make_gradcam_heatmap=heatmaps, tqdm=tqdm,
                    plt=types.SimpleNamespace(figure=noop, subplot=noop, imshow=noop, axis=noop, title=noop, savefig=noop, clf=noop, close=noop))
        preprocess, run = w.run(os.path.join("..", "tools", "grad_cam_CvT.py"), "preprocess_data", "test_and_save_results")
        labels_dict, proc_scaled, images, valid_dict, count = preprocess(w.excel_data, w.excel_process, 1, GROUPS, 1, PIECES, LAYERS, H, W)
        # the weights the tool loads: a model of the same variant, saved under the name the tool expects (:513)
        trained = w.ns["create_cvt_model"](H, W, 1, 5, 1)
        wdir = w.result("Weight", "Images & Parameters")
        wdir.mkdir(parents=True)
        trained.save_weights(str(wdir / ("cvt_model_weights_%s.h5" % FREQ)))

        run(FREQ, labels_dict, proc_scaled, images, valid_dict, count)

        assert len(calls) == 1 and calls[0][0] == "stage3_transformer"
        h = calls[0][1]
        n_val = 2 * LAYERS
        assert h.shape[0] == n_val and h.ndim == 3 and np.nanmax(h) <= 1.0 + 1e-6      # one normalised map per validation image, [n, h3, w3]
        assert w.result("Plots", "Heatmap").is_dir()
        assert len(w.made) == 2 and w.made[1].cfg.stages[0].with_cls_token == w.made[0].cfg.stages[0].with_cls_token
    finally:
        K.configure(projection="dw_bn", cls_token=True, variant="reference")
