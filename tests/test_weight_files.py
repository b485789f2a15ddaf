"""Keras-2 legacy HDF5 weight files (models/CvT(Par).py:489 save_weights, models/CvT_test(Par).py:513 load_weights) without h5py:
container round trips of transformer_stm_b200/h5lite.py and the structural name map of SURVEY 9.3."""
import struct

import numpy as np
import pytest

from tests.fake_engine import OracleEngine
from transformer_stm_b200 import h5lite
from transformer_stm_b200 import keras_api as K


def test_container_roundtrip_many_links_nested_groups_and_attributes(tmp_path):
    rng = np.random.default_rng(0)
    root = h5lite.Group()
    root.attrs["backend"] = np.bytes_(b"tensorflow")
    root.attrs["numbers"] = np.arange(5, dtype=np.int64)
    for i in range(70):                                           # > 8 links: several symbol-table nodes under one B-tree node
        g = root.require("layer_%03d/inner/deeper" % i)
        g["kernel:0"] = rng.standard_normal((3, 4, 5)).astype(np.float32)
        g["bias:0"] = rng.standard_normal((5,)).astype(np.float32)
        g["empty"] = np.zeros((0, 4), np.float32)
        g["scalar"] = np.float32(i)
        g.attrs["weight_names"] = np.asarray([b"a/kernel:0", b"a/bias:0"])
    p = str(tmp_path / "t.h5")
    h5lite.write_h5(p, root)
    raw = open(p, "rb").read()
    assert raw[:8] == h5lite.SIGNATURE and raw[8] == 0            # superblock version 0, as h5py's default writes it
    assert struct.unpack_from("<Q", raw, 40)[0] == len(raw)       # end-of-file address
    back = h5lite.read_h5(p)
    assert bytes(back.attrs["backend"]) == b"tensorflow" and np.array_equal(back.attrs["numbers"], np.arange(5))
    assert sorted(back) == sorted(root)
    for name in root:
        a, b = root[name]["inner"]["deeper"], back[name]["inner"]["deeper"]
        for k in ("kernel:0", "bias:0", "empty", "scalar"):
            assert np.array_equal(np.asarray(a[k]), b[k]) and b[k].dtype == np.float32 and b[k].shape == np.asarray(a[k]).shape
        assert list(b.attrs["weight_names"]) == [b"a/kernel:0", b"a/bias:0"]
    with pytest.raises(h5lite.H5Error):
        open(str(tmp_path / "bad.h5"), "wb").write(b"not hdf5" * 100)
        h5lite.read_h5(str(tmp_path / "bad.h5"))


def test_long_attribute_lists_are_split_like_keras(tmp_path):
    names = [("layer_with_a_rather_long_name_%05d" % i) for i in range(2500)]          # > 64512 bytes as one attribute
    layers = [(n, []) for n in names]
    layers[3] = (names[3], [(names[3] + "/kernel:0", np.ones((2, 2), np.float32))])
    p = str(tmp_path / "many.h5")
    h5lite.save_keras_weights(p, layers)
    root = h5lite.read_h5(p)
    assert "layer_names" not in root.attrs and "layer_names0" in root.attrs and "layer_names1" in root.attrs
    back = h5lite.load_keras_weights(p)
    assert list(back) == names and back[names[3]][0][1].shape == (2, 2)


def _model(cls=True, method="dw_bn", proc=5):
    K.configure(projection=method, cls_token=cls)
    try:
        if proc:
            return K.create_cvt_model(32, 32, 1, proc, 1, engine_factory=OracleEngine, seed=1, max_batch=2)
        return K.create_cvt_model(32, 32, 1, 1, engine_factory=OracleEngine, seed=1, max_batch=2)
    finally:
        K.configure(projection="dw_bn", cls_token=True)


def test_layout_is_keras_order_trainable_first_moving_statistics_last():
    m = _model()
    lay = dict(m.keras_layout())
    assert list(dict(m.keras_layout()))[:3] == ["Image_inputs", "stage1_ConvEmbed", "stage1_transformer"]
    t3 = [c for _, c in lay["stage3_transformer"]]
    assert t3[0] == "stage3_transformer/cls_token" and t3[1:3] == ["stage3_transformer/norm1/gamma", "stage3_transformer/norm1/beta"]
    assert t3[3:6] == ["stage3_transformer/attn/q_proj/depthwise_kernel", "stage3_transformer/attn/q_proj/bn/gamma", "stage3_transformer/attn/q_proj/bn/beta"]
    assert t3[-6:] == ["stage3_transformer/attn/%s_proj/bn/%s" % (b, s) for b in "qkv" for s in ("moving_mean", "moving_variance")]
    assert t3[-8:-6] == ["stage3_transformer/mlp/fc2/kernel", "stage3_transformer/mlp/fc2/bias"]
    # every parameter of the model appears exactly once
    flat = [c for _, ws in m.keras_layout() for _, c in ws]
    assert sorted(flat) == sorted(p["name"] for p in m.engine.param_infos())
    # keras numbers anonymous Dense layers in creation order: proj_q/k/v = dense, dense_1, dense_2, proj = dense_3, Mlp = dense_4, dense_5
    s1 = [k for k, _ in lay["stage1_transformer"]]
    assert "stage1_transformer/conv_attention/dense_3/kernel:0" in s1 and "stage1_transformer/sequential/dense_5/bias:0" in s1


@pytest.mark.parametrize("cls,method,proc", [(True, "dw_bn", 5), (False, "avg", 0), (True, "linear", 5)])
def test_h5_weights_roundtrip_between_models(tmp_path, cls, method, proc):
    m = _model(cls, method, proc)
    rng = np.random.default_rng(5)
    m.set_weights_dict({p["name"]: rng.standard_normal(p["shape"]).astype(np.float32) for p in m.engine.param_infos()})
    p = str(tmp_path / "w.h5")
    m.save_weights(p)
    m2 = _model(cls, method, proc)
    m2.load_weights(p)
    a, b = m.get_weights_dict(), m2.get_weights_dict()
    assert all(np.array_equal(a[k], b[k]) for k in a)
    # the file lists one group per keras layer with trainable-then-non-trainable weight_names
    f = h5lite.load_keras_weights(p)
    assert "Image_inputs" in f and f["Image_inputs"] == []
    if method == "dw_bn":
        assert [n.rsplit("/", 1)[-1] for n, _ in f["stage2_transformer"]][-2:] == ["moving_mean:0", "moving_variance:0"]
    # a different architecture refuses the file, naming the first mismatch (as Keras does)
    other = _model(not cls, method, proc)
    with pytest.raises(ValueError):
        other.load_weights(p)


@pytest.mark.parametrize("variant,method,cls", [("gradcam", "dw_bn", False), ("gradcam", "linear", True), ("model_plot", "dw_bn", True), ("model_plot", "avg", False)])
def test_h5_weights_roundtrip_of_the_tools_model_variants(tmp_path, variant, method, cls):
    """tools/grad_cam_CvT.py:315-349 / tools/model_plot.py:254-300: proj_q/k/v built but never applied (Keras never creates their weights), a Dense
    inside the 'linear' Projection, heads without a LayerNormalization.  The layout used to list the reference model's weights only and
    save_weights raised KeyError on these models (found by running the tool script itself, tests/test_reference_script_dropin.py)."""
    def make(seed):
        K.configure(projection=method, cls_token=cls, variant=variant)
        try:
            return K.create_cvt_model(32, 32, 1, 5, 1, engine_factory=OracleEngine, seed=seed, max_batch=2)
        finally:
            K.configure(projection="dw_bn", cls_token=True, variant="reference")
    m = make(1)
    rng = np.random.default_rng(6)
    m.set_weights_dict({p["name"]: rng.standard_normal(p["shape"]).astype(np.float32) for p in m.engine.param_infos()})
    flat = [c for _, ws in m.keras_layout() for _, c in ws]
    assert sorted(flat) == sorted(p["name"] for p in m.engine.param_infos())      # everything the engine holds, once
    p = str(tmp_path / "w.h5")
    m.save_weights(p)
    m2 = make(2)
    m2.load_weights(p)
    a, b = m.get_weights_dict(), m2.get_weights_dict()
    assert a.keys() == b.keys() and all(np.array_equal(a[k], b[k]) for k in a)
    ref = _model()                                                                 # the reference model is another architecture: refused
    with pytest.raises(ValueError):
        ref.load_weights(p)


def test_loading_matches_by_position_not_by_auto_numbered_names(tmp_path):
    """A file written by a Keras process whose global name counters stood elsewhere: dense_17.., layer_normalization_41."""
    m = _model()
    rng = np.random.default_rng(6)
    m.set_weights_dict({p["name"]: rng.standard_normal(p["shape"]).astype(np.float32) for p in m.engine.param_infos()})
    w = m.get_weights_dict()
    layers = []
    for ln, ws in m.keras_layout():
        if ln.startswith("layer_normalization"):
            ln2 = "layer_normalization_41"
            layers.append((ln2, [(kn.replace(ln, ln2), w[cn]) for kn, cn in ws]))
        else:
            layers.append((ln, [(kn.replace("dense_", "dense_1").replace("/dense/", "/dense_10/"), w[cn]) for kn, cn in ws]))
    layers.insert(5, ("tf.reshape", []))
    p = str(tmp_path / "foreign.h5")
    h5lite.save_keras_weights(p, layers, keras_version="2.8.0")
    m2 = _model()
    m2.load_weights(p)
    b = m2.get_weights_dict()
    assert all(np.array_equal(w[k], b[k]) for k in w)


def test_npz_archive_still_loads(tmp_path):
    m = _model()
    p = str(tmp_path / "w.weights")
    m.save_weights(p)
    m2 = _model()
    m2.load_weights(p)
    a, b = m.get_weights_dict(), m2.get_weights_dict()
    assert all(np.array_equal(a[k], b[k]) for k in a)
    with pytest.raises(FileNotFoundError):
        m2.load_weights(str(tmp_path / "missing.h5"))
