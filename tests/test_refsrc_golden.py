"""oracle/cvt_oracle.py against the committed vectors of tests/golden/refsrc_*.npz: outputs of the reference's own model source executed over
oracle/mini_keras.py (generated in the build container by tests/golden/make_refsrc_golden.py, where /root/reference exists).  Runs anywhere -
also on the GPU box, where the reference tree is absent: the oracle the CUDA parity tests compare against is the one these vectors pin."""
import glob
import os

import numpy as np
import pytest

from oracle import cvt_oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
FILES = sorted(glob.glob(os.path.join(HERE, "golden", "refsrc_*.npz")))


def test_the_vectors_are_there():
    assert len(FILES) >= 8


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f)[7:-4] for f in FILES])
def test_oracle_reproduces_the_reference_source_vectors(path):
    d = np.load(path)
    script, kind, method, cls, h, w, proc_dim, nc, batch, seed = [str(x) for x in d["meta"]]
    cls, h, w, proc_dim, nc, batch, seed = cls == "True", int(h), int(w), int(proc_dim), int(nc), int(batch), int(seed)
    if kind == "gradcam":
        spec = O.gradcam_spec(cls, method, proc_dim=proc_dim, num_classes=nc, image_h=h, image_w=w)
    else:
        spec = O.ref_spec(method, cls, proc_dim=proc_dim, num_classes=nc, image_h=h, image_w=w)
        if kind == "model_plot":
            spec.head = "cls" if cls else "mean"
    params = O.init_params(spec, seed, "parity")
    img, proc, y = O.synthetic_batch(spec, batch, seed % 97)
    pred, _ = O.forward(spec, params, img, proc if proc_dim else None, training=False)
    np.testing.assert_allclose(pred, d["pred"], rtol=1e-9, atol=1e-9)
    masks = O.make_dropout_masks(spec, batch, seed=seed, step=0)
    r = O.loss_and_grads(spec, params, img, proc if proc_dim else None, y, masks)
    assert r["loss"] == pytest.approx(float(d["loss"]), rel=1e-11)
    np.testing.assert_allclose(r["pred"], d["train_pred"], rtol=1e-9, atol=1e-9)
    keys = [k for k in d.files if k.startswith("g/")]
    assert sorted(k[2:] for k in keys) == sorted(r["grads"])                      # every trainable tensor, nothing else
    scale = max(float(np.abs(g).max()) for g in r["grads"].values())
    for k in keys:
        g = np.asarray(r["grads"][k[2:]], dtype=np.float64).ravel()
        want = d[k]
        got = np.concatenate([[g.sum(), (g * g).sum()], g[:16]])
        assert np.abs(got[2:] - want[2:]).max() <= 1e-9 * scale, k
        assert abs(got[0] - want[0]) <= 1e-8 * scale * max(1.0, np.sqrt(g.size)), k
        assert abs(got[1] - want[1]) <= 1e-8 * max(want[1], scale * scale * 1e-6), k
