"""Input pipeline (models/CvT(Par).py:418-423): cv2.resize(INTER_LINEAR) + cv2.cvtColor(BGR2GRAY) + / 255 on the device,
the resident dataset, and fit epochs as permutations gathered on the device.  Integer work: bit-exact.
Goldens (tests/golden/preprocess_*.npz) are outputs of OpenCV itself (make_preprocess_golden.py)."""
import glob
import os

import numpy as np
import pytest

from oracle import preprocess_oracle as PO

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = sorted(glob.glob(os.path.join(HERE, "golden", "preprocess_*.npz")))


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p)[11:-4] for p in CASES])
def test_oracle_reproduces_opencv_golden(path):
    g = np.load(path)
    r = PO.resize_linear_u8(g["src"], 128, 128)
    assert np.array_equal(r, g["resized"])
    assert np.array_equal(PO.bgr2gray_u8(r), g["gray"])
    assert np.array_equal(PO.preprocess(g["src"], 128, 128), g["out"].astype(np.float32))


def test_oracle_against_live_opencv_when_present():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(5)
    for (h, w, dh, dw) in [(345, 340, 128, 128), (97, 211, 64, 48), (128, 128, 128, 128), (60, 60, 128, 128), (345, 340, 340, 345)]:
        im = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        ref = cv2.cvtColor(cv2.resize(im, (dw, dh)), cv2.COLOR_BGR2GRAY) / 255.0
        assert np.array_equal(PO.preprocess(im, dh, dw), ref.astype(np.float32)), (h, w, dh, dw)


@pytest.mark.gpu
def test_device_preprocess_is_bit_exact_and_dataset_roundtrips():
    import transformer_stm_b200 as T
    eng = T.CvtEngine(T.CvtConfig.reference(image_h=128, image_w=128, max_batch=4))
    try:
        eng.dataset_alloc(2, 16)
        off = 0
        for path in CASES:                                   # different source sizes into one dataset, like the real folders
            g = np.load(path)
            eng.dataset_write_u8(2, off, np.stack([g["src"], g["src"][::-1].copy()]))
            img, _, _ = eng.dataset_read(2, off, 2)
            assert np.array_equal(img[0, ..., 0], g["out"].astype(np.float32)), os.path.basename(path)
            assert np.array_equal(img[1, ..., 0], PO.preprocess(g["src"][::-1], 128, 128))
            off += 2
        # grey single-channel sources skip the colour conversion
        g = np.load(CASES[0])
        eng.dataset_write_u8(2, off, g["src"][None, :, :, 0])
        img, _, _ = eng.dataset_read(2, off, 1)
        assert np.array_equal(img[0, ..., 0], PO.preprocess(g["src"][:, :, 0], 128, 128))
        # float64 / float32, pageable host arrays -> resident rows
        rng = np.random.default_rng(0)
        x64 = rng.random((5, 128, 128)); p64 = rng.standard_normal((5, 5)); y64 = 1000 + rng.standard_normal(5)
        eng.dataset_write(2, 9, x64, p64, y64)
        img, proc, y = eng.dataset_read(2, 9, 5)
        assert np.array_equal(img[..., 0], x64.astype(np.float32)) and np.array_equal(proc, p64.astype(np.float32))
        assert np.array_equal(y[:, 0], y64.astype(np.float32))
        with pytest.raises(T.CvtError):
            eng.dataset_write(2, 14, x64, p64, y64)          # past the end
        with pytest.raises(T.CvtError):
            eng.train_epoch_resident(3, None, 4, 1e-3)        # empty slot
    finally:
        eng.close()


@pytest.mark.gpu
@pytest.mark.parametrize("precision", ["fp32", "tf32"])
def test_stream_resident_and_host_epochs_walk_the_same_trajectory(precision):
    """One shuffled epoch three ways - the caller materialises the shuffled arrays (round-1 path), the library gathers host rows by
    index into its pinned ring (and fills the resident set), the library gathers resident rows on the device - gives the same
    per-batch losses, and the resident set holds the data at its own indices."""
    import transformer_stm_b200 as T
    from tests.parity_utils import make_pair, rel_err
    n, bs = 22, 4
    spec, e1, params, img, proc, y = make_pair(batch=n, precision=precision, hw=(32, 32), drop_rate=0.1, max_batch=bs)
    engines = [e1] + [T.CvtEngine(e1.cfg) for _ in range(2)]
    try:
        for e in engines:
            e.set_params(params); e.set_dropout("philox", 7)
        order = np.random.default_rng(3).permutation(n).astype(np.int32)
        x64, p64, y64 = img.astype(np.float64)[..., 0], proc.astype(np.float64), y.astype(np.float64)[:, 0]
        l0, a0 = engines[0].train_epoch(img[order], proc[order], y[order], bs, 1e-3)
        engines[1].dataset_alloc(0, n)
        l1, a1 = engines[1].train_epoch_stream(x64, p64, y64, order, bs, 1e-3, fill_slot=0)          # float64, pageable, as the scripts pass
        ri, rp, ry = engines[1].dataset_read(0, 0, n)
        assert np.array_equal(ri, img) and np.array_equal(rp, proc) and np.array_equal(ry, y)
        engines[2].dataset_alloc(0, n)
        engines[2].dataset_write(0, 0, img, proc, y)
        l2, a2 = engines[2].train_epoch_resident(0, order, bs, 1e-3)
        tol = 1e-5 if precision == "fp32" else 2e-3
        assert len(l0) == len(l1) == len(l2) == 6
        assert rel_err(l1, l0) <= tol and rel_err(l2, l0) <= tol and rel_err(a1, a0) <= tol and rel_err(a2, a0) <= tol
        # a second epoch from the set the stream epoch left behind == the same epoch on the uploaded set
        order2 = np.random.default_rng(4).permutation(n).astype(np.int32)
        l1b, _ = engines[1].train_epoch_resident(0, order2, bs, 1e-3)
        l2b, _ = engines[2].train_epoch_resident(0, order2, bs, 1e-3)
        assert rel_err(l1b, l2b) <= (1e-4 if precision == "fp32" else 5e-3)
        # evaluate / predict from the resident set == from host arrays
        le, ae = engines[2].evaluate_resident(0)
        lh, ah = engines[2].evaluate(img, proc, y)
        assert abs(le - lh) <= 1e-6 * abs(lh) and abs(ae - ah) <= 1e-6 * abs(ah)
        assert np.array_equal(engines[2].predict_resident(0), engines[2].predict(x64, p64))
    finally:
        for e in engines:
            e.close()
